"""GPU box: where the default plugin's observe -> deepcopy -> suggest cycle spends its time at n_obs = 2000, d = 20, 1e6 candidates."""
import copy, json, os, sys, time
import numpy as np
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import bench
from metaopt_b200 import hyperopt
from metaopt_b200.algo import BayesianOptimizer
from metaopt_b200.space import build_required_space

X, y = bench.problem()
tspace = build_required_space("real", bench.real_space(bench.DIM))
algo = BayesianOptimizer(tspace, seed=1, n_initial_points=10, acq_func="gp_hedge", alpha=1e-10, noise="gaussian", noise_level=1e-4,
                         normalize_y=True, n_points=1000000, device=0)
algo.observe([tuple(float(v) for v in r) for r in X], [{"objective": float(v)} for v in y])
T = {"hyperopt": [], "evals": [], "score": [], "total": [], "final_fit": []}
real_opt = hyperopt.optimise
def timed_opt(engine, *a, **k):
    t0 = time.perf_counter(); res = real_opt(engine, *a, **k); T["hyperopt"].append((time.perf_counter() - t0) * 1e3); T["evals"].append(res.n_evals); return res
hyperopt.optimise = timed_opt
real_score = BayesianOptimizer._score_range
def timed_score(self, *a, **k):
    t0 = time.perf_counter(); r = real_score(self, *a, **k); T["score"].append((time.perf_counter() - t0) * 1e3); return r
BayesianOptimizer._score_range = timed_score
algo.suggest(1)
rs = np.random.RandomState(5)
for k in T: T[k].clear()
for _ in range(30):
    x_new = tuple(float(v) for v in rs.uniform(size=bench.DIM)); y_new = float(((np.asarray(x_new) - 0.5) ** 2).sum() + 0.01 * rs.randn())
    t0 = time.perf_counter()
    algo.observe([x_new], [{"objective": y_new}]); naive = copy.deepcopy(algo); naive.suggest(1); algo.set_state(naive.state_dict)
    T["total"].append((time.perf_counter() - t0) * 1e3)
print(json.dumps({k: (float(np.median(v)) if v else None) for k, v in T.items()}), "evals", T["evals"], "theta", algo._theta[0], float(np.mean(algo._theta[1])), algo._theta[2], "stats", algo._engine().prune_stats())
