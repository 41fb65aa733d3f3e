"""GPU box: cProfile of BayesianOptimizer.suggest(1) with the fit cached (host-side overhead around the scoring call)."""
import cProfile, pstats, sys, os, io
import numpy as np
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import bench
from metaopt_b200.algo import BayesianOptimizer
from metaopt_b200.space import build_required_space
X, y = bench.problem()
ts = build_required_space("real", bench.real_space(bench.DIM))
a = BayesianOptimizer(ts, seed=1, acq_func="EI", n_points=10000, fit_hyperparameters=False, normalize_y=True, noise_level=1e-4, device=0)
a.observe([tuple(float(v) for v in r) for r in X], [{"objective": float(v)} for v in y])
for _ in range(20): a.suggest(1)
pr = cProfile.Profile(); pr.enable()
for _ in range(300): a.suggest(1)
pr.disable()
s = io.StringIO(); pstats.Stats(pr, stream=s).sort_stats("tottime").print_stats(18); print(s.getvalue()[:3500])
