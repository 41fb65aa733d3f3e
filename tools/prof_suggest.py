import sys, time, cProfile, pstats, io
import numpy as np
sys.path.insert(0, "."); sys.path.insert(0, "tests")
from helpers import gp_problem
from metaopt_b200.algo import BayesianOptimizer
from metaopt_b200.space import Real, Space, build_required_space
s = Space()
for k in range(20): s.register(Real("x%02d" % k, "uniform", 0, 1))
t = build_required_space("real", s)
X, y = gp_problem(2000, 20, seed=0)
for acq in ("EI", "gp_hedge"):
    a = BayesianOptimizer(t, seed=1, n_initial_points=1, acq_func=acq, n_points=1000000, normalize_y=True, length_scale=1.0, fit_hyperparameters=False)
    a._X, a._y = [r for r in X], list(y)
    for _ in range(3): a.suggest(1)
    ts = []
    for _ in range(30):
        t0 = time.perf_counter(); a.suggest(1); ts.append(time.perf_counter() - t0)
    print(acq, "p50 ms", 1e3 * float(np.median(ts)))
pr = cProfile.Profile(); pr.enable()
for _ in range(20): a.suggest(1)
pr.disable()
st = io.StringIO(); pstats.Stats(pr, stream=st).sort_stats("cumulative").print_stats(14); print(st.getvalue()[:2600])
