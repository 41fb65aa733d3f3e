"""GPU box: latency of the plugin's default mode (hyper-parameters fitted by L-BFGS-B on the device LML gradient):
first suggest (cold search) and suggest after one new observation (warm-started search), with the search's
evaluation count; scikit-learn's own fit (same kernel family, optimizer on, all host cores) beside it where it is
affordable."""
import json
import sys
import time

import numpy as np

sys.path.insert(0, ".")
sys.path.insert(0, "tests")
from helpers import gp_problem  # noqa: E402
from metaopt_b200 import hyperopt  # noqa: E402
from metaopt_b200.algo import BayesianOptimizer  # noqa: E402
from metaopt_b200.space import Real, Space, build_required_space  # noqa: E402

out = {}
for n, d in ((50, 2), (500, 20), (2000, 20)):
    s = Space()
    for k in range(d):
        s.register(Real("x%02d" % k, "uniform", 0, 1))
    t = build_required_space("real", s)
    X, y = gp_problem(n + 1, d, seed=0)
    a = BayesianOptimizer(t, seed=1, n_initial_points=1, acq_func="EI", n_points=1000000, normalize_y=True)
    a._X, a._y = [r for r in X[:n]], list(y[:n])
    evals = []
    real_opt = hyperopt.optimise

    def counted(*args, **kwargs):
        res = real_opt(*args, **kwargs)
        evals.append(res.n_evals)
        return res
    hyperopt.optimise = counted
    t0 = time.perf_counter()
    a.suggest(1)
    cold = time.perf_counter() - t0
    a.observe([tuple(X[n])], [{"objective": float(y[n])}])
    t0 = time.perf_counter()
    a.suggest(1)
    warm = time.perf_counter() - t0
    hyperopt.optimise = real_opt
    rec = dict(cold_suggest_ms=1e3 * cold, warm_suggest_ms=1e3 * warm, lml_evaluations=evals)
    eng = a._engine()
    t0 = time.perf_counter()
    _, n_scored = eng.score_pruned(None, M=1000000, seed=1, first_index=0, acq="EI", y_opt=float(min(a._y)))
    rec["pruned_ms"] = 1e3 * (time.perf_counter() - t0)
    rec["pruned_fully_scored"] = int(n_scored)
    rec["theta"] = [float(a._theta[0]), [float(v) for v in np.atleast_1d(a._theta[1])][:4], float(a._theta[2])]
    if n <= 500:
        from sklearn.gaussian_process import GaussianProcessRegressor
        from sklearn.gaussian_process.kernels import ConstantKernel, Matern, WhiteKernel
        import warnings
        k = ConstantKernel(1.0, (0.01, 1000.0)) * Matern(length_scale=np.ones(d), length_scale_bounds=(0.01, 100.0), nu=2.5) \
            + WhiteKernel(1e-4, (1e-5, 1e5))
        with warnings.catch_warnings():
            warnings.simplefilter("ignore")
            t0 = time.perf_counter()
            GaussianProcessRegressor(kernel=k, alpha=1e-10, normalize_y=True, n_restarts_optimizer=0).fit(X[:n], y[:n])
            rec["sklearn_fit_with_optimizer_ms"] = 1e3 * (time.perf_counter() - t0)
    out["n%d_d%d" % (n, d)] = rec
    print(json.dumps(out), flush=True)
