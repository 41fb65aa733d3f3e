"""Latency of one posterior-gradient evaluation, and of a plugin suggest() with acq_optimizer = sampling / lbfgs
(1M candidates, synthetic observations injected, fixed hyper-parameters)."""
import json
import sys
import time

import numpy as np

sys.path.insert(0, ".")
from metaopt_b200.algo import BayesianOptimizer  # noqa: E402
from metaopt_b200.engine import get_engine  # noqa: E402
from metaopt_b200.space import Real, Space, build_required_space  # noqa: E402

eng = get_engine(0)
out = {}
for n, d in ((50, 2), (2000, 20), (8000, 100)):
    s = Space()
    for k in range(d):
        s.register(Real("x%03d" % k, "uniform", 0, 1))
    t = build_required_space("real", s)
    rng = np.random.RandomState(0)
    X = rng.rand(n, d)
    y = ((X - 0.4) ** 2).sum(1) + 0.01 * rng.randn(n)
    rec = {}
    for mode in ("sampling", "lbfgs"):
        a = BayesianOptimizer(t, seed=1, n_initial_points=1, acq_func="EI", n_points=1000000, normalize_y=True,
                              length_scale=0.5 * np.sqrt(d), fit_hyperparameters=False, acq_optimizer=mode)
        a._X, a._y = [r for r in X], list(y)
        a.suggest(1)
        ts = []
        for _ in range(5):
            t0 = time.perf_counter()
            (p,) = a.suggest(1)
            ts.append(time.perf_counter() - t0)
        rec[mode + "_suggest_ms"] = 1e3 * float(np.median(ts))
    P = rng.rand(1, d)
    eng.predict_grad(P)
    t0 = time.perf_counter()
    for _ in range(50):
        eng.predict_grad(P)
    rec["predict_grad_ms"] = (time.perf_counter() - t0) / 50 * 1e3
    out["n%d_d%d" % (n, d)] = rec
print(json.dumps(out, indent=1))
