"""GPU box: one more observation by refit vs by ob200_gp_append, and a constant-liar batch suggest(8)."""
import json
import sys
import time

import numpy as np

sys.path.insert(0, ".")
sys.path.insert(0, "tests")
from helpers import gp_problem  # noqa: E402
from metaopt_b200.algo import BayesianOptimizer  # noqa: E402
from metaopt_b200.engine import get_engine  # noqa: E402
from metaopt_b200.space import Real, Space, build_required_space  # noqa: E402

eng = get_engine(0)
out = {}
for n, d in ((50, 2), (1000, 20), (2000, 20), (8000, 100)):
    X, y = gp_problem(n + 12, d, seed=1)
    kw = dict(length_scale=0.5 * np.sqrt(d), alpha=1e-4, normalize_y=True)
    rec = {}
    for mode in (False, True):
        eng.fit(X[:n], y[:n], incremental=False, **kw)
        ts = []
        for k in range(1, 11):
            t0 = time.perf_counter()
            eng.fit(X[:n + k], y[:n + k], incremental=mode, **kw)
            ts.append(time.perf_counter() - t0)
        rec["append_ms" if mode else "refit_ms"] = 1e3 * float(np.median(ts))
    out["n%d_d%d" % (n, d)] = rec
s = Space()
for k in range(20):
    s.register(Real("x%02d" % k, "uniform", 0, 1))
t = build_required_space("real", s)
X, y = gp_problem(2000, 20, seed=0)
for label, inc in (("refit", False), ("append", True)):
    a = BayesianOptimizer(t, seed=1, n_initial_points=1, acq_func="EI", n_points=1000000, normalize_y=True,
                          length_scale=1.0, fit_hyperparameters=False)
    a._X, a._y = [r for r in X], list(y)
    if not inc:
        real_fit = eng.fit
        eng.fit = lambda *args, **kwargs: real_fit(*args, **dict(kwargs, incremental=False))
    a.suggest(8)
    t0 = time.perf_counter()
    a.suggest(8)
    out["suggest8_1M_n2000_%s_ms" % label] = 1e3 * (time.perf_counter() - t0)
    if not inc:
        del eng.fit
print(json.dumps(out, indent=1))
