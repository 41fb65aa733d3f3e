"""Kernel hyper-parameter optimisation: host L-BFGS-B over log-parameters, every objective/gradient
evaluation on the device (`ob200_gp_fit` + `ob200_gp_lml_grad`).

Counterpart of what the reference's plugin gets from scikit-learn
(`GaussianProcessRegressor.fit`: `_constrained_optimization` with `fmin_l_bfgs_b` on the negative
log marginal likelihood, `n_restarts_optimizer` extra starts drawn log-uniformly inside the bounds,
sklearn `_gpr.py:299-340`), with skopt's kernel `ConstantKernel(1, (0.01, 1000)) * Matern|RBF(ls,
(0.01, 100)) [+ WhiteKernel(1, (1e-5, 1e5))]`.  theta = [log amplitude, log ls_1..D (or one), log noise].
"""
import numpy as np
from scipy.optimize import minimize

from . import _lib

AMP_BOUNDS = (1e-2, 1e3)
LS_BOUNDS = (1e-2, 1e2)
NOISE_BOUNDS = (1e-5, 1e5)


class FitResult:
    def __init__(self, amplitude, length_scale, noise, lml, n_evals):
        self.amplitude, self.length_scale, self.noise, self.lml, self.n_evals = amplitude, length_scale, noise, lml, n_evals


def optimise(engine, X, y, kind, amplitude, length_scale, noise, alpha, normalize_y, anisotropic=True,
             optimise_noise=True, n_restarts=0, rng=None, max_iter=200):
    """Maximise the log marginal likelihood; leaves `engine` fitted at the best hyper-parameters.

    X, y: torch CUDA tensors (kept on the device across evaluations) or numpy arrays."""
    D = X.shape[1]
    n_ls = D if anisotropic else 1
    ls0 = np.broadcast_to(np.asarray(length_scale, dtype=float), (n_ls,)).copy()
    theta0 = np.concatenate([[np.log(amplitude)], np.log(ls0), [np.log(max(noise, NOISE_BOUNDS[0]))] if optimise_noise else []])
    bounds = [tuple(np.log(AMP_BOUNDS))] + [tuple(np.log(LS_BOUNDS))] * n_ls + ([tuple(np.log(NOISE_BOUNDS))] if optimise_noise else [])
    n_evals = [0]
    # search evaluations need L, alpha and inv(L) only: `fit_search` skips what only the score kernel reads
    fit = engine.fit_search if hasattr(engine, "fit_search") and hasattr(X, "is_cuda") else engine.fit

    def unpack(theta):
        amp = float(np.exp(theta[0]))
        ls = np.exp(theta[1:1 + n_ls])
        nz = float(np.exp(theta[1 + n_ls])) if optimise_noise else float(noise)
        return amp, ls, nz

    def objective(theta):
        amp, ls, nz = unpack(theta)
        n_evals[0] += 1
        try:
            fit(X, y, kind=kind, amplitude=amp, length_scale=ls, alpha=alpha + nz, normalize_y=normalize_y)
        except _lib.NotPositiveDefinite:
            return 1e25, np.zeros_like(theta)  # sklearn returns -inf for the likelihood here
        lml, g = engine.lml_grad()
        grad = [g[0]]
        grad += list(g[1:1 + D]) if anisotropic else [float(np.sum(g[1:1 + D]))]
        if optimise_noise:
            grad.append(nz * g[1 + D])
        return -lml, -np.asarray(grad)

    starts = [theta0]
    rng = rng or np.random.RandomState(0)
    for _ in range(int(n_restarts)):
        starts.append(np.array([rng.uniform(lo, hi) for lo, hi in bounds]))
    best = None
    for start in starts:
        res = minimize(objective, np.clip(start, [b[0] for b in bounds], [b[1] for b in bounds]), jac=True,
                       method="L-BFGS-B", bounds=bounds, options={"maxiter": max_iter})
        if best is None or res.fun < best.fun:
            best = res
    amp, ls, nz = unpack(best.x)
    engine.fit(X, y, kind=kind, amplitude=amp, length_scale=ls, alpha=alpha + nz, normalize_y=normalize_y)
    return FitResult(amp, ls, nz, engine.lml(), n_evals[0])
