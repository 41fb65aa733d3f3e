"""Acquisition refinement: host L-BFGS-B from the best sampled candidates, every value/gradient evaluation on
the device (`ob200_gp_predict_grad_host`).

Counterpart of skopt's `acq_optimizer="lbfgs"` (what `"auto"` resolves to for the all-real space Orion hands
the plugin): `Optimizer._tell` scores `n_points` sampled candidates, takes the `n_restarts_optimizer` (5) best as
starting points and runs `fmin_l_bfgs_b(gaussian_acquisition_1D, x0, bounds=space.transformed_bounds,
maxiter=20)` from each, keeping the best end point clipped to the bounds (scikit-optimize
`skopt/optimizer/optimizer.py`; third-party, absent from the reference tree, no version pinned -- SURVEY.md row
19).  The utilities (to MAXIMISE) and their gradients, with improve = y_opt - xi - mu and z = improve / sigma:

  EI   = improve Phi(z) + sigma phi(z)        dEI  = -Phi(z) dmu + phi(z) dsigma
  PI   = Phi(z)                               dPI  = -phi(z) (dmu + z dsigma) / sigma
  -LCB = kappa sigma - mu                     d    = kappa dsigma - dmu

(`skopt/acquisition.py`, `gaussian_ei/pi/lcb(..., return_grad=True)`); sigma == 0 gives EI = PI = 0 with a zero
gradient, like the masked branches there.
"""
import math

import numpy as np
from scipy.optimize import fmin_l_bfgs_b

_SQRT2 = math.sqrt(2.0)
_INV_SQRT_2PI = 1.0 / math.sqrt(2.0 * math.pi)


def utility_and_grad(name, mu, sigma, dmu, dsigma, y_opt, xi=0.01, kappa=1.96):
    """(value, gradient [D]) of the utility that is maximised, from one point's posterior and its gradients."""
    dmu = np.asarray(dmu, dtype=float)
    dsigma = np.asarray(dsigma, dtype=float)
    if name == "LCB":
        return kappa * sigma - mu, kappa * dsigma - dmu
    if not sigma > 0.0:
        return 0.0, np.zeros_like(dmu)
    improve = y_opt - xi - mu
    z = improve / sigma
    cdf = 0.5 * math.erfc(-z / _SQRT2)
    pdf = _INV_SQRT_2PI * math.exp(-0.5 * z * z)
    if name == "EI":
        return improve * cdf + sigma * pdf, -cdf * dmu + pdf * dsigma
    if name == "PI":
        return cdf, -pdf * (dmu + z * dsigma) / sigma
    raise ValueError("unknown acquisition %r" % (name,))


class Refined:
    def __init__(self, x, value, mu, sigma, start, n_evals):
        self.x, self.value, self.mu, self.sigma, self.start, self.n_evals = x, value, mu, sigma, start, n_evals


def refine(engine, name, starts, y_opt, xi=0.01, kappa=1.96, lower=0.0, upper=1.0, maxiter=20):
    """L-BFGS-B on the utility from every row of `starts` (normalised coordinates, inside the box).

    Returns the best `Refined` end point -- never worse than the best start, whose own utility is evaluated first
    through the same entry point."""
    starts = np.atleast_2d(np.asarray(starts, dtype=np.float64))
    d = starts.shape[1]
    lo = np.broadcast_to(np.asarray(lower, dtype=float), (d,))
    hi = np.broadcast_to(np.asarray(upper, dtype=float), (d,))
    bounds = list(zip(lo.tolist(), hi.tolist()))
    n_evals = [0]

    def evaluate(x):
        mu, sigma, dmu, dsigma = engine.predict_grad(x[None, :])
        n_evals[0] += 1
        val, grad = utility_and_grad(name, float(mu[0]), float(sigma[0]), dmu[0], dsigma[0], y_opt, xi, kappa)
        return val, grad, float(mu[0]), float(sigma[0])

    best = None
    for k, x0 in enumerate(starts):
        x0 = np.clip(x0, lo, hi)
        v0, _, m0, s0 = evaluate(x0)
        if best is None or v0 > best.value:
            best = Refined(x0.copy(), v0, m0, s0, k, 0)
        x1, _, _ = fmin_l_bfgs_b(lambda x: tuple(-t for t in evaluate(x)[:2]), x0, bounds=bounds, maxiter=int(maxiter))
        x1 = np.clip(x1, lo, hi)
        v1, _, m1, s1 = evaluate(x1)
        if v1 > best.value:
            best = Refined(x1.copy(), v1, m1, s1, k, 0)
    best.n_evals = n_evals[0]
    return best
