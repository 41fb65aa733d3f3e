"""skopt acquisition functions restated (test infrastructure).

`skopt.acquisition.gaussian_ei / gaussian_pi / gaussian_lcb` are NOT in /root/reference and
skopt is not installed here (unpinned third-party dependency of the un-vendored
`orion.algo.skopt`, `docs/src/install/plugins.rst:18-46`).  Published formulas
(scikit-optimize 0.5-0.9, `skopt/acquisition.py`), for minimisation:

  EI : mask sigma>0 ; improve = y_opt - xi - mu ; z = improve/sigma ;
       EI = improve*Phi(z) + sigma*phi(z) ; 0 where sigma == 0          (xi = 0.01)
  PI : Phi((y_opt - xi - mu)/sigma) where sigma>0 else 0
  LCB: mu - kappa*sigma                                                   (kappa = 1.96)

The optimiser minimises -EI, -PI, +LCB over the sampled candidates and takes
`X[np.argmin(values)]` (first occurrence wins ties).
"""
import numpy as np
from scipy.stats import norm

ACQ_EI, ACQ_PI, ACQ_LCB = 0, 1, 2


def gaussian_ei(mu, std, y_opt, xi=0.01):
    values = np.zeros_like(mu)
    mask = std > 0
    improve = y_opt - xi - mu[mask]
    scaled = improve / std[mask]
    values[mask] = improve * norm.cdf(scaled) + std[mask] * norm.pdf(scaled)
    return values


def gaussian_pi(mu, std, y_opt, xi=0.01):
    values = np.zeros_like(mu)
    mask = std > 0
    values[mask] = norm.cdf((y_opt - xi - mu[mask]) / std[mask])
    return values


def gaussian_lcb(mu, std, kappa=1.96):
    return mu - kappa * std


def utility(which, mu, std, y_opt, xi=0.01, kappa=1.96):
    """The quantity that is MAXIMISED (EI, PI, -LCB)."""
    if which == ACQ_EI:
        return gaussian_ei(mu, std, y_opt, xi)
    if which == ACQ_PI:
        return gaussian_pi(mu, std, y_opt, xi)
    return -gaussian_lcb(mu, std, kappa)


def pick(which, mu, std, y_opt, xi=0.01, kappa=1.96):
    """argmin of the minimised acquisition == first argmax of the utility."""
    if which == ACQ_LCB:
        vals = gaussian_lcb(mu, std, kappa)
    else:
        vals = -utility(which, mu, std, y_opt, xi, kappa)
    return int(np.argmin(vals))
