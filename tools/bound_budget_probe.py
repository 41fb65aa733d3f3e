"""GPU box: the hyper-parameters the default plugin mode ends up with at cfg2, and what they do to the FP32 bound
pass's error budget: S = sum_j |alpha_j| k_j (the budget's scale), the same with the kernel centred on the
amplitude, S_c = sum_j |alpha_j| |k_j - amp|, and the spread of the posterior mean they have to resolve."""
import json
import os
import sys

import numpy as np
import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import bench  # noqa: E402
from metaopt_b200.algo import BayesianOptimizer  # noqa: E402
from metaopt_b200.space import build_required_space  # noqa: E402

X, y = bench.problem()
tspace = build_required_space("real", bench.real_space(bench.DIM))
algo = BayesianOptimizer(tspace, seed=1, n_initial_points=10, acq_func="gp_hedge", alpha=1e-10, noise="gaussian",
                         noise_level=1e-4, normalize_y=True, n_points=1000000, device=0)
algo.observe([tuple(float(v) for v in r) for r in X], [{"objective": float(v)} for v in y])
algo.suggest(1)
amp, ls, noise = algo._theta
ls = np.asarray(ls, dtype=np.float64)
eng = algo._engine()
L, alpha = eng.get_factor()
Xs = np.asarray(X) / ls
rs = np.random.RandomState(3)
C = rs.uniform(size=(2000, bench.DIM)) / ls
d2 = np.maximum((C ** 2).sum(1)[:, None] + (Xs ** 2).sum(1)[None, :] - 2.0 * C @ Xs.T, 0.0)
t = np.sqrt(5.0 * d2)
k = amp * (1.0 + t + t * t / 3.0) * np.exp(-t)
mu = k @ alpha
S = k @ np.abs(alpha)
Sc = np.abs(k - amp) @ np.abs(alpha)
kbar = k.mean()
Sm = np.abs(k - kbar) @ np.abs(alpha)
out = dict(amplitude=float(amp), noise=float(noise), ls_min=float(ls.min()), ls_mean=float(ls.mean()), ls_max=float(ls.max()),
           alpha_l1=float(np.abs(alpha).sum()), alpha_sum=float(alpha.sum()), mu_std_over_candidates=float(mu.std()),
           mu_min=float(mu.min()), mu_max=float(mu.max()), S_max=float(S.max()), S_centred_on_amp_max=float(Sc.max()),
           S_centred_on_mean_k_max=float(Sm.max()), k_over_amp_min=float(k.min() / amp), k_over_amp_max=float(k.max() / amp),
           t_max=float(t.max()), stats=eng.prune_stats())
print(json.dumps(out))
os.makedirs("gpurun_out", exist_ok=True)
json.dump(out, open("gpurun_out/bound_budget_probe.json", "w"), indent=1)
