"""Timing of the pruned path and of its bound pass alone (cfg2 shape by default): FP32 bound kernel vs the FP64 one."""
import json, os, sys, time
import numpy as np, torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from metaopt_b200.engine import get_engine

n, d, M = [int(v) for v in (sys.argv[1:4] + [2000, 20, 1000000][len(sys.argv) - 1:])]
eng = get_engine(0)
rs = np.random.RandomState(0)
X = rs.uniform(size=(n, d)); y = ((X - 0.5) ** 2).sum(1) + 0.01 * rs.randn(n)
eng.fit(X, y, alpha=1e-10 + 1e-4, normalize_y=True)
C = torch.rand((M, d), dtype=torch.float64, device="cuda", generator=torch.Generator(device="cuda").manual_seed(1))
y_opt = float(y.min())
out = {"n": n, "d": d, "M": M}

def timed(fn, reps=10):
    fn(); torch.cuda.synchronize()
    a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    a.record()
    for _ in range(reps): fn()
    b.record(); torch.cuda.synchronize()
    return a.elapsed_time(b) / reps

for mode in ("1", "0"):
    os.environ["OB200_BOUND_F32"] = mode
    tag = "f32" if mode == "1" else "f64"
    out["bounds_ms_" + tag] = timed(lambda: eng.score_bounds(C))
    for mask, name in ((1, "EI"), (7, "hedge")):
        out["pruned_ms_%s_%s" % (name, tag)] = timed(lambda: eng.score_pruned_multi(C, acq_mask=mask, y_opt=y_opt))
        out["stats_%s_%s" % (name, tag)] = eng.prune_stats()
dense = eng.score(C, acq_mask=7, y_opt=y_opt)
os.environ["OB200_BOUND_F32"] = "1"
pr, _ = eng.score_pruned_multi(C, acq_mask=7, y_opt=y_opt)
out["same_winner"] = pr == dense
mu_lb, sg_ub, used = eng.score_bounds(C[:200000])
ex = eng.score(C[:200000], acq_mask=1, y_opt=y_opt, dense=True)
_, y_std, _ = eng.ystats()
out["gap_over_ystd"] = [float(((ex["mu"] - mu_lb) / y_std).min()), float(((ex["mu"] - mu_lb) / y_std).max())]
out["rigorous"] = bool(torch.all(mu_lb <= ex["mu"])) and bool(torch.all(sg_ub >= ex["sigma"]))
print(json.dumps(out))
