"""GPU box: pruned entry point at small candidate counts (the plugin's default n_points is 10 000): where does the dense
fallback stop paying?  OB200_PRUNE_MIN_BATCHES = k sends M <= k * 4736 to the dense kernel."""
import json, os, sys, time
import numpy as np, torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from metaopt_b200.engine import get_engine

n, d = (int(sys.argv[1]) if len(sys.argv) > 1 else 2000), 20
eng = get_engine(0)
rs = np.random.RandomState(0)
X = rs.uniform(size=(n, d)); y = ((X - 0.5) ** 2).sum(1) + 0.01 * rs.randn(n)
eng.fit(X, y, alpha=1e-10 + 1e-4, normalize_y=True)
y_opt = float(y.min())
out = {}
for M in (4736, 6000, 10000, 15000, 25000, 50000, 100000, 200000):
    C = torch.rand((M, d), dtype=torch.float64, device="cuda", generator=torch.Generator(device="cuda").manual_seed(1))
    row = {}
    for k in ("64", "1"):
        os.environ["OB200_PRUNE_MIN_BATCHES"] = k
        for mask, name in ((1, "EI"), (7, "hedge")):
            eng.score_pruned_multi(C, acq_mask=mask, y_opt=y_opt); torch.cuda.synchronize()
            t0 = time.perf_counter()
            for _ in range(20): r = eng.score_pruned_multi(C, acq_mask=mask, y_opt=y_opt)
            row[("dense_" if k == "64" else "pruned_") + name] = round((time.perf_counter() - t0) / 20 * 1e3, 3)
    out[M] = row
    print(M, row, flush=True)
