"""Timing of the bound pass alone (ob200_score_bounds) over n_obs, to separate per-tile fixed cost from per-evaluation cost."""
import json, os, sys
import numpy as np, torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from metaopt_b200.engine import get_engine
eng = get_engine(0)
d, M = int(sys.argv[1]) if len(sys.argv) > 1 else 20, 1000000
C = torch.rand((M, d), dtype=torch.float64, device="cuda", generator=torch.Generator(device="cuda").manual_seed(1))
out = {}
for n in (100, 250, 500, 1000, 2000, 4000):
    rs = np.random.RandomState(0)
    X = rs.uniform(size=(n, d)); y = ((X - 0.5) ** 2).sum(1) + 0.01 * rs.randn(n)
    eng.fit(X, y, alpha=1e-4, normalize_y=True)
    eng.score_bounds(C); torch.cuda.synchronize()
    a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    a.record()
    for _ in range(5): eng.score_bounds(C)
    b.record(); torch.cuda.synchronize()
    out[n] = a.elapsed_time(b) / 5
print(json.dumps(out))
