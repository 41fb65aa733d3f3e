"""GPU box: the FP64 mean-only bound pass (score_kernel<1>) on cfg2's shape, alone (target for ncu)."""
import os, sys
import numpy as np, torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
os.environ["OB200_BOUND_F32"] = "0"
from metaopt_b200.engine import get_engine

n, d, M = 2000, 20, 1000000
eng = get_engine(0)
rs = np.random.RandomState(0)
X = rs.uniform(size=(n, d)); y = ((X - 0.5) ** 2).sum(1) + 0.01 * rs.randn(n)
eng.fit(X, y, alpha=1e-10 + 1e-4, normalize_y=True)
C = torch.rand((M, d), dtype=torch.float64, device="cuda", generator=torch.Generator(device="cuda").manual_seed(1))
for _ in range(3):
    a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    a.record(); eng.score_bounds(C); b.record(); torch.cuda.synchronize()
    print("bounds ms", a.elapsed_time(b))
