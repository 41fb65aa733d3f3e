"""GPU box: where the host-buffer entry point (ob200_score_host) spends its time beyond the device-resident call."""
import os, sys, time
import numpy as np, torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from metaopt_b200.engine import get_engine

n, d, M = 2000, 20, 1_000_000
eng = get_engine(0)
rs = np.random.RandomState(0)
X = rs.uniform(size=(n, d)); y = ((X - 0.5) ** 2).sum(1) + 0.01 * rs.randn(n)
eng.fit(X, y, alpha=1e-10 + 1e-4, normalize_y=True)
host = torch.rand((M, d), dtype=torch.float64).pin_memory()
dev = host.cuda()
hn = host.numpy()
y_opt = float(y.min())

def wall(fn, reps=5):
    fn(); torch.cuda.synchronize()
    t0 = time.perf_counter()
    for _ in range(reps): fn()
    torch.cuda.synchronize()
    return (time.perf_counter() - t0) / reps * 1e3

print("device-resident score  ms", round(wall(lambda: eng.score(dev, acq_mask=1, y_opt=y_opt)), 3))
print("host-buffer score      ms", round(wall(lambda: eng.score_host(hn, acq_mask=1, y_opt=y_opt)), 3))
print("H2D alone (160 MB)     ms", round(wall(lambda: dev.copy_(host, non_blocking=True)), 3))
head = 148 * 32 * 16
print("device score of head   ms", round(wall(lambda: eng.score(dev[:head], acq_mask=1, y_opt=y_opt)), 3))
print("device score of tail   ms", round(wall(lambda: eng.score(dev[head:], acq_mask=1, y_opt=y_opt)), 3))
