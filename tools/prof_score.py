"""GPU box: a short program that launches the score kernel a few times (target for ncu)."""
import sys

import numpy as np
import torch

sys.path.insert(0, ".")
from metaopt_b200.engine import get_engine  # noqa: E402

nobs, d = (int(sys.argv[1]), int(sys.argv[2])) if len(sys.argv) > 2 else (2000, 20)
M = int(sys.argv[3]) if len(sys.argv) > 3 else 148 * 64 * 8
eng = get_engine(0)
rs = np.random.RandomState(0)
X = rs.uniform(size=(nobs, d))
y = ((X - 0.5) ** 2).sum(1) + 0.01 * rs.randn(nobs)
eng.fit(X, y, alpha=1e-4, normalize_y=True)
C = torch.as_tensor(np.random.RandomState(1).uniform(size=(M, d))).cuda()
for _ in range(3):
    a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    a.record()
    r = eng.score(C, acq_mask=1, y_opt=float(y.min()))
    b.record()
    torch.cuda.synchronize()
    print("score ms", a.elapsed_time(b), r["EI"])
