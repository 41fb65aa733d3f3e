"""GPU box: a few hyper-parameter-search evaluations (fit_search + LML gradient) at one size (target for ncu launch lists)."""
import sys
import numpy as np
import torch
sys.path.insert(0, ".")
from metaopt_b200.engine import get_engine
n, d = int(sys.argv[1]), int(sys.argv[2])
eng = get_engine(0)
rs = np.random.RandomState(0)
X = torch.as_tensor(rs.uniform(size=(n, d))).cuda(); y = torch.as_tensor(((rs.uniform(size=(n, d)) - 0.5) ** 2).sum(1)).cuda()
for _ in range(3):
    eng.fit_search(X, y, alpha=1e-4, normalize_y=True, length_scale=np.ones(d))
    eng.lml_grad()
eng.sync()
print("ok")
