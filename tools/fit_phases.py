import sys, numpy as np, torch
sys.path.insert(0, ".")
from metaopt_b200.engine import get_engine
eng = get_engine(0)
for n, d in ((2000, 20), (500, 20), (8000, 100)):
    rs = np.random.RandomState(0)
    X = torch.as_tensor(rs.uniform(size=(n, d))).cuda()
    y = torch.as_tensor(((rs.uniform(size=(n, d)) - 0.5) ** 2).sum(1)).cuda()
    for _ in range(3):
        eng.fit(X, y, alpha=1e-4, normalize_y=True)
