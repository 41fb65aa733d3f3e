"""GPU box: a few fits at one size (target for ncu launch lists of the fit kernels).  usage: fit_once.py n d"""
import sys
import numpy as np
sys.path.insert(0, ".")
from metaopt_b200.engine import get_engine
n, d = int(sys.argv[1]), int(sys.argv[2])
eng = get_engine(0)
rs = np.random.RandomState(0)
X = rs.uniform(size=(n, d)); y = ((X - 0.5) ** 2).sum(1)
for _ in range(3):
    eng.fit(X, y, alpha=1e-4, normalize_y=True)
eng.sync()
print("ok")
