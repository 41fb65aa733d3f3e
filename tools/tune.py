"""GPU box: time the score kernel for every libob200*.so variant found (one subprocess each)."""
import glob
import json
import os
import subprocess
import sys

CHILD = r'''
import json, sys, numpy as np, torch
sys.path.insert(0, ".")
from metaopt_b200.engine import get_engine
eng = get_engine(0)
out = {}
for (nobs, d, M) in ((2000, 20, 400_000), (4000, 60, 100_000)):
    rs = np.random.RandomState(0)
    X = rs.uniform(size=(nobs, d)); y = ((X - 0.5) ** 2).sum(1) + 0.01 * rs.randn(nobs)
    eng.fit(X, y, alpha=1e-4, normalize_y=True)
    C = torch.rand((M, d), dtype=torch.float64, device="cuda")
    r0 = eng.score(C[:20000], acq_mask=1, y_opt=float(y.min()))
    best = 1e30
    for _ in range(3):
        a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        a.record(); r = eng.score(C, acq_mask=1, y_opt=float(y.min())); b.record(); torch.cuda.synchronize()
        best = min(best, a.elapsed_time(b))
    out["n%d" % nobs] = dict(ms=round(best, 3), tflops=round(M * (nobs * nobs + 2 * nobs * d + 2 * nobs) / best / 1e9, 2), idx=r["EI"]["index"])
print(json.dumps(out))
'''

if __name__ == "__main__":
    libs = sorted(glob.glob("metaopt_b200/libob200*.so"))
    envs = [dict(os.environ, OB200_LIB=os.path.abspath(lib)) for lib in libs]
    if "--p1" in sys.argv:  # sweep the phase-1 ring depth on the main library
        libs = ["p1_stages=%d" % k for k in (1, 2, 3, 4, 6)]
        envs = [dict(os.environ, OB200_P1_STAGES=str(k)) for k in (1, 2, 3, 4, 6)]
    for lib, env in zip(libs, envs):
        try:
            res = subprocess.run([sys.executable, "-c", CHILD], env=env, capture_output=True, text=True, timeout=120)
            print(os.path.basename(lib), res.stdout.strip().splitlines()[-1] if res.stdout.strip() else res.stderr[-400:], flush=True)
        except subprocess.TimeoutExpired:
            print(os.path.basename(lib), "TIMEOUT", flush=True)
