"""Development probe (GPU box): FP64 peaks and a first timing of the score kernel at cfg2."""
import json
import sys
import time

import numpy as np
import torch

sys.path.insert(0, ".")
from metaopt_b200.engine import get_engine  # noqa: E402


def ev_time(fn, reps=3):
    torch.cuda.synchronize()
    best = 1e30
    for _ in range(reps):
        a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        a.record()
        fn()
        b.record()
        torch.cuda.synchronize()
        best = min(best, a.elapsed_time(b))
    return best


def main():
    out = {}
    n = 8192
    A = torch.randn(n, n, dtype=torch.float64, device="cuda")
    B = torch.randn(n, n, dtype=torch.float64, device="cuda")
    torch.matmul(A, B)
    ms = ev_time(lambda: torch.matmul(A, B), 5)
    out["cublas_dgemm_tflops"] = 2 * n ** 3 / ms / 1e9
    del A, B
    eng = get_engine(0)
    for (nobs, d, M) in ((2000, 20, 200_000), (2000, 20, 1_000_000), (4000, 60, 200_000), (8000, 100, 100_000)):
        rs = np.random.RandomState(0)
        X = rs.uniform(size=(nobs, d))
        y = ((X - 0.5) ** 2).sum(1) + 0.01 * rs.randn(nobs)
        t0 = time.time()
        eng.fit(X, y, alpha=1e-10 + 1e-4, normalize_y=True)
        eng.sync()
        fit_s = time.time() - t0
        t0 = time.time()
        eng.fit(X, y, alpha=1e-10 + 1e-4, normalize_y=True)
        eng.sync()
        fit2_s = time.time() - t0
        C = torch.rand((M, d), dtype=torch.float64, device="cuda")
        eng.score(C[:10000], acq_mask=1, y_opt=float(y.min()))
        ms = ev_time(lambda: eng.score(C, acq_mask=1, y_opt=float(y.min())), 2)
        flops = M * (nobs ** 2 + 2 * nobs * d + 2 * nobs)
        out["score_n%d_d%d_M%d" % (nobs, d, M)] = dict(ms=ms, cand_per_s=M / ms * 1e3, tflops=flops / ms / 1e9,
                                                       fit_first_s=fit_s, fit_s=fit2_s)
        print(json.dumps(out), flush=True)
    print(json.dumps(out))


if __name__ == "__main__":
    main()
