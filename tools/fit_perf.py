"""GPU box: wall time of one fit and of one fit + LML gradient at several n (after warm-up)."""
import json
import sys
import time

import numpy as np
import torch

sys.path.insert(0, ".")
from metaopt_b200.engine import get_engine  # noqa: E402

eng = get_engine(0)
out = {}
for n, d in ((50, 2), (200, 10), (500, 20), (2000, 20), (4000, 60), (8000, 100)):
    rs = np.random.RandomState(0)
    X = torch.as_tensor(rs.uniform(size=(n, d))).cuda()
    y = torch.as_tensor(((rs.uniform(size=(n, d)) - 0.5) ** 2).sum(1)).cuda()
    for _ in range(2):
        eng.fit(X, y, alpha=1e-4, normalize_y=True)
        eng.lml_grad()
    reps = 5 if n <= 2000 else 2
    t0 = time.perf_counter()
    for _ in range(reps):
        eng.fit(X, y, alpha=1e-4, normalize_y=True)
    fit_ms = (time.perf_counter() - t0) / reps * 1e3
    t0 = time.perf_counter()
    for _ in range(reps):
        eng.fit(X, y, alpha=1e-4, normalize_y=True)
        eng.lml_grad()
    both_ms = (time.perf_counter() - t0) / reps * 1e3
    t0 = time.perf_counter()
    for _ in range(reps):
        eng.fit_search(X, y, alpha=1e-4, normalize_y=True)
        eng.lml_grad()
    search_ms = (time.perf_counter() - t0) / reps * 1e3
    eng.fit(X, y, alpha=1e-4, normalize_y=True)
    out["n%d_d%d" % (n, d)] = dict(fit_ms=round(fit_ms, 3), fit_plus_grad_ms=round(both_ms, 3), search_eval_ms=round(search_ms, 3))
    print(json.dumps(out), flush=True)

# vendor reference for the factorisation alone: torch.linalg.cholesky (cuSOLVER potrf) on a Gram matrix of the same size
ref = {}
for n in (500, 2000, 4000, 8000):
    A = torch.randn(n, n, dtype=torch.float64, device="cuda")
    K = A @ A.T / n + torch.eye(n, dtype=torch.float64, device="cuda")
    for _ in range(2):
        torch.linalg.cholesky(K)
    torch.cuda.synchronize()
    a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    a.record()
    for _ in range(5):
        L = torch.linalg.cholesky(K)
    b.record()
    torch.cuda.synchronize()
    t_chol = a.elapsed_time(b) / 5
    a.record()
    for _ in range(5):
        torch.linalg.solve_triangular(L, torch.eye(n, dtype=torch.float64, device="cuda"), upper=False)
    b.record()
    torch.cuda.synchronize()
    ref["n%d" % n] = dict(cusolver_potrf_ms=round(t_chol, 3), cublas_trsm_identity_ms=round(a.elapsed_time(b) / 5, 3))
print(json.dumps({"vendor_reference": ref}), flush=True)
