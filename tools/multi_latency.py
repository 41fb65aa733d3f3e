"""GPU box (N GPUs, one process): latency of ONE suggest-sized scoring call through MultiEngine."""
import json
import sys
import time

import numpy as np
import torch

sys.path.insert(0, ".")
sys.path.insert(0, "tests")
from helpers import real_space, transformed  # noqa: E402
from metaopt_b200 import compiler  # noqa: E402
from metaopt_b200.multi import get_multi_engine  # noqa: E402

n_dev = torch.cuda.device_count()
out = {"devices": n_dev}
rs = np.random.RandomState(0)
X = rs.uniform(size=(2000, 20))
y = ((X - 0.5) ** 2).sum(axis=1) + 0.01 * rs.randn(2000)
comp = compiler.compile_space(transformed(real_space(20)))
for nd in sorted({1, 2, 4, n_dev} & set(range(1, n_dev + 1))):
    eng = get_multi_engine(list(range(nd)))
    compiler.upload(eng, comp)
    eng._space_key = None
    eng.fit(X, y, alpha=1e-4, normalize_y=True)
    y_opt, M = float(y.min()), 1_000_000
    ref = eng.score(None, M=M, seed=1, acq_mask=1, y_opt=y_opt)
    pr, n_scored = eng.score_pruned(None, M=M, seed=1, acq="EI", y_opt=y_opt)
    assert pr == ref
    lat_d, lat_p = [], []
    for _ in range(10):
        t0 = time.perf_counter(); eng.score(None, M=M, seed=1, acq_mask=1, y_opt=y_opt); lat_d.append((time.perf_counter() - t0) * 1e3)
    for _ in range(30):
        t0 = time.perf_counter(); eng.score_pruned(None, M=M, seed=1, acq="EI", y_opt=y_opt); lat_p.append((time.perf_counter() - t0) * 1e3)
    out["gpus_%d" % nd] = dict(dense_ms_p50=float(np.median(lat_d)), pruned_ms_p50=float(np.median(lat_p)), winner=ref["EI"]["index"])
    print(json.dumps(out), flush=True)
json.dump(out, open("gpurun_out/multi_latency.json", "w"), indent=1)
