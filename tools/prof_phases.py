"""GPU box: where a score_kernel tile spends its cycles (needs the SC_PROFILE_CYCLES build:
`python tools/build_variants.py prof`, then OB200_LIB=metaopt_b200/libob200_prof.so python tools/prof_phases.py)."""
import ctypes as C
import json
import sys

import numpy as np
import torch

sys.path.insert(0, ".")
sys.path.insert(0, "tests")
from helpers import gp_problem, real_space, transformed  # noqa: E402
from metaopt_b200 import _lib, compiler  # noqa: E402
from metaopt_b200.engine import get_engine  # noqa: E402

eng = get_engine(0)
n, d, M = 2000, 20, 1000000
X, y = gp_problem(n, d, seed=0)
eng.fit(X, y, length_scale=1.0, alpha=1e-4 + 1e-10, normalize_y=True)
comp = compiler.compile_space(transformed(real_space(d)))
compiler.upload(eng, comp)
cand = eng.sample(1, 0, M)
for _ in range(2):
    eng.score(cand, acq_mask=1, y_opt=float(y.min()))
a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
a.record()
eng.score(cand, acq_mask=1, y_opt=float(y.min()))
b.record()
torch.cuda.synchronize()
lib = _lib.load()
n_cta = 148
buf = (C.c_ulonglong * (n_cta * 8))()
lib.ob200_debug_score_cycles.argtypes = [C.c_void_p, C.c_int]
assert lib.ob200_debug_score_cycles(buf, n_cta) == 0
cyc = np.asarray(list(buf), dtype=np.float64).reshape(n_cta, 8)
names = ["stage candidates", "phase 1 loop", "wait (A)", "phase 2", "reduce + phase 3", "wait (B)", "-", "loop overhead"]
tot = cyc.sum(axis=1).mean()
out = {"ms": a.elapsed_time(b), "cycles_per_cta": tot,
       "share": {nm: float(cyc[:, k].mean() / tot) for k, nm in enumerate(names) if nm != "-"}}
print(json.dumps(out, indent=1))
