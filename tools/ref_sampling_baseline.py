"""Build container only (the reference tree is not on the GPU box): time the UNMODIFIED reference's own
`Space.sample(n)` + `TransformedSpace.transform` loop on BASELINE config 5's space (SURVEY.md 8(d) CPU baseline
(ii)); single core, pure Python, as the reference runs it.  Writes one JSON object to stdout."""
import json
import os
import sys
import time
import warnings

import numpy as np

sys.path.insert(0, ".")
from oracle import ref_loader  # noqa: E402

warnings.simplefilter("ignore")
ref_loader.load()
real_meshgrid = np.meshgrid
np.meshgrid = lambda *a, **k: list(real_meshgrid(*a, **k))  # numpy-1.x return type for transformer.py:320-322
from orion.algo.space import Categorical, Integer, Real, Space  # noqa: E402
from orion.core.worker.transformer import build_required_space  # noqa: E402

s = Space()
for i in range(9):
    s.register(Real("ru%02d" % i, "uniform", -1.0 - i, 2.0 + i))
for i in range(8):
    s.register(Real("rl%02d" % i, "reciprocal", 1e-5, 1.0))
for i in range(17):
    s.register(Integer("in%02d" % i, "uniform", 0, 10 + i))
for i in range(16):
    probs = [0.25] * 4 if i % 2 == 0 else [0.1, 0.2, 0.3, 0.4]
    s.register(Categorical("ca%02d" % i, dict(zip("abcd", probs))))
t = build_required_space("real", s)

n = int(sys.argv[1]) if len(sys.argv) > 1 else 2000
s.sample(50, seed=7)
t0 = time.perf_counter()
pts = s.sample(n, seed=7)
t_sample = time.perf_counter() - t0
t0 = time.perf_counter()
flat = [t.transform(p) for p in pts]
t_transform = time.perf_counter() - t0
print(json.dumps(dict(workload="cfg5 space: 9 uniform + 8 loguniform + 17 integer + 16 categorical(k=4), D_in=50, D_out=98",
                      points=n, cores=1, host_cpus=os.cpu_count(),
                      space_sample_s=t_sample, transform_s=t_transform,
                      points_per_s_sample_only=n / t_sample, points_per_s_sample_plus_transform=n / (t_sample + t_transform),
                      note="reference code unmodified, imported with oracle/ref_loader.py shims; measured in the build "
                           "container (the reference tree does not travel to the GPU box)")))
