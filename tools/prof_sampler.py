"""GPU box: launch the sampler kernel on the cfg5 space and on 20 reals (target for ncu)."""
import sys

import torch

sys.path.insert(0, ".")
sys.path.insert(0, "tests")
from helpers import cfg5_space, real_space, transformed  # noqa: E402
from metaopt_b200 import compiler  # noqa: E402
from metaopt_b200.engine import get_engine  # noqa: E402

eng = get_engine(0)
n = 1 << 22
for space in (cfg5_space(), real_space(20)):
    comp = compiler.compile_space(transformed(space))
    compiler.upload(eng, comp)
    buf = torch.empty((n, comp.d_out), dtype=torch.float64, device="cuda")
    for _ in range(2):
        a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        a.record()
        eng.sample(7, 0, n, out=buf, check=False)
        b.record()
        torch.cuda.synchronize()
        print(comp.d_out, "ms", a.elapsed_time(b), "GB/s", n * comp.d_out * 8 / a.elapsed_time(b) / 1e6)
