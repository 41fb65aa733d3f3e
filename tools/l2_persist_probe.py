import sys, json, torch
sys.path.insert(0, "."); sys.path.insert(0, "tests")
from helpers import cfg5_space, transformed
from metaopt_b200 import compiler
from metaopt_b200.engine import get_engine
def ev(fn, reps=5):
    best = 1e30
    for _ in range(reps):
        a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        torch.cuda.synchronize(); a.record(); fn(); b.record(); torch.cuda.synchronize()
        best = min(best, a.elapsed_time(b))
    return best
n = 1 << 23
x = torch.empty((n, 98), dtype=torch.float64, device="cuda")
out = {"fill_before_ctx": n * 784 / ev(lambda: x.fill_(1.0)) / 1e6}
eng = get_engine(0)
out["fill_after_ctx"] = n * 784 / ev(lambda: x.fill_(1.0)) / 1e6
comp = compiler.compile_space(transformed(cfg5_space())); compiler.upload(eng, comp)
eng.sample(7, 0, n, out=x, check=False)
out["sampler"] = n * 784 / ev(lambda: eng.sample(7, n, n, out=x, check=False)) / 1e6
out["fill_after_sampler"] = n * 784 / ev(lambda: x.fill_(1.0)) / 1e6
print(json.dumps(out))
