"""GPU box: sampler throughput for every libob200*.so variant."""
import glob, os, subprocess, sys
CHILD = r'''
import sys, json, torch
sys.path.insert(0, "."); sys.path.insert(0, "tests")
from helpers import cfg5_space, real_space, cfg3_space, transformed
from metaopt_b200 import compiler
from metaopt_b200.engine import get_engine
eng = get_engine(0); out = {}
for name, space in (("cfg5", cfg5_space()), ("real20", real_space(20)), ("cfg3", cfg3_space())):
    comp = compiler.compile_space(transformed(space)); compiler.upload(eng, comp)
    n = 1 << 23
    buf = torch.empty((n, comp.d_out), dtype=torch.float64, device="cuda")
    eng.sample(7, 0, n, out=buf, check=False); best = 1e30
    for _ in range(3):
        a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        a.record(); eng.sample(7, n, n, out=buf, check=False); b.record(); torch.cuda.synchronize()
        best = min(best, a.elapsed_time(b))
    out[name] = round(n * comp.d_out * 8 / best / 1e6, 1)
    del buf
print(json.dumps(out))
'''
for lib in sorted(glob.glob("metaopt_b200/libob200*.so")):
    env = dict(os.environ, OB200_LIB=os.path.abspath(lib))
    res = subprocess.run([sys.executable, "-c", CHILD], env=env, capture_output=True, text=True, timeout=120)
    print(os.path.basename(lib), res.stdout.strip().splitlines()[-1] if res.stdout.strip() else res.stderr[-300:], flush=True)
