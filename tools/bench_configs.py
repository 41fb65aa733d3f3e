"""GPU box: throughput of the other BASELINE configs (cfg3/cfg4 scoring shapes, cfg5 sampler) on ONE GPU.
Not the contract bench line (bench.py is); results go to profiles/ as supporting measurements."""
import json
import sys
import time

import numpy as np
import torch

sys.path.insert(0, ".")
sys.path.insert(0, "tests")
from helpers import cfg3_space, cfg5_space, real_space, transformed  # noqa: E402
from metaopt_b200 import compiler  # noqa: E402
from metaopt_b200.engine import get_engine  # noqa: E402

PEAKS = json.load(open("MEASURED_PEAKS.json")) if __import__("os").path.exists("MEASURED_PEAKS.json") else {"hbm_gbs": 6650.0}


def ev(fn, reps=3):
    best = 1e30
    for _ in range(reps):
        a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        torch.cuda.synchronize()
        a.record()
        fn()
        b.record()
        torch.cuda.synchronize()
        best = min(best, a.elapsed_time(b))
    return best


def main():
    eng = get_engine(0)
    out = {"hbm_gbs_measured_peak": PEAKS.get("hbm_gbs")}
    # ---- cfg5: Space.sample + transform, 50-D mixed -> 98 columns, chunks of 2^24 points
    comp = compiler.compile_space(transformed(cfg5_space()))
    compiler.upload(eng, comp)
    chunk = 1 << 24
    buf = torch.empty((chunk, comp.d_out), dtype=torch.float64, device="cuda")
    eng.sample(7, 0, chunk, out=buf)
    ms = ev(lambda: eng.sample(7, chunk, chunk, out=buf, check=False), 3)
    gbs = chunk * comp.d_out * 8 / ms / 1e6
    out["cfg5_sampler"] = dict(points=chunk, d_out=comp.d_out, ms=ms, points_per_s=chunk / ms * 1e3, achieved_gbs=gbs,
                               frac_of_measured_hbm=gbs / PEAKS.get("hbm_gbs", 6650.0),
                               extrapolated_1e9_points_s=1e9 / (chunk / ms * 1e3))
    # pure-write reference: torch fill of the same buffer
    ms_fill = ev(lambda: buf.fill_(1.0), 3)
    out["cfg5_sampler"]["torch_fill_gbs"] = chunk * comp.d_out * 8 / ms_fill / 1e6
    # 20-D real (cfg2 candidates)
    comp2 = compiler.compile_space(transformed(real_space(20)))
    compiler.upload(eng, comp2)
    buf2 = torch.empty((1 << 24, 20), dtype=torch.float64, device="cuda")
    eng.sample(1, 0, 1 << 24, out=buf2)
    ms2 = ev(lambda: eng.sample(1, 0, 1 << 24, out=buf2, check=False), 3)
    out["real20_sampler"] = dict(points=1 << 24, ms=ms2, achieved_gbs=(1 << 24) * 160 / ms2 / 1e6)
    del buf, buf2
    print(json.dumps(out), flush=True)
    # CPU restatement (numpy, vectorised Philox) of the same sampler on a bounded sample, single core
    from oracle import sampling as osamp
    t0 = time.perf_counter()
    osamp.sample_transform(comp.slots, comp.cdf, 7, 0, 20000, True)
    out["cfg5_sampler"]["cpu_port_points_per_s"] = 20000 / (time.perf_counter() - t0)
    # small-problem latency (BASELINE config 1 shape): n_obs = 50, D = 2, 10k candidates, through the plugin
    from metaopt_b200.algo import BayesianOptimizer
    t2 = transformed(real_space(2))
    algo = BayesianOptimizer(t2, seed=1, n_initial_points=5, acq_func="EI", n_points=10000, normalize_y=True)
    rs = np.random.RandomState(0)
    pts = [tuple(float(v) for v in rs.uniform(size=2)) for _ in range(50)]
    algo.observe(pts, [{"objective": float((p[0] - 0.3) ** 2 + (p[1] - 0.6) ** 2)} for p in pts])
    t0 = time.perf_counter()
    algo.suggest(1)
    first = (time.perf_counter() - t0) * 1e3
    lat = []
    for _ in range(30):
        t0 = time.perf_counter()
        algo.suggest(1)
        lat.append((time.perf_counter() - t0) * 1e3)
    algo.observe(pts[:1], [{"objective": 0.5}])
    t0 = time.perf_counter()
    algo.suggest(1)
    refit = (time.perf_counter() - t0) * 1e3
    out["cfg1_plugin_latency_ms"] = dict(first_suggest_with_hyperfit=first, p50_cached_fit=float(np.median(lat)),
                                         suggest_after_one_new_observation=refit, n_obs=50, n_points=10000)
    print(json.dumps(out), flush=True)
    if "--sampler-only" in sys.argv:
        return
    # ---- cfg3 / cfg4 scoring shapes (reduced candidate counts, same n and D)
    for tag, space, nobs, M in (("cfg3", cfg3_space(), 4000, 400_000), ("cfg4", real_space(100), 8000, 200_000)):
        comp = compiler.compile_space(transformed(space))
        compiler.upload(eng, comp)
        X = eng.sample(3, 0, nobs).cpu().numpy()
        y = ((X - 0.5) ** 2).sum(axis=1) + 0.01 * np.random.RandomState(0).randn(nobs)
        t0 = time.perf_counter()
        eng.fit(X, y, alpha=1e-10 + 1e-4, normalize_y=True)
        eng.sync()
        fit_ms = (time.perf_counter() - t0) * 1e3
        eng.score(None, M=20000, seed=5, acq_mask=1, y_opt=float(y.min()))
        ms = ev(lambda: eng.score(None, M=M, seed=5, first_index=0, acq_mask=1, y_opt=float(y.min())), 2)
        d = comp.d_out
        flops = M * (nobs * nobs + 2 * nobs * d + 2 * nobs)
        out[tag] = dict(n_obs=nobs, D=d, candidates=M, ms=ms, cand_per_s=M / ms * 1e3, tflops=flops / ms / 1e9, fit_ms=fit_ms,
                        includes="Philox candidate generation + transform + normalise inside the timed call")
        print(json.dumps(out), flush=True)
    json.dump(out, open("gpurun_out/bench_configs.json", "w"), indent=1)


if __name__ == "__main__":
    main()
