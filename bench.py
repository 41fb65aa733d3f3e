#!/usr/bin/env python
"""Benchmark of the GP-EI candidate-scoring hot path (BASELINE.json metric) on N B200s.

    python bench.py [--gpus N] [--steps K] [--warmup W] [--impl ours|reference]

A "step" = one pass of the hot path over one batch: score M = 1e6 candidates per GPU against a GP
fitted on n_obs = 2000 observations in d = 20 (BASELINE config 2: SURVEY.md 8d recipe) with EI and
return the argmax.  Weak scaling: every rank scores its own 1e6-candidate shard of the global
index range (no data-path collective); the winners are merged with one tiny all_gather.

Prints ONE JSON line (rank 0).  `value` = candidates/s with candidates resident in HBM; `e2e` = the
same through the host-buffer entry point (`ob200_score_host`: pinned host candidates -> H2D ->
score -> D2H winner inside the timed region).  `--impl reference` times the reference's CPU path
(scikit-learn GaussianProcessRegressor.predict + restated skopt gaussian_ei + argmin) on the
host cores, on a bounded sample of the same workload.
"""
import argparse
import json
import os
import statistics
import subprocess
import sys
import tempfile
import time

import numpy as np

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

N_OBS, DIM, M_PER_GPU = 2000, 20, 1_000_000
ALPHA, NOISE = 1e-10, 1e-4
XI = 0.01
METRIC = "GP-EI candidates/sec at n_obs=2k,d=20"
WORKLOAD = ("cfg2: GP-EI scoring, 20-D real space, n_obs=2000, 1M candidates per GPU, "
            "1.0*Matern52(ls=1), alpha=1e-10+noise 1e-4, normalize_y, xi=0.01")


def problem():
    rs = np.random.RandomState(0)
    X = rs.uniform(size=(N_OBS, DIM))
    y = ((X - 0.5) ** 2).sum(axis=1) + 0.01 * rs.randn(N_OBS)
    return X, y


def flops_per_candidate(n=N_OBS, d=DIM):
    """n^2 (triangular variance product) + 2 n D (distance GEMM) + 2 n (mean)."""
    return n * n + 2 * n * d + 2 * n


# ------------------------------------------------------------------------------------------------
class ClockSampler:
    QUERY = ("index,clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.active,"
             "clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown,"
             "clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap")

    def __init__(self, gpu_index=0):
        self.tmp = tempfile.NamedTemporaryFile("w+", suffix=".csv", delete=False)
        self.proc = None
        try:
            self.proc = subprocess.Popen(["nvidia-smi", "-i", str(gpu_index), "--query-gpu=" + self.QUERY,
                                          "--format=csv,noheader,nounits", "-lms", "100"],
                                         stdout=self.tmp, stderr=subprocess.DEVNULL)
        except Exception:
            self.proc = None

    def stop(self):
        out = {"sm_mhz": None, "sm_max_mhz": None, "reasons": [], "samples": 0}
        if self.proc is None:
            return out
        self.proc.terminate()
        try:
            self.proc.wait(timeout=5)
        except Exception:
            self.proc.kill()
        self.tmp.flush()
        self.tmp.seek(0)
        sm, mx, reasons = [], [], set()
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        for line in self.tmp.read().splitlines():
            parts = [p.strip() for p in line.split(",")]
            if len(parts) < 9:
                continue
            try:
                sm.append(float(parts[1]))
                mx.append(float(parts[2]))
            except ValueError:
                continue
            for name, val in zip(names, parts[5:9]):
                if val.lower().startswith("active"):
                    reasons.add(name)
        try:
            os.unlink(self.tmp.name)
        except OSError:
            pass
        if sm:
            out.update(sm_mhz=statistics.median(sm), sm_max_mhz=max(mx), reasons=sorted(reasons), samples=len(sm))
        return out


def cpu_reference_rate(X, y, candidates, repeats=1):
    """The reference's CPU path on `candidates`: sklearn predict(return_std) + restated EI + argmin."""
    from oracle import acquisition as oacq
    from oracle import gp as ogp
    import sklearn  # noqa: F401  (loads the BLAS/OpenMP runtimes that threadpoolctl then finds)
    try:
        from threadpoolctl import threadpool_info, threadpool_limits
        # torchrun exports OMP_NUM_THREADS=1; the CPU baseline is entitled to every host core
        threadpool_limits(limits=os.cpu_count() or 1)
        threads = max([p.get("num_threads", 1) for p in threadpool_info()] or [1])
    except Exception:
        threads = os.cpu_count() or 1
    model = ogp.SklearnGP(ogp.KERNEL_MATERN52, 1.0, np.ones(DIM), ALPHA + NOISE, True).fit(X, y)
    best = float("inf")
    pick = -1
    for _ in range(repeats):
        t0 = time.perf_counter()
        mu, sd = model.predict(candidates)
        pick = oacq.pick(oacq.ACQ_EI, mu, sd, float(y.min()), XI)
        best = min(best, time.perf_counter() - t0)
    return len(candidates) / best, threads, pick


def run_reference(args):
    """`--impl reference`: the CPU path with all host threads; rank 0 only."""
    if int(os.environ.get("RANK", "0")) != 0:
        return
    X, y = problem()
    m_step = 20_000  # bounded sample per step (~2 s of CPU work)
    cand = np.random.RandomState(1).uniform(size=(m_step, DIM))
    times = []
    threads = os.cpu_count() or 1
    for i in range(args.warmup + args.steps):
        rate, threads, _ = cpu_reference_rate(X, y, cand)
        if i >= args.warmup:
            times.append(m_step / rate)
    per_step = sum(times) / len(times)
    value = m_step / per_step
    sample = "%d of the 1e6 candidates per step (same X_obs, y, kernel); sklearn %s predict(return_std) + restated " \
             "skopt gaussian_ei + argmin" % (m_step, __import__("sklearn").__version__)
    line = {
        "impl": "reference", "metric": METRIC, "value": value, "unit": "candidates/s", "n_gpus": args.gpus,
        "steps": args.steps, "warmup": args.warmup, "ms_per_step": per_step * 1e3, "higher_is_better": True,
        "scaling": "weak", "vs_baseline": None, "dtype": "f64", "data": "synthetic",
        "config": {"workload": WORKLOAD, "sample_per_step": m_step},
        "cpu_baseline": {"value": value, "unit": "candidates/s", "cores": threads, "kind": "reference", "sample": sample},
        "e2e": {"value": value, "unit": "candidates/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "gpu_launches": 0,
    }
    emit(line)


def run_ours(args):
    import torch
    import torch.distributed as dist
    from metaopt_b200 import dist as obdist
    from metaopt_b200.engine import get_engine

    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    if not torch.cuda.is_available():
        raise SystemExit("bench.py needs a CUDA device: there is no CPU fallback (use --impl reference for the CPU path)")
    torch.cuda.set_device(local)
    multi = world > 1
    if multi:
        dist.init_process_group("nccl", device_id=torch.device("cuda", local))
    eng = get_engine(local)
    dev = eng.tdev

    X, y = problem()
    y_opt = float(y.min())
    # rank 0 fits, everyone else receives inv(L)/alpha/X by broadcast
    t0 = time.perf_counter()
    if rank == 0 or not multi:
        eng.fit(X, y, amplitude=1.0, length_scale=np.ones(DIM), alpha=ALPHA + NOISE, normalize_y=True)
    if multi:
        obdist.broadcast_fit(eng, N_OBS, DIM, src=0)
    eng.sync()
    fit_ms = (time.perf_counter() - t0) * 1e3

    # the space (for Philox candidate generation): 20 x Real uniform(0, 1)
    from metaopt_b200 import compiler
    from metaopt_b200.space import Real, Space, build_required_space
    space = Space()
    for i in range(DIM):
        space.register(Real("x%02d" % i, "uniform", 0, 1))
    comp = compiler.compile_space(build_required_space("real", space))
    compiler.upload(eng, comp)

    M = M_PER_GPU
    total = M * world
    start, count = obdist.shard_range(total, rank, world)
    seed = 1
    cand = eng.sample(seed, start, count)  # resident in HBM before the timed region (160 MB > L2)
    torch.cuda.synchronize()

    def step():
        rec = eng.score(cand, acq_mask=1, y_opt=y_opt, xi=XI, first_index=start)
        if multi:
            rec = obdist.allgather_best(rec, device=dev)
        return rec

    for _ in range(args.warmup):
        rec = step()
    if multi:
        dist.barrier()
    torch.cuda.synchronize()
    sampler = ClockSampler(local) if rank == 0 else None
    launches0 = eng.launches
    kern_ms = []
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(args.steps):
        a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        a.record()
        rec = step()
        b.record()
        kern_ms.append((a, b))
    e1.record()
    if multi:
        dist.barrier()
    torch.cuda.synchronize()
    elapsed_ms = e0.elapsed_time(e1)
    launches = eng.launches - launches0
    clocks = sampler.stop() if sampler else None
    if multi:
        elapsed_ms = obdist.max_over_ranks(elapsed_ms, device=dev)
    ms_per_step = elapsed_ms / args.steps
    value = total / (ms_per_step * 1e-3)
    k_ms = sum(a.elapsed_time(b) for a, b in kern_ms) / len(kern_ms)

    # ---- end to end through the host-buffer entry point (pinned host candidates) -----------------
    host_cand = torch.empty((count, DIM), dtype=torch.float64).pin_memory()
    host_cand.copy_(cand)
    host_np = host_cand.numpy()
    eng.score_host(host_np[:4096], acq_mask=1, y_opt=y_opt, xi=XI)
    e2e_steps = max(1, min(args.steps, 5))
    if multi:
        dist.barrier()
    torch.cuda.synchronize()
    t0 = time.perf_counter()
    for _ in range(e2e_steps):
        rec_h = eng.score_host(host_np, acq_mask=1, y_opt=y_opt, xi=XI)
        if multi:
            rec_h = obdist.allgather_best({k: dict(v, index=v["index"] + start) for k, v in rec_h.items()}, device=dev)
    torch.cuda.synchronize()
    e2e_s = (time.perf_counter() - t0) / e2e_steps
    if multi:
        e2e_s = obdist.max_over_ranks(e2e_s, device=dev)
    e2e_value = total / e2e_s

    # ---- exact pruned scoring (same winner, reported separately: `value` stays the dense number) --------
    pr_rec, n_scored = eng.score_pruned(cand, acq="EI", y_opt=y_opt, xi=XI, first_index=start)
    pr_steps = max(1, min(args.steps, 5))
    if multi:
        dist.barrier()
    torch.cuda.synchronize()
    p0, p1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    p0.record()
    for _ in range(pr_steps):
        pr_rec, n_scored = eng.score_pruned(cand, acq="EI", y_opt=y_opt, xi=XI, first_index=start)
        if multi:
            pr_rec = obdist.allgather_best(pr_rec, device=dev)
    p1.record()
    torch.cuda.synchronize()
    pr_ms = p0.elapsed_time(p1) / pr_steps
    if multi:
        pr_ms = obdist.max_over_ranks(pr_ms, device=dev)
    pruned = {"value": total / (pr_ms * 1e-3), "unit": "candidates/s", "ms_per_step": pr_ms,
              "fully_scored_fraction_rank0": n_scored / count, "same_winner_as_dense": pr_rec.get("EI") == rec.get("EI"),
              "note": "exact mean-only upper-bound pruning (ob200_score_pruned): identical argmax and values; "
                      "data-dependent, therefore NOT the headline value"}

    # ---- north_star target shape: ONE suggest over 1e6 candidates in total, sharded over the N GPUs (strong) -----
    strong = None
    if multi:
        s_start, s_count = obdist.shard_range(M_PER_GPU, rank, world)
        s_cand = cand[:s_count]  # any 1e6/N candidates of this rank: same work as its shard of a 1e6-candidate suggest

        def timed(fn, reps):
            fn()
            dist.barrier()
            torch.cuda.synchronize()
            a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            a.record()
            for _ in range(reps):
                fn()
            b.record()
            torch.cuda.synchronize()
            return obdist.max_over_ranks(a.elapsed_time(b) / reps, device=dev)
        dense_ms = timed(lambda: obdist.allgather_best(eng.score(s_cand, acq_mask=1, y_opt=y_opt, xi=XI, first_index=start),
                                                       device=dev), 10)
        pruned_ms = timed(lambda: obdist.allgather_best(eng.score_pruned(s_cand, acq="EI", y_opt=y_opt, xi=XI,
                                                                         first_index=start)[0], device=dev), 20)
        strong = {"candidates_total": M_PER_GPU, "candidates_per_gpu": s_count, "dense_ms_per_suggest": dense_ms,
                  "pruned_ms_per_suggest": pruned_ms,
                  "note": "1e6 candidates in total over all GPUs incl. the winner all_gather; target of north_star: < 10 ms on 8 GPUs"}

    # ---- suggest() latency through the plugin API (BaseAlgorithm.suggest), rank 0 only --------------------
    suggest_latency = None
    if rank == 0:
        from metaopt_b200.algo import BayesianOptimizer
        tspace = build_required_space("real", space)

        def plugin_latency(calls, acq_func="EI", **extra):
            algo = BayesianOptimizer(tspace, seed=1, n_initial_points=10, acq_func=acq_func, alpha=ALPHA, noise="gaussian",
                                     noise_level=NOISE, normalize_y=True, n_points=M, xi=XI, device=local,
                                     fit_hyperparameters=False, **extra)
            pts = [tuple(float(v) for v in row) for row in X]
            algo.observe(pts, [{"objective": float(v)} for v in y])
            algo.suggest(1)  # fit + first suggest
            lat = []
            for _ in range(calls):
                t1 = time.perf_counter()
                algo.suggest(1)
                lat.append((time.perf_counter() - t1) * 1e3)
            return statistics.median(lat), len(lat)
        p50_dense, n_dense = plugin_latency(max(5, min(20, args.steps * 2)), prune=False)
        p50_pruned, n_calls = plugin_latency(50, prune=True)
        p50_hedge, _ = plugin_latency(50, acq_func="gp_hedge", prune=True)  # the plugin's default portfolio (EI, LCB, PI)
        suggest_latency = {"p50_ms_dense": p50_dense, "p50_ms_pruned": p50_pruned, "p50_ms_pruned_gp_hedge": p50_hedge,
                           "calls_dense": n_dense, "calls_pruned": n_calls, "n_points": M,
                           "n_obs": N_OBS, "api": "BayesianOptimizer.suggest(1) (host wall clock, fit cached)"}

    if rank != 0:
        if multi:
            dist.destroy_process_group()
        return

    # ---- roofline of the dominant kernel (score_kernel: FP64 tensor-core bound) -------------------
    n = 8192
    A = torch.randn(n, n, dtype=torch.float64, device=dev)
    B = torch.randn(n, n, dtype=torch.float64, device=dev)
    torch.matmul(A, B)
    best = 1e30
    for _ in range(5):
        a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        a.record()
        torch.matmul(A, B)
        b.record()
        torch.cuda.synchronize()
        best = min(best, a.elapsed_time(b))
    dgemm_tflops = 2 * n ** 3 / best / 1e9
    del A, B
    achieved = count * flops_per_candidate() / (k_ms * 1e-3) / 1e12
    traffic = None
    tpath = os.path.join(ROOT, "profiles", "score_kernel_traffic.json")
    if os.path.exists(tpath):
        try:
            traffic = json.load(open(tpath)).get("dram_bytes_per_launch")
        except Exception:
            traffic = None
    roofline = {"bound": "tensor", "achieved": achieved, "peak": dgemm_tflops, "unit": "TFLOP/s",
                "frac": achieved / dgemm_tflops, "traffic": traffic,
                "kernel": "score_kernel (FP64 DMMA)", "flops_per_candidate": flops_per_candidate(),
                "peak_source": "cuBLAS DGEMM 8192^3 burst measured live in this run (MEASURED_PEAKS.json holds "
                               "only HBM and bf16 figures; nominal B200 FP64 is 37-40 TFLOP/s)"}

    # ---- CPU baseline on a bounded sample, rank 0 at N=1 only ---------------------------------------
    cpu_baseline = None
    if world == 1 and not args.no_cpu_baseline:
        m_cpu = 100_000
        sub = cand[:m_cpu].cpu().numpy()
        rate, threads, pick = cpu_reference_rate(X, y, sub)
        sub_rec = eng.score(cand[:m_cpu], acq_mask=1, y_opt=y_opt, xi=XI, first_index=start)
        cpu_baseline = {"value": rate, "unit": "candidates/s", "cores": threads, "kind": "reference",
                        "sample": "first %d of the 1e6 candidates of the timed step; sklearn %s "
                                  "GaussianProcessRegressor.predict(return_std) + restated skopt gaussian_ei + argmin"
                                  % (m_cpu, __import__("sklearn").__version__),
                        "argmax_matches_gpu": bool(pick + start == sub_rec["EI"]["index"])}

    line = {
        "metric": METRIC, "value": value, "unit": "candidates/s", "n_gpus": world, "steps": args.steps,
        "warmup": args.warmup, "ms_per_step": ms_per_step, "higher_is_better": True, "scaling": "weak",
        "vs_baseline": None, "dtype": "f64", "data": "synthetic",
        "config": {"workload": WORKLOAD, "n_obs": N_OBS, "d": DIM, "candidates_per_gpu": M, "candidates_total": total,
                   "acq": "EI", "parallelism": "candidate shards x%d, fit broadcast, all_gather argmax" % world,
                   "l2": "candidate inputs (160 MB/GPU) exceed L2; inv(L) (32 MB) is L2-resident by design",
                   "fit_ms": fit_ms},
        "clocks": clocks,
        "e2e": {"value": e2e_value, "unit": "candidates/s", "h2d_bytes_per_step": int(count * DIM * 8),
                "d2h_bytes_per_step": 96, "steps": e2e_steps, "ms_per_step": e2e_s * 1e3},
        "gpu_launches": int(launches),
        "roofline": roofline,
        "best": rec.get("EI"),
        "pruned": pruned,
        "suggest_latency": suggest_latency,
        "strong_scaling_1M": strong,
    }
    if cpu_baseline:
        line["cpu_baseline"] = cpu_baseline
    emit(line)
    if multi:
        dist.destroy_process_group()


_REAL_STDOUT = None


def emit(line):
    """Print the ONE JSON line on the real stdout (fd 1 is redirected to stderr while running, so that
    banners printed by native libraries such as NCCL cannot end up in the result stream)."""
    data = (json.dumps(line) + "\n").encode()
    if _REAL_STDOUT is not None:
        os.write(_REAL_STDOUT, data)
    else:
        sys.stdout.write(data.decode())
        sys.stdout.flush()


def main():
    global _REAL_STDOUT
    sys.stdout.flush()
    _REAL_STDOUT = os.dup(1)
    os.dup2(2, 1)
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=10)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--no-cpu-baseline", action="store_true")
    args = ap.parse_args()
    if args.warmup < 3 and args.impl == "ours":
        args.warmup = max(args.warmup, 0)
    if args.gpus > 1 and "WORLD_SIZE" not in os.environ:
        # convenience: `python bench.py --gpus N` relaunches itself under torchrun (one rank per GPU)
        cmd = [sys.executable, "-m", "torch.distributed.run", "--nnodes=1", "--nproc-per-node", str(args.gpus),
               "--master-addr", "127.0.0.1", "--master-port", os.environ.get("MASTER_PORT", "29531"),
               os.path.abspath(__file__)] + sys.argv[1:]
        raise SystemExit(subprocess.call(cmd))
    if args.impl == "reference":
        run_reference(args)
    else:
        run_ours(args)


if __name__ == "__main__":
    main()
