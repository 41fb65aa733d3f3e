#!/usr/bin/env python
"""Benchmark of the GP-EI candidate-scoring hot path (BASELINE.json metric) on N B200s.

    python bench.py [--gpus N] [--steps K] [--warmup W] [--impl ours|reference]

A "step" = one pass of the hot path over one batch: score M = 1e6 candidates per GPU against a GP
fitted on n_obs = 2000 observations in d = 20 (BASELINE config 2: SURVEY.md 8d recipe) with EI and
return the argmax.  Weak scaling: every rank scores its own 1e6-candidate shard of the global
index range (no data-path collective); the winners are merged with one tiny all_gather.

Prints ONE JSON line (rank 0).  `value` = candidates/s with candidates resident in HBM; `e2e` = the
same through the host-buffer entry point (`ob200_score_host`: pinned host candidates -> H2D ->
score -> D2H winner inside the timed region).  The other legs of the line (all measured in this
run, all driver-visible): `generation_ms` (the cand == NULL path's candidate generation), `pruned`
and `pruned_worst_case` (exact pruned scoring), `suggest_latency` (plugin API, incl. the
producer's deepcopy -> observe -> suggest cycle), `configs` (BASELINE configs 3, 4, 5 with their own
rooflines), and at N > 1 `sharding_proof`, `multi_engine_selftest`, `strong_scaling_1M`.
`--impl reference` times the reference's CPU path (scikit-learn GaussianProcessRegressor.predict +
restated skopt gaussian_ei + argmin) on the host cores, on a bounded sample of the same workload.
"""
import argparse
import copy
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
M_REFERENCE_STEP = 20_000   # candidates per step of the CPU arm (bounded sample; the CPU path is linear in M)
METRIC = "GP-EI candidates/sec at n_obs=2k,d=20"
WORKLOAD = ("cfg2: GP-EI scoring, 20-D real space, n_obs=2000, 1M candidates per GPU, "
            "1.0*Matern52(ls=1), alpha=1e-10+noise 1e-4, normalize_y, xi=0.01")


def problem(n=N_OBS, d=DIM):
    rs = np.random.RandomState(0)
    X = rs.uniform(size=(n, d))
    y = ((X - 0.5) ** 2).sum(axis=1) + 0.01 * rs.randn(n)
    return X, y


def flops_per_candidate(n=N_OBS, d=DIM):
    """n^2 (triangular variance product) + 2 n D (distance GEMM) + 2 n (mean)."""
    return n * n + 2 * n * d + 2 * n


def measured_peaks():
    path = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(path):
        try:
            return json.load(open(path)), "MEASURED_PEAKS.json (driver-measured copy bandwidth)"
        except Exception:
            pass
    return {"hbm_gbs": 6650.0}, "fallback stated in B200_PROFILING.md (MEASURED_PEAKS.json absent)"


# ---- the BASELINE spaces (mirror classes of orion.algo.space) -------------------------------------------------
def real_space(d):
    from metaopt_b200.space import Real, Space
    s = Space()
    for i in range(d):
        s.register(Real("x%03d" % i, "uniform", 0, 1))
    return s


def cfg3_space():
    """10 Integer uniform(0,100), 10 Categorical k=4, 10 Real uniform(0,1) -> D = 60 after one-hot."""
    from metaopt_b200.space import Categorical, Integer, Real, Space
    s = Space()
    for i in range(10):
        s.register(Integer("i%02d" % i, "uniform", 0, 100))
        s.register(Categorical("c%02d" % i, ["a", "b", "c", "d"]))
        s.register(Real("r%02d" % i, "uniform", 0, 1))
    return s


def cfg5_space():
    """9 uniform + 8 loguniform reals, 17 integers, 16 categoricals k=4 (half non-uniform) -> D_in 50, D_out 98."""
    from metaopt_b200.space import Categorical, Integer, Real, Space
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
    return s


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


def workload_config(sample_per_step, **extra):
    """The `config` both arms print: same workload string, same keys; `sample_per_step` says how many of the
    1e6 candidates of a step the arm actually scores per step (the CPU arm scores a bounded sample)."""
    cfg = {"workload": WORKLOAD, "n_obs": N_OBS, "d": DIM, "candidates_per_gpu": M_PER_GPU, "acq": "EI",
           "sample_per_step": int(sample_per_step)}
    cfg.update(extra)
    return cfg


def run_reference(args):
    """`--impl reference`: the CPU path with all host threads; rank 0 only."""
    if int(os.environ.get("RANK", "0")) != 0:
        return
    X, y = problem()
    m_step = M_REFERENCE_STEP
    cand = np.random.RandomState(1).uniform(size=(m_step, DIM))
    times = []
    threads = os.cpu_count() or 1
    for i in range(args.warmup + args.steps):
        rate, threads, _ = cpu_reference_rate(X, y, cand)
        if i >= args.warmup:
            times.append(m_step / rate)
    per_step = sum(times) / len(times)
    value = m_step / per_step
    sample = "%d of the 1e6 candidates per step (same X_obs, y, kernel; the CPU path is linear in the candidate count); " \
             "sklearn %s predict(return_std) + restated skopt gaussian_ei + argmin" % (m_step, __import__("sklearn").__version__)
    line = {
        "impl": "reference", "metric": METRIC, "value": value, "unit": "candidates/s", "n_gpus": args.gpus,
        "steps": args.steps, "warmup": args.warmup, "ms_per_step": per_step * 1e3, "higher_is_better": True,
        "scaling": "weak", "vs_baseline": None, "dtype": "f64", "data": "synthetic",
        "config": workload_config(m_step),
        "cpu_baseline": {"value": value, "unit": "candidates/s", "cores": threads, "kind": "reference", "sample": sample},
        "e2e": {"value": value, "unit": "candidates/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "gpu_launches": 0,
    }
    emit(line)


# ------------------------------------------------------------------------------------------------
def dgemm_tflops(torch, dev):
    """Live cuBLAS DGEMM rate (8192^3, best of 5): the FP64 tensor roofline denominator of this run."""
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
    del A, B
    return 2 * n ** 3 / best / 1e9


def event_ms(torch, fn, reps=1):
    torch.cuda.synchronize()
    a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    a.record()
    for _ in range(reps):
        fn()
    b.record()
    torch.cuda.synchronize()
    return a.elapsed_time(b) / reps


def run_ours(args):
    import torch
    import torch.distributed as dist
    from metaopt_b200 import compiler
    from metaopt_b200 import dist as obdist
    from metaopt_b200.engine import get_engine
    from metaopt_b200.space import build_required_space

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

    def barrier():
        if multi:
            dist.barrier()
        torch.cuda.synchronize()

    def rmax(v):
        return obdist.max_over_ranks(v, device=dev) if multi else v

    X, y = problem()
    y_opt = float(y.min())
    # rank 0 fits, everyone else receives inv(L)/alpha/X by broadcast
    fit_ms = None
    if rank == 0 or not multi:
        eng.fit(X, y, amplitude=1.0, length_scale=np.ones(DIM), alpha=ALPHA + NOISE, normalize_y=True)  # warm (allocations)
        eng.sync()
        t0 = time.perf_counter()
        eng.fit(X, y, amplitude=1.0, length_scale=np.ones(DIM), alpha=ALPHA + NOISE, normalize_y=True)
        eng.sync()
        fit_ms = (time.perf_counter() - t0) * 1e3
    first_broadcast_ms = fit_broadcast_ms = None
    if multi:
        t0 = time.perf_counter()
        obdist.broadcast_fit(eng, N_OBS, DIM, src=0)
        eng.sync()
        first_broadcast_ms = (time.perf_counter() - t0) * 1e3     # includes NCCL communicator start-up
        barrier()
        t0 = time.perf_counter()
        obdist.broadcast_fit(eng, N_OBS, DIM, src=0)
        eng.sync()
        fit_broadcast_ms = rmax((time.perf_counter() - t0) * 1e3)

    # the space (for Philox candidate generation): 20 x Real uniform(0, 1)
    space = real_space(DIM)
    tspace = build_required_space("real", space)
    comp = compiler.compile_space(tspace)
    compiler.upload(eng, comp)

    M = M_PER_GPU
    total = M * world
    start, count = obdist.shard_range(total, rank, world)
    seed = 1
    cand = eng.sample(seed, start, count)  # resident in HBM before the timed region (160 MB > L2)
    torch.cuda.synchronize()
    generation_ms = rmax(event_ms(torch, lambda: eng.sample(seed, start, count, out=cand, check=False), 5))

    def step():
        rec = eng.score(cand, acq_mask=1, y_opt=y_opt, xi=XI, first_index=start)
        if multi:
            rec = obdist.allgather_best(rec, device=dev)
        return rec

    for _ in range(args.warmup):
        rec = step()
    barrier()
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
    barrier()
    elapsed_ms = e0.elapsed_time(e1)
    launches = eng.launches - launches0
    clocks = sampler.stop() if sampler else None
    elapsed_ms = rmax(elapsed_ms)
    ms_per_step = elapsed_ms / args.steps
    value = total / (ms_per_step * 1e-3)
    k_ms = sum(a.elapsed_time(b) for a, b in kern_ms) / len(kern_ms)
    my_winner = eng.score(cand, acq_mask=1, y_opt=y_opt, xi=XI, first_index=start)["EI"]

    # ---- end to end through the host-buffer entry point (pinned host candidates) -----------------
    host_cand = torch.empty((count, DIM), dtype=torch.float64).pin_memory()
    host_cand.copy_(cand)
    host_np = host_cand.numpy()
    # untimed warm-up at the timed size: the first full-size call grows the context's candidate buffer (160 MB) and creates
    # its copy stream, which used to land inside the timed steps (+3 ms per step over five steps)
    eng.score_host(host_np[:4096], acq_mask=1, y_opt=y_opt, xi=XI)
    for _ in range(min(args.warmup, 2)):
        eng.score_host(host_np, acq_mask=1, y_opt=y_opt, xi=XI)
    e2e_steps = max(1, min(args.steps, 5))
    barrier()
    t0 = time.perf_counter()
    for _ in range(e2e_steps):
        rec_h = eng.score_host(host_np, acq_mask=1, y_opt=y_opt, xi=XI)
        if multi:
            rec_h = obdist.allgather_best({k: dict(v, index=v["index"] + start) for k, v in rec_h.items()}, device=dev)
    torch.cuda.synchronize()
    e2e_s = rmax((time.perf_counter() - t0) / e2e_steps)
    e2e_value = total / e2e_s

    # ---- exact pruned scoring (same winner, reported separately: `value` stays the dense number) --------
    pr_rec, n_scored = eng.score_pruned(cand, acq="EI", y_opt=y_opt, xi=XI, first_index=start)
    pr_steps = max(3, min(args.steps, 10))
    barrier()

    def pruned_step():
        r, _ = eng.score_pruned(cand, acq="EI", y_opt=y_opt, xi=XI, first_index=start)
        return obdist.allgather_best(r, device=dev) if multi else r
    pr_ms = rmax(event_ms(torch, pruned_step, pr_steps))
    pr_rec = pruned_step()
    st = eng.prune_stats()
    os.environ["OB200_BOUND_F32"] = "0"
    eng.score_pruned(cand, acq="EI", y_opt=y_opt, xi=XI, first_index=start)
    pr64_ms = rmax(event_ms(torch, pruned_step, pr_steps))
    st64 = eng.prune_stats()
    os.environ.pop("OB200_BOUND_F32")
    pruned = {"value": total / (pr_ms * 1e-3), "unit": "candidates/s", "ms_per_step": pr_ms,
              "fully_scored_fraction_rank0": n_scored / count, "same_winner_as_dense": pr_rec.get("EI") == rec.get("EI"),
              "bound_pass": "fp32/tf32 tensor cores with a rigorous error budget; survivors re-scored in FP64",
              "batches_rank0": st, "ms_per_step_fp64_bound_pass": pr64_ms, "batches_rank0_fp64_bound_pass": st64,
              "note": "exact upper-bound pruning (ob200_score_pruned): identical argmax and values; "
                      "data-dependent, therefore NOT the headline value"}

    # ---- N > 1: the sharded winner equals the single-GPU winner of the whole global range (rank 0 re-scores it) ----
    sharding_proof = None
    if multi:
        gathered = [None] * world
        dist.all_gather_object(gathered, my_winner)
        if rank == 0:
            t0 = time.perf_counter()
            whole, _ = eng.score_pruned(None, M=total, seed=seed, first_index=0, acq="EI", y_opt=y_opt, xi=XI)
            whole_ms = (time.perf_counter() - t0) * 1e3
            merged = obdist.merge_best([{"EI": g} for g in gathered])
            sharding_proof = {"sharded_winner_equals_single": merged["EI"] == whole["EI"] and rec.get("EI") == whole["EI"],
                              "single_gpu_winner_of_global_range": whole["EI"], "per_rank_winners": gathered,
                              "single_gpu_rescore_ms": whole_ms,
                              "how": "rank 0 regenerates all %d candidates from (seed, index) and scores them alone "
                                     "(exact pruned path, bit-identical to dense)" % total}
        barrier()

    # ---- north_star target shape: ONE suggest over 1e6 candidates in total, sharded over the N GPUs (strong) -----
    strong = None
    if multi:
        s_start, s_count = obdist.shard_range(M_PER_GPU, rank, world)
        s_cand = cand[:s_count]  # any 1e6/N candidates of this rank: same work as its shard of a 1e6-candidate suggest

        def timed(fn, reps):
            fn()
            barrier()
            return rmax(event_ms(torch, fn, reps))
        dense_ms = timed(lambda: obdist.allgather_best(eng.score(s_cand, acq_mask=1, y_opt=y_opt, xi=XI, first_index=start),
                                                       device=dev), 10)
        pruned_ms = timed(lambda: obdist.allgather_best(eng.score_pruned(s_cand, acq="EI", y_opt=y_opt, xi=XI,
                                                                         first_index=start)[0], device=dev), 20)
        strong = {"candidates_total": M_PER_GPU, "candidates_per_gpu": s_count, "dense_ms_per_suggest": dense_ms,
                  "pruned_ms_per_suggest": pruned_ms,
                  "note": "1e6 candidates in total over all GPUs incl. the winner all_gather; target of north_star: < 10 ms on 8 GPUs"}

    # ---- the plugin's own multi-GPU mode (one process, MultiEngine) against the single engine, rank 0 drives all GPUs ----
    multi_selftest = None
    if multi:
        if rank == 0:
            try:
                from metaopt_b200.multi import get_multi_engine
                me = get_multi_engine(list(range(world)))
                compiler.upload(me, comp)
                me.fit(X, y, amplitude=1.0, length_scale=np.ones(DIM), alpha=ALPHA + NOISE, normalize_y=True)
                single = eng.score(None, M=M, seed=7, first_index=123, acq_mask=7, y_opt=y_opt, xi=XI)
                sharded = me.score(None, M=M, seed=7, first_index=123, acq_mask=7, y_opt=y_opt, xi=XI)
                sp, _ = eng.score_pruned_multi(None, M=M, seed=7, first_index=123, acq_mask=7, y_opt=y_opt, xi=XI)
                mp, _ = me.score_pruned_multi(None, M=M, seed=7, first_index=123, acq_mask=7, y_opt=y_opt, xi=XI)
                multi_selftest = {"devices": world, "dense_equal": sharded == single, "pruned_equal": mp == sp and sp == single,
                                  "what": "MultiEngine (one process, one context per GPU, thread pool) vs the single engine; "
                                          "correctness only -- the other GPUs are shared with this job's idle ranks, so no "
                                          "latency is quoted from here (tools/multi_latency.py measures it alone)"}
            except Exception as exc:  # the bench line must still be printed
                multi_selftest = {"devices": world, "error": repr(exc)}
        barrier()

    # ---- worst cases of the pruned path (rank 0): weak bounds, saturated PI, wide LCB ----------------------------
    worst = None
    if rank == 0:
        worst = {}
        sub = cand[:500_000]

        def leg(name, acq, **kw):
            d_ms = event_ms(torch, lambda: eng.score(sub, acq_mask=1 << {"EI": 0, "PI": 1, "LCB": 2}[acq], y_opt=y_opt, xi=XI,
                                                     first_index=start, **kw), 1)
            eng.score_pruned(sub, acq=acq, y_opt=y_opt, xi=XI, first_index=start, **kw)
            p_ms = event_ms(torch, lambda: eng.score_pruned(sub, acq=acq, y_opt=y_opt, xi=XI, first_index=start, **kw), 2)
            p, ns = eng.score_pruned(sub, acq=acq, y_opt=y_opt, xi=XI, first_index=start, **kw)
            d = eng.score(sub, acq_mask=1 << {"EI": 0, "PI": 1, "LCB": 2}[acq], y_opt=y_opt, xi=XI, first_index=start, **kw)
            worst[name] = {"dense_ms": d_ms, "pruned_ms": p_ms, "fully_scored_fraction": ns / sub.shape[0],
                           "same_winner": p[acq] == d[acq], "batches": eng.prune_stats()}
        leg("PI_cfg2", "PI")
        leg("LCB_kappa10_cfg2", "LCB", kappa=10.0)
        # long length scales (what a fitted GP on a smooth objective looks like): the nearest-observation bound is weak
        eng.fit(X, y, amplitude=1.0, length_scale=4.0 * np.ones(DIM), alpha=ALPHA + NOISE, normalize_y=True)
        leg("EI_length_scale_4", "EI")
        worst["candidates"] = int(sub.shape[0])
        eng.fit(X, y, amplitude=1.0, length_scale=np.ones(DIM), alpha=ALPHA + NOISE, normalize_y=True)
    if multi:
        obdist.broadcast_fit(eng, N_OBS, DIM, src=0)
        barrier()

    # ---- suggest() latency through the plugin API (BaseAlgorithm.suggest), rank 0 only --------------------
    suggest_latency = None
    if rank == 0:
        from metaopt_b200.algo import BayesianOptimizer
        pts = [tuple(float(v) for v in row) for row in X]
        rs = np.random.RandomState(5)

        def make(acq_func="EI", n_points=M, **extra):
            algo = BayesianOptimizer(tspace, seed=1, n_initial_points=10, acq_func=acq_func, alpha=ALPHA, noise="gaussian",
                                     noise_level=NOISE, normalize_y=True, n_points=n_points, xi=XI, device=local, **extra)
            algo.observe(pts, [{"objective": float(v)} for v in y])
            return algo

        def cached(calls, **kw):
            algo = make(fit_hyperparameters=False, **kw)
            algo.suggest(1)  # fit + first suggest
            lat = []
            for _ in range(calls):
                t1 = time.perf_counter()
                algo.suggest(1)
                lat.append((time.perf_counter() - t1) * 1e3)
            return statistics.median(lat), len(lat)

        def producer_cycle(calls, **kw):
            """What Orion's producer does for every suggestion (`producer.py:103-132,159-174`): the real algorithm
            observes the newly completed trial, a deep copy of it suggests, the copy's state comes back."""
            algo = make(**kw)
            algo.suggest(1)  # cold: first fit (and hyper-parameter search)
            lat = []
            for _ in range(calls):
                x_new = tuple(float(v) for v in rs.uniform(size=DIM))
                y_new = float(((np.asarray(x_new) - 0.5) ** 2).sum() + 0.01 * rs.randn())
                t1 = time.perf_counter()
                algo.observe([x_new], [{"objective": y_new}])
                naive = copy.deepcopy(algo)
                naive.suggest(1)
                algo.set_state(naive.state_dict)
                lat.append((time.perf_counter() - t1) * 1e3)
            return statistics.median(lat), max(lat), len(lat)
        p50_dense, n_dense = cached(max(5, min(20, args.steps * 2)), prune=False)
        p50_pruned, n_calls = cached(50, prune=True)
        p50_hedge, _ = cached(50, acq_func="gp_hedge", prune=True)  # the plugin's default portfolio (EI, LCB, PI)
        # the plugin's DEFAULT candidate count (n_points = 10 000, skopt's default), EI and the default portfolio
        p50_10k, _ = cached(50, n_points=10_000, prune=True)
        p50_10k_hedge, _ = cached(50, n_points=10_000, acq_func="gp_hedge", prune=True)
        p50_fixed, max_fixed, n_cycle = producer_cycle(50, fit_hyperparameters=False, prune=True)
        p50_default, max_default, _ = producer_cycle(50, acq_func="gp_hedge", fit_hyperparameters=True, prune=True)
        suggest_latency = {"p50_ms_dense": p50_dense, "p50_ms_pruned": p50_pruned, "p50_ms_pruned_gp_hedge": p50_hedge,
                           "p50_ms_n_points_10000": p50_10k, "p50_ms_n_points_10000_gp_hedge": p50_10k_hedge,
                           "p50_ms_observe_then_suggest_fixed_theta": p50_fixed, "max_ms_observe_then_suggest_fixed_theta": max_fixed,
                           "p50_ms_observe_then_suggest_default": p50_default, "max_ms_observe_then_suggest_default": max_default,
                           "calls_dense": n_dense, "calls_pruned": n_calls, "calls_observe_then_suggest": n_cycle, "n_points": M,
                           "n_obs": N_OBS,
                           "api": "BayesianOptimizer (host wall clock). p50_ms_*: suggest(1) with the fit cached. "
                                  "observe_then_suggest: observe(1 new trial) -> copy.deepcopy -> copy.suggest(1) -> "
                                  "set_state, as Orion's producer does; fixed_theta = fit_hyperparameters false (O(n^2) "
                                  "append), default = the plugin's defaults (gp_hedge, hyper-parameters re-fitted by L-BFGS-B "
                                  "on the device LML gradient, warm-started)"}

    # ---- the other BASELINE configs, each with its own roofline --------------------------------------------------
    peaks, peak_src = measured_peaks()
    dgemm = dgemm_tflops(torch, dev) if rank == 0 or multi else None
    configs = bench_other_configs(torch, eng, comp, rank, world, multi, dgemm, peaks, peak_src, barrier, rmax, obdist, compiler,
                                  build_required_space)
    # restore the cfg2 space and fit on the engine (the legs above registered other spaces and fits)
    compiler.upload(eng, comp)
    eng.fit(X, y, amplitude=1.0, length_scale=np.ones(DIM), alpha=ALPHA + NOISE, normalize_y=True)

    if rank != 0:
        if multi:
            dist.destroy_process_group()
        return

    # ---- roofline of the dominant kernel (score_kernel: FP64 tensor-core bound) -------------------
    achieved = count * flops_per_candidate() / (k_ms * 1e-3) / 1e12
    traffic = None
    tpath = os.path.join(ROOT, "profiles", "score_kernel_traffic.json")
    if os.path.exists(tpath):
        try:
            traffic = json.load(open(tpath)).get("dram_bytes_per_launch")
        except Exception:
            traffic = None
    roofline = {"bound": "tensor", "achieved": achieved, "peak": dgemm, "unit": "TFLOP/s",
                "frac": achieved / dgemm, "traffic": traffic,
                "kernel": "score_kernel (FP64 DMMA)", "flops_per_candidate": flops_per_candidate(),
                "peak_source": "of live cuBLAS DGEMM: 8192^3 burst measured in this run (MEASURED_PEAKS.json holds "
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
        "config": workload_config(M, candidates_total=total,
                                  parallelism="candidate shards x%d, fit broadcast, all_gather argmax" % world,
                                  l2="candidate inputs (160 MB/GPU) exceed L2; inv(L) (32 MB) is L2-resident by design",
                                  fit_ms=fit_ms, fit_broadcast_ms=fit_broadcast_ms,
                                  nccl_init_plus_first_broadcast_ms=first_broadcast_ms),
        "clocks": clocks,
        "e2e": {"value": e2e_value, "unit": "candidates/s", "h2d_bytes_per_step": int(count * DIM * 8),
                "d2h_bytes_per_step": 96, "steps": e2e_steps, "ms_per_step": e2e_s * 1e3},
        "gpu_launches": int(launches),
        "roofline": roofline,
        "best": rec.get("EI"),
        "generation_ms": generation_ms,
        "value_incl_generation": total / ((ms_per_step + generation_ms) * 1e-3),
        "pruned": pruned,
        "pruned_worst_case": worst,
        "suggest_latency": suggest_latency,
        "configs": configs,
        "sharding_proof": sharding_proof,
        "multi_engine_selftest": multi_selftest,
        "strong_scaling_1M": strong,
    }
    if cpu_baseline:
        line["cpu_baseline"] = cpu_baseline
    emit(line)
    if multi:
        dist.destroy_process_group()


def bench_other_configs(torch, eng, comp2, rank, world, multi, dgemm, peaks, peak_src, barrier, rmax, obdist, compiler,
                        build_required_space):
    """BASELINE configs 3, 4 (scoring shapes, FP64 tensor roofline) and 5 (sampling + transform, HBM roofline), one timed
    pass each.  Every rank runs the same per-GPU work (weak scaling, like the headline); times are the max over ranks."""
    out = {}
    hbm = float(peaks.get("hbm_gbs", 6650.0))
    # ---- cfg5: Space.sample + transform, 50-D mixed -> 98 columns; 1e9 points over all GPUs in chunks of 2^24 -----
    c5 = compiler.compile_space(build_required_space("real", cfg5_space()))
    compiler.upload(eng, c5)
    chunk = 1 << 24
    buf = torch.empty((chunk, c5.d_out), dtype=torch.float64, device=eng.tdev)
    n_points = 1_000_000_000
    p_start, p_count = obdist.shard_range(n_points, rank, world)
    eng.sample(7, 0, chunk, out=buf)
    barrier()

    def all_points():
        done = 0
        while done < p_count:
            c = min(chunk, p_count - done)
            eng.sample(7, p_start + done, c, out=buf[:c], check=False)
            done += c
    ms5 = rmax(event_ms(torch, all_points, 1))
    eng.sample(7, 0, 16, out=buf[:16], check=True)  # the latched status of the whole run surfaces here
    bytes5 = n_points * c5.d_out * 8
    gbs_per_gpu = bytes5 / world / ms5 / 1e6
    traffic5 = None   # DRAM bytes per launch (one 2^24-point chunk), scaled from the ncu capture's bytes per point
    tp = os.path.join(ROOT, "profiles", "sampler_traffic.json")
    if os.path.exists(tp):
        try:
            traffic5 = float(json.load(open(tp))["dram_bytes_per_point"]) * min(chunk, p_count)
        except Exception:
            traffic5 = None
    out["cfg5"] = {"workload": "cfg5: Space.sample + transform + normalise, 50-D mixed space (17 Real, 17 Integer, 16 Categorical "
                               "k=4) -> 98 float64 columns, 1e9 points over %d GPU(s) in chunks of 2^24" % world,
                   "points": n_points, "ms": ms5, "points_per_s": n_points / ms5 * 1e3,
                   "roofline": {"bound": "hbm", "achieved": gbs_per_gpu, "peak": hbm, "unit": "GB/s", "frac": gbs_per_gpu / hbm,
                                "traffic": traffic5, "kernel": "sample_transform_kernel", "bytes_per_point": c5.d_out * 8,
                                "algorithmic_bytes_per_launch": c5.d_out * 8 * min(chunk, p_count),
                                "per": "GPU", "peak_source": peak_src}}
    del buf
    # ---- cfg3 / cfg4 shape: dense scoring, candidates generated on the device inside the timed call ------------------
    for tag, space, nobs, m_gpu, label in (
            ("cfg3", cfg3_space(), 4000, 4_000_000 // max(world, 1) if multi else 1_000_000,
             "cfg3: mixed 30-D space (10 Integer, 10 Categorical one-hot, 10 Real) -> D=60, n_obs=4000"),
            ("cfg4_shape", real_space(100), 8000, 2_000_000 if multi else 500_000,
             "cfg4: 100-D real space, n_obs=8000, 2e6 candidates per GPU (16e6 over 8)")):
        c = compiler.compile_space(build_required_space("real", space))
        compiler.upload(eng, c)
        Xo = eng.sample(3, 0, nobs).cpu().numpy()
        yo = ((Xo - 0.5) ** 2).sum(axis=1) + 0.01 * np.random.RandomState(0).randn(nobs)
        if multi:  # identical observations everywhere (same sampler), fit redundantly: no broadcast needed for a bench leg
            pass
        t0 = time.perf_counter()
        eng.fit(Xo, yo, alpha=ALPHA + NOISE, normalize_y=True)
        eng.sync()
        fit_first_ms = (time.perf_counter() - t0) * 1e3   # includes growing the context's workspaces to this size
        t0 = time.perf_counter()
        eng.fit(Xo, yo, alpha=ALPHA + NOISE, normalize_y=True)
        eng.sync()
        fit_ms = (time.perf_counter() - t0) * 1e3         # host arrays in, factor + inverse + alpha on the device
        yo_opt = float(yo.min())
        s0, _ = obdist.shard_range(m_gpu * world, rank, world)
        eng.score(None, M=20_000, seed=5, first_index=s0, acq_mask=1, y_opt=yo_opt)
        barrier()
        holder = {}

        def one():
            r = eng.score(None, M=m_gpu, seed=5, first_index=s0, acq_mask=1, y_opt=yo_opt)
            holder["rec"] = obdist.allgather_best(r, device=eng.tdev) if multi else r
        ms = rmax(event_ms(torch, one, 1))
        pr, ns = eng.score_pruned(None, M=m_gpu, seed=5, first_index=s0, acq="EI", y_opt=yo_opt)
        ms_p = rmax(event_ms(torch, lambda: eng.score_pruned(None, M=m_gpu, seed=5, first_index=s0, acq="EI", y_opt=yo_opt), 1))
        local_dense = eng.score(None, M=min(m_gpu, 200_000), seed=5, first_index=s0, acq_mask=1, y_opt=yo_opt)
        local_pruned, _ = eng.score_pruned(None, M=min(m_gpu, 200_000), seed=5, first_index=s0, acq="EI", y_opt=yo_opt)
        d = c.d_out
        tfl = m_gpu * flops_per_candidate(nobs, d) / ms / 1e9
        out[tag] = {"workload": label, "n_obs": nobs, "D": d, "candidates_per_gpu": m_gpu, "candidates_total": m_gpu * world,
                    "ms": ms, "candidates_per_s": m_gpu * world / ms * 1e3, "fit_ms": fit_ms, "fit_first_call_ms": fit_first_ms,
                    "includes": "Philox candidate generation + transform + normalise inside the timed call",
                    "roofline": {"bound": "tensor", "achieved": tfl, "peak": dgemm, "unit": "TFLOP/s",
                                 "frac": tfl / dgemm if dgemm else None, "per": "GPU", "flops_per_candidate": flops_per_candidate(nobs, d),
                                 "peak_source": "of live cuBLAS DGEMM (8192^3 burst, this run)"},
                    "pruned_ms": ms_p, "pruned_fully_scored_fraction": ns / m_gpu,
                    "pruned_equals_dense_on_200k": local_dense["EI"] == local_pruned["EI"]}
    return out


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
    if args.impl == "ours":
        args.warmup = max(args.warmup, 3)   # timing rule: at least three warm-up steps
    if args.gpus > 1 and "WORLD_SIZE" not in os.environ:
        # convenience: `python bench.py --gpus N` relaunches itself under torchrun (one rank per GPU)
        cmd = [sys.executable, "-m", "torch.distributed.run", "--nnodes=1", "--nproc-per-node", str(args.gpus),
               "--master-addr", "127.0.0.1", "--master-port", os.environ.get("MASTER_PORT", "29531"),
               os.path.abspath(__file__)] + sys.argv[1:]
        raise SystemExit(subprocess.call(cmd, stdout=_REAL_STDOUT))   # the ranks inherit the REAL stdout for the JSON line
    if args.impl == "reference":
        run_reference(args)
    else:
        run_ours(args)


if __name__ == "__main__":
    main()
