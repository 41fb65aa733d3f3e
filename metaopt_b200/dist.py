"""Multi-GPU sharding of the candidate-scoring path: one process per GPU (torchrun), candidates
partitioned by global index, the fit broadcast once, one tiny exchange for the argmax.

Candidates are independent, so the data path has no collective: rank r scores the contiguous index
range `shard_range(M, r, world)`; candidate i is a pure function of (seed, i) (Philox), hence the
candidate set -- and the argmax -- is identical for 1/2/4/8 GPUs.  The only exchanges are
  * `broadcast_fit`: rank 0's fit state (inv(L) tiles, alpha, X/ls, y statistics) as one blob
    through `torch.distributed.broadcast` (NCCL over NVLink on GPUs);
  * `allgather_best`: 3 x (value, index, mu, sigma) per rank through `all_gather`, followed by the
    same deterministic merge on every rank (NCCL has no MAXLOC; ties -> lowest global index, i.e.
    numpy argmin/argmax first-occurrence semantics of the reference path).
The helpers only need an initialised process group; with the `gloo` backend they run on CPU
tensors, which is how the N>1 host logic is tested without GPUs.
"""
import numpy as np
import torch
import torch.distributed as dist

ACQ_NAMES = ("EI", "PI", "LCB")


def shard_range(total, rank, world):
    """Contiguous split of [0, total) into `world` parts whose sizes differ by at most one."""
    if world < 1 or not 0 <= rank < world:
        raise ValueError("bad rank/world: %r/%r" % (rank, world))
    base, extra = divmod(int(total), world)
    start = rank * base + min(rank, extra)
    return start, base + (1 if rank < extra else 0)


def better(a, b):
    """True if winner record `a` beats `b`: larger utility, ties to the lower global index."""
    return a["value"] > b["value"] or (a["value"] == b["value"] and a["index"] < b["index"])


def merge_best(records):
    """Deterministic reduction of per-shard winners {acq: {value,index,mu,sigma}} (None = empty shard)."""
    out = {}
    for rec in records:
        if not rec:
            continue
        for name, r in rec.items():
            if name not in ACQ_NAMES:
                continue
            if name not in out or better(r, out[name]):
                out[name] = dict(r)
    return out


def pack_best(rec, device="cpu"):
    """{acq: record} -> float64 tensor [3, 4]; the int64 index travels bit-exactly in a float64 slot."""
    t = torch.zeros((3, 4), dtype=torch.float64)
    idx = torch.full((3,), -1, dtype=torch.int64)
    t[:, 0] = -float("inf")
    for k, name in enumerate(ACQ_NAMES):
        if rec and name in rec:
            r = rec[name]
            t[k, 0], t[k, 2], t[k, 3] = r["value"], r["mu"], r["sigma"]
            idx[k] = int(r["index"])
    t[:, 1] = idx.view(torch.float64)
    return t.to(device)


def unpack_best(t):
    t = t.detach().cpu().contiguous()
    idx = t[:, 1].contiguous().view(torch.int64)
    rec = {}
    for k, name in enumerate(ACQ_NAMES):
        if int(idx[k]) >= 0:
            rec[name] = dict(value=float(t[k, 0]), index=int(idx[k]), mu=float(t[k, 2]), sigma=float(t[k, 3]))
    return rec


def allgather_best(rec, device="cpu", group=None):
    """The 'argmax all-reduce': gather every rank's winners and merge identically everywhere."""
    mine = pack_best(rec, device)
    world = dist.get_world_size(group)
    gathered = [torch.empty_like(mine) for _ in range(world)]
    dist.all_gather(gathered, mine, group=group)
    return merge_best([unpack_best(g) for g in gathered])


def broadcast_fit(engine, n, D, src=0, group=None):
    """Rank `src` holds a fit; every other rank receives it (inv(L), alpha, X/ls, y stats)."""
    rank = dist.get_rank(group)
    if rank == src:
        blob = engine.export_state()
    else:
        blob = torch.empty(engine.state_bytes(n, D), dtype=torch.uint8, device=engine.tdev)
    dist.broadcast(blob, src=src, group=group)
    if rank != src:
        engine.import_state(blob, n, D)
    return blob.numel()


def sharded_score(engine, total, seed, first_index=0, acq_mask=1, y_opt=0.0, xi=0.01, kappa=1.96, group=None):
    """Score candidates first_index .. first_index+total-1 (generated from `seed`) across all ranks."""
    rank, world = dist.get_rank(group), dist.get_world_size(group)
    start, count = shard_range(total, rank, world)
    rec = None
    if count > 0:
        rec = engine.score(None, M=count, seed=seed, first_index=first_index + start, acq_mask=acq_mask,
                           y_opt=y_opt, xi=xi, kappa=kappa)
    return allgather_best(rec, device=engine.tdev, group=group)


def max_over_ranks(value, device="cpu", group=None):
    t = torch.tensor([float(value)], dtype=torch.float64, device=device)
    dist.all_reduce(t, op=dist.ReduceOp.MAX, group=group)
    return float(t.item())


def to_numpy_seed(seed):
    return int(np.uint64(seed))
