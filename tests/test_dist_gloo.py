"""N>1 host logic on CPU: world_size-2 `gloo` process group (127.0.0.1), exercising the shard split,
the winner exchange (`allgather_best`, the 'argmax all-reduce') and the max-over-ranks timing helper."""
import os
import socket

import numpy as np
import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

from metaopt_b200 import dist as obdist


def test_shard_range_partitions_exactly():
    for total in (0, 1, 7, 1_000_000, 16_000_001):
        for world in (1, 2, 3, 4, 8):
            parts = [obdist.shard_range(total, r, world) for r in range(world)]
            assert parts[0][0] == 0 and sum(c for _, c in parts) == total
            for (s0, c0), (s1, _) in zip(parts, parts[1:]):
                assert s0 + c0 == s1
            assert max(c for _, c in parts) - min(c for _, c in parts) <= 1
    with pytest.raises(ValueError):
        obdist.shard_range(10, 2, 2)


def test_merge_prefers_value_then_lowest_index():
    a = {"EI": dict(value=1.0, index=50, mu=0.1, sigma=0.2), "LCB": dict(value=-3.0, index=7, mu=0, sigma=1)}
    b = {"EI": dict(value=1.0, index=20, mu=0.3, sigma=0.4), "LCB": dict(value=-2.0, index=900, mu=0, sigma=1)}
    out = obdist.merge_best([a, None, b])
    assert out["EI"]["index"] == 20 and out["LCB"]["index"] == 900 and "PI" not in out


def test_pack_unpack_is_bit_exact():
    rec = {"EI": dict(value=0.1 + 0.2, index=(1 << 62) + 12345, mu=-1e-300, sigma=3.5),
           "PI": dict(value=1.0, index=0, mu=0.0, sigma=0.0)}
    assert obdist.unpack_best(obdist.pack_best(rec)) == rec
    assert obdist.unpack_best(obdist.pack_best(None)) == {}


def _free_port():
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    port = s.getsockname()[1]
    s.close()
    return port


def _worker(rank, world, port, out_dir):
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        total = 1001
        start, count = obdist.shard_range(total, rank, world)
        # a fake per-shard scorer: utility = a fixed pseudo-random function of the global index,
        # with a deliberate cross-shard tie between indices 3 and 700
        util = np.random.RandomState(0).rand(total)
        util[3] = util[700] = 2.0
        local = util[start:start + count]
        i = int(np.argmax(local))
        rec = {"EI": dict(value=float(local[i]), index=start + i, mu=float(rank), sigma=1.0)}
        if rank == 1:
            rec["LCB"] = dict(value=5.0, index=start, mu=0.0, sigma=0.0)
        merged = obdist.allgather_best(rec)
        slow = obdist.max_over_ranks(10.0 * (rank + 1))
        torch.save(dict(merged=merged, slow=slow, shard=(start, count)), os.path.join(out_dir, "r%d.pt" % rank))
    finally:
        dist.destroy_process_group()


def test_world_size_2_gloo_exchange(tmp_path):
    world, port = 2, _free_port()
    mp.spawn(_worker, args=(world, port, str(tmp_path)), nprocs=world, join=True)
    res = [torch.load(tmp_path / ("r%d.pt" % r)) for r in range(world)]
    assert res[0]["merged"] == res[1]["merged"]            # identical decision on every rank
    assert res[0]["merged"]["EI"]["index"] == 3             # tie across shards -> lowest global index
    assert res[0]["merged"]["EI"]["value"] == 2.0
    assert res[0]["merged"]["LCB"]["index"] == res[1]["shard"][0]  # an acquisition only one rank reported
    assert res[0]["slow"] == res[1]["slow"] == 20.0
    assert res[0]["shard"] == (0, 501) and res[1]["shard"] == (501, 500)
