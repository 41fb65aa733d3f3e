"""GPU box: write-only vs copy bandwidth of this B200 (what bounds the sampler, which only writes)."""
import json

import torch


def ev(fn, reps=5):
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


out = {}
for gib in (1, 4, 12):
    n = gib * (1 << 30) // 8
    x = torch.empty(n, dtype=torch.float64, device="cuda")
    y = torch.empty(n, dtype=torch.float64, device="cuda")
    x.fill_(1.0); y.fill_(2.0)
    rec = {}
    rec["memset_zero_gbs"] = n * 8 / ev(lambda: x.zero_()) / 1e6
    rec["fill_gbs"] = n * 8 / ev(lambda: x.fill_(1.5)) / 1e6
    rec["copy_gbs_read_plus_write"] = 2 * n * 8 / ev(lambda: y.copy_(x)) / 1e6
    rec["read_only_sum_gbs"] = n * 8 / ev(lambda: x.sum()) / 1e6
    xb = x.view(torch.uint8)
    rec["memset_u8_gbs"] = n * 8 / ev(lambda: xb.fill_(3)) / 1e6
    out["%dGiB" % gib] = rec
    del x, y, xb
print(json.dumps(out, indent=1))
