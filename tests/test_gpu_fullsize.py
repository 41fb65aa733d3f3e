"""BASELINE.json's full configuration sizes, checked through size-independent properties (the CPU oracle
cannot finish these in seconds): sharding invariance of the argmax, pruned == dense, bounds that must
hold for every candidate, linearity of the posterior mean in y, and distribution checksums of 1e9 samples."""
import numpy as np
import pytest
import torch

from metaopt_b200 import compiler
from metaopt_b200 import dist as obdist

from helpers import cfg3_space, cfg5_space, gp_problem, real_space, transformed

pytestmark = pytest.mark.gpu


def _fit_cfg(engine, space, n_obs, seed=3):
    comp = compiler.compile_space(transformed(space))
    compiler.upload(engine, comp)
    X = engine.sample(seed, 0, n_obs).cpu().numpy()
    y = ((X - 0.5) ** 2).sum(axis=1) + 0.01 * np.random.RandomState(0).randn(n_obs)
    engine.fit(X, y, alpha=1e-10 + 1e-4, normalize_y=True)
    return comp, X, y


def test_cfg2_one_million_candidates_properties(engine):
    comp, X, y = _fit_cfg(engine, real_space(20), 2000)
    M, y_opt = 1_000_000, float(y.min())
    out = engine.score(None, M=M, seed=1, first_index=0, acq_mask=7, y_opt=y_opt, dense=True)
    mu, sd, acq = out["mu"], out["sigma"], out["acq"]
    _, y_std, _ = engine.ystats()
    assert bool(torch.all(sd >= 0)) and bool(torch.all(sd <= y_std * (1 + 1e-12)))     # 0 <= sigma <= prior std
    assert bool(torch.all(acq[0] >= 0)) and bool(torch.all((acq[1] >= 0) & (acq[1] <= 1)))  # EI >= 0, PI in [0, 1]
    torch.testing.assert_close(acq[2], -(mu - 1.96 * sd), rtol=1e-15, atol=1e-15)     # -LCB = -(mu - kappa sigma), fused on the device
    for k, name in enumerate(("EI", "PI", "LCB")):                                    # winner = first argmax of the dense values
        assert out[name]["index"] == int(torch.argmax(acq[k]))
        assert out[name]["value"] == float(acq[k][out[name]["index"]])
    # sharding invariance (what 8 GPUs do) and exact pruning
    parts = []
    for r in range(8):
        s0, c = obdist.shard_range(M, r, 8)
        parts.append(engine.score(None, M=c, seed=1, first_index=s0, acq_mask=7, y_opt=y_opt))
    merged = obdist.merge_best(parts)
    for name in ("EI", "PI", "LCB"):
        assert merged[name] == {k: out[name][k] for k in ("value", "index", "mu", "sigma")}
    for name in ("EI", "LCB"):
        pruned, n_scored = engine.score_pruned(None, M=M, seed=1, first_index=0, acq=name, y_opt=y_opt)
        assert pruned[name] == merged[name] and n_scored < M
    # permutation invariance: the winner's values do not depend on where the candidate sits
    cand = engine.sample(1, 0, M)
    perm = torch.randperm(M, device=cand.device, generator=torch.Generator(device=cand.device).manual_seed(0))
    shuffled = engine.score(cand[perm].contiguous(), acq_mask=1, y_opt=y_opt)
    assert int(perm[shuffled["EI"]["index"]]) == out["EI"]["index"]
    assert shuffled["EI"]["value"] == out["EI"]["value"] and shuffled["EI"]["sigma"] == out["EI"]["sigma"]


def test_posterior_mean_is_linear_in_y_and_sigma_independent_of_y(engine):
    X, y = gp_problem(2000, 20, seed=0)
    C = torch.rand((200_000, 20), dtype=torch.float64, device="cuda", generator=torch.Generator(device="cuda").manual_seed(1))
    engine.fit(X, y, alpha=1e-4, normalize_y=False)
    a = engine.score(C, acq_mask=4, y_opt=0.0, dense=True)
    engine.fit(X, 3.0 * y, alpha=1e-4, normalize_y=False)
    b = engine.score(C, acq_mask=4, y_opt=0.0, dense=True)
    torch.testing.assert_close(b["mu"], 3.0 * a["mu"], rtol=1e-10, atol=1e-12)
    assert torch.equal(a["sigma"], b["sigma"])


def test_cfg3_four_million_candidates_sharded_equals_whole(engine):
    comp, X, y = _fit_cfg(engine, cfg3_space(), 4000)
    assert comp.d_out == 60
    M, y_opt = 4_000_000, float(y.min())
    whole, n_scored = engine.score_pruned(None, M=M, seed=11, first_index=0, acq="EI", y_opt=y_opt)
    parts = []
    for r in range(4):
        s0, c = obdist.shard_range(M, r, 4)
        parts.append(engine.score_pruned(None, M=c, seed=11, first_index=s0, acq="EI", y_opt=y_opt)[0])
    assert obdist.merge_best(parts)["EI"] == whole["EI"]
    # the winner really is a point of the mixed space: regenerate it from its index and re-score it alone
    row = engine.sample(11, whole["EI"]["index"], 1)
    alone = engine.score(row, acq_mask=1, y_opt=y_opt, first_index=whole["EI"]["index"])
    assert alone["EI"] == whole["EI"]
    raw = engine.sample(11, whole["EI"]["index"], 1, normalise=False).cpu().numpy()[0]
    assert compiler.regroup_row(comp, raw, transformed=True) in transformed(cfg3_space())


def test_cfg4_shape_one_shard_of_two_million(engine):
    comp, X, y = _fit_cfg(engine, real_space(100), 8000)
    y_opt = float(y.min())
    M = 2_000_000  # one GPU's shard of the 16M candidates
    pruned, n_scored = engine.score_pruned(None, M=M, seed=21, first_index=6_000_000, acq="EI", y_opt=y_opt)
    sub = engine.score(None, M=300_000, seed=21, first_index=6_000_000, acq_mask=1, y_opt=y_opt)
    assert pruned["EI"]["value"] >= sub["EI"]["value"]
    if pruned["EI"]["index"] < 6_300_000:
        assert pruned["EI"] == sub["EI"]
    row = engine.sample(21, pruned["EI"]["index"], 1)
    assert engine.score(row, acq_mask=1, y_opt=y_opt, first_index=pruned["EI"]["index"])["EI"] == pruned["EI"]


def test_cfg5_one_billion_points_checksums(engine):
    """1e9 points of the 50-D mixed space streamed through one 2^24-row buffer: per-column sums (a
    checksum of checksums) against the analytic means, exact structural invariants on every chunk."""
    space = cfg5_space()
    comp = compiler.compile_space(transformed(space))
    compiler.upload(engine, comp)
    chunk, total = 1 << 24, 1_000_000_000
    buf = torch.empty((chunk, comp.d_out), dtype=torch.float64, device="cuda")
    sums = torch.zeros(comp.d_out, dtype=torch.float64, device="cuda")
    lo = torch.full((comp.d_out,), float("inf"), dtype=torch.float64, device="cuda")
    hi = torch.full((comp.d_out,), -float("inf"), dtype=torch.float64, device="cuda")
    onehot_cols = [(s["out_col"], s["width"]) for s in comp.slots if s["ncat"] > 2]
    int_cols = [s["out_col"] for s in comp.slots if s["kind"] == 1]
    done = 0
    while done < total:
        n = min(chunk, total - done)
        view = buf[:n]
        engine.sample(7, done, n, normalise=False, out=view, check=False)
        sums += view.sum(dim=0)
        lo = torch.minimum(lo, view.min(dim=0).values)
        hi = torch.maximum(hi, view.max(dim=0).values)
        if done % (chunk * 16) == 0:  # structural checks on a subset of chunks (each is a full pass over 13 GB)
            for c0, w in onehot_cols[:4]:
                assert bool(torch.all(view[:, c0:c0 + w].sum(dim=1) == 1.0))
            iv = view[:, int_cols]
            assert bool(torch.all(iv == torch.floor(iv)))
        done += n
    engine._check(engine.lib.ob200_sample_status(engine.ctx))
    mean = (sums / total).cpu().numpy()
    lo, hi = lo.cpu().numpy(), hi.cpu().numpy()
    for item, dim in zip(comp.layout, transformed(space).values()):
        c0 = item["col"]
        orig = dim.original_dimension
        if item["kind"] == "categorical":
            np.testing.assert_allclose(mean[c0:c0 + 4], orig._probs, atol=2e-4)
            continue
        low, high = orig.interval()
        assert lo[c0] >= low and hi[c0] < high
        if orig._prior_name == "uniform":
            a, width = orig._args
            want = a + width / 2 if item["kind"] == "real" else a + (width - 1) / 2
            np.testing.assert_allclose(mean[c0], want, atol=5 * width / np.sqrt(12 * total) + 1e-9)
        else:  # reciprocal(a, b): E = (b - a) / ln(b / a)
            a, b = orig._args
            np.testing.assert_allclose(mean[c0], (b - a) / np.log(b / a), rtol=2e-4)
