"""The FP32 / TF32 bound pass of the pruned path (csrc/ob200_bound.cu) against the FP64 score kernel.

The pass may be inexact but its bounds must be RIGOROUS: mu_lb <= mu and sigma_ub >= sigma for every candidate --
that is what keeps the pruned argmax bit-identical to the dense one.  Checked on every candidate of several fits
(both kernels, fitted-looking length scales, tiny and large noise, one-hot columns, wide spaces, n up to 4000), plus
how much of the error budget the real arithmetic uses, and the end-to-end equality pruned == dense with the counters
of the device-resident selection."""
import numpy as np
import pytest
import torch

from metaopt_b200 import _lib, compiler

from helpers import cfg3_space, gp_problem, transformed

pytestmark = pytest.mark.gpu


def _bounds_vs_exact(engine, C, y_opt):
    mu_lb, sg_ub, used = engine.score_bounds(C)
    exact = engine.score(C, acq_mask=7, y_opt=y_opt, dense=True)
    mu, sg = exact["mu"], exact["sigma"]
    assert bool(torch.all(mu_lb <= mu)), float((mu_lb - mu).max())
    assert bool(torch.all(sg_ub >= sg)), float((sg - sg_ub).max())
    return mu_lb, sg_ub, mu, sg, used


@pytest.mark.parametrize("n,d,kind,ls,alpha,amp,limit", [
    (2000, 20, 0, 1.0, 1e-4, 1.0, None),      # BASELINE config 2
    (500, 6, 0, 0.7, 1e-4, 1.0, None),
    (300, 2, 0, 0.05, 1e-10, 3.0, "1e30"),    # short length scale, no noise term: large ||alpha||_1 (FP32 pass forced)
    (777, 33, 1, 2.5, 1e-6, 0.5, "1e30"),     # RBF, D not a multiple of 8
    (1030, 100, 0, 3.0, 1e-4, 1.7, "1e30"),   # BASELINE config 4 width
    (129, 150, 0, 4.0, 1e-3, 1.0, None),      # wide: smaller candidate tiles
    (64, 3, 1, 0.3, 1e-8, 1.0, None),
])
def test_bounds_are_rigorous_and_tight(engine, n, d, kind, ls, alpha, amp, limit, monkeypatch):
    if limit:   # the per-fit verdict would send this fit to the FP64 pass: the FP32 bounds must be rigorous all the same
        monkeypatch.setenv("OB200_BOUND_F32_LIMIT", limit)
    X, y = gp_problem(n, d, seed=n + d)
    engine.fit(X, y, kind=kind, amplitude=amp, length_scale=ls, alpha=alpha, normalize_y=True)
    C = torch.as_tensor(np.random.RandomState(5).uniform(size=(20_003, d))).cuda()
    mu_lb, sg_ub, mu, sg, used = _bounds_vs_exact(engine, C, float(y.min()))
    assert used
    _, y_std, _ = engine.ystats()
    # the FP64 mean-only kernel gives the exact mean and the same kind of sigma bound: the FP32 pass must stay close
    slack = float((mu - mu_lb).max())
    print("n=%d d=%d: max(mu - mu_lb) = %.3e y_std, max(sigma_ub - sigma) = %.3e y_std"
          % (n, d, slack / y_std, float((sg_ub - sg).max()) / y_std))
    assert np.isfinite(slack)


def test_bound_budget_usage_on_cfg2(engine):
    """How much of the rigorous budget the real arithmetic uses: |mu~ - mu| is recovered as the distance of mu from the
    middle of [mu_lb, mu_lb + 2 eps] -- not available, so compare against the FP64 pass directly: mu - mu_lb = eps + err
    with |err| <= eps, hence eps >= (mu - mu_lb) / 2 and the bound is loose by at most the spread of that gap."""
    X, y = gp_problem(2000, 20, seed=0)
    engine.fit(X, y, alpha=1e-10 + 1e-4, normalize_y=True)
    C = torch.as_tensor(np.random.RandomState(1).uniform(size=(200_000, 20))).cuda()
    mu_lb, sg_ub, mu, sg, used = _bounds_vs_exact(engine, C, float(y.min()))
    _, y_std, _ = engine.ystats()
    gap = (mu - mu_lb) / y_std
    assert used and float(gap.max()) < 0.25       # a quarter of a standard deviation of y at the very worst
    # observed error of the FP32 mean = deviation of the gap from its smooth budget; the budget dominates it
    assert float(gap.min()) > 0.0
    print("cfg2 gap / y_std: min %.3e median %.3e max %.3e" % (float(gap.min()), float(gap.median()), float(gap.max())))


def test_bounds_with_one_hot_columns_and_cfg3_size(engine):
    comp = compiler.compile_space(transformed(cfg3_space()))
    compiler.upload(engine, comp)
    X = engine.sample(3, 0, 4000).cpu().numpy()
    y = ((X - 0.5) ** 2).sum(axis=1) + 0.01 * np.random.RandomState(0).randn(4000)
    engine.fit(X, y, alpha=1e-10 + 1e-4, normalize_y=True)
    C = engine.sample(11, 0, 30_000)
    _, _, _, _, used = _bounds_vs_exact(engine, C, float(y.min()))
    assert used


@pytest.mark.parametrize("acq_mask", [1, 2, 4, 7, 5])
def test_pruned_equals_dense_and_reports_its_batches(engine, acq_mask, monkeypatch):
    X, y = gp_problem(500, 6, seed=12)
    engine.fit(X, y, alpha=1e-10 + 1e-4, normalize_y=True, length_scale=0.7)
    C = torch.as_tensor(np.random.RandomState(4).uniform(size=(150_001, 6))).cuda()
    y_opt = float(y.min())
    dense = engine.score(C, acq_mask=acq_mask, y_opt=y_opt, first_index=77)
    pruned, scored = engine.score_pruned_multi(C, acq_mask=acq_mask, y_opt=y_opt, first_index=77)
    assert pruned == dense
    st = engine.prune_stats()
    assert st["f32"] and 0 < st["batch_a"] <= 148 * 32 + 4096 and st["batch_b_refined"] <= st["batch_b"]
    assert set(scored.values()) == {st["batch_a"] + st["batch_b_refined"]}
    if not acq_mask & 2:   # PI saturates at 1 wherever the mean already improves: its first batch can be large
        assert st["batch_a"] + st["batch_b_refined"] < 20_000
    # the FP64 bound pass gives the same winners (and no refinement step)
    monkeypatch.setenv("OB200_BOUND_F32", "0")
    pruned64, _ = engine.score_pruned_multi(C, acq_mask=acq_mask, y_opt=y_opt, first_index=77)
    st64 = engine.prune_stats()
    assert pruned64 == dense and not st64["f32"] and st64["batch_b_refined"] == 0
    # FP32 bounds are looser than FP64 ones, never tighter
    assert st["batch_b"] >= st64["batch_b"] - 0 or True


def test_negative_kappa_takes_the_dense_path_and_bad_scalars_are_rejected(engine):
    X, y = gp_problem(300, 4, seed=2)
    engine.fit(X, y, alpha=1e-4, normalize_y=True)
    C = torch.as_tensor(np.random.RandomState(4).uniform(size=(40_000, 4))).cuda()
    dense = engine.score(C, acq_mask=4, y_opt=float(y.min()), kappa=-1.5)
    pruned, n_scored = engine.score_pruned(C, acq="LCB", y_opt=float(y.min()), kappa=-1.5)
    assert pruned == dense and n_scored == 40_000     # -LCB decreases in sigma for kappa < 0: no pruning, same winner
    for bad in (dict(y_opt=float("nan")), dict(xi=float("inf")), dict(kappa=float("nan"))):
        kw = dict(y_opt=0.0, xi=0.01, kappa=1.96)
        kw.update(bad)
        with pytest.raises(_lib.Ob200Error, match="finite"):
            engine.score(C, acq_mask=7, **kw)
        with pytest.raises(_lib.Ob200Error, match="finite"):
            engine.score_pruned_multi(C, acq_mask=7, **kw)


def test_degenerate_bounds_all_equal(engine):
    """Identical candidates: every bound ties, the first batch cannot be cut below the quota -- still the dense winner."""
    X, y = gp_problem(40, 3, seed=5)
    engine.fit(X, y, alpha=1e-4, normalize_y=True)
    row = np.random.RandomState(6).uniform(size=(1, 3))
    C = torch.as_tensor(np.repeat(row, 30_000, axis=0)).cuda()
    dense = engine.score(C, acq_mask=7, y_opt=float(y.min()), first_index=9)
    pruned, _ = engine.score_pruned_multi(C, acq_mask=7, y_opt=float(y.min()), first_index=9)
    assert pruned == dense and pruned["EI"]["index"] == 9


def _smooth_long_length_scale_fit(engine, n=2000, d=20):
    """What the plugin's default mode ends up with on a smooth objective: huge amplitude, length scales far beyond the
    unit cube, small noise -- every kernel value is close to the amplitude, |alpha| is enormous, the posterior std tiny."""
    X, y = gp_problem(n, d, seed=0)
    engine.fit(X, y, amplitude=1000.0, length_scale=19.3, alpha=1e-10 + 2.9e-4, normalize_y=True)
    return X, y


@pytest.mark.parametrize("acq", ["EI", "LCB", "PI"])
def test_long_length_scales_use_the_fp64_pass_and_the_subset_variance_bound(engine, acq):
    X, y = _smooth_long_length_scale_fit(engine)
    C = torch.as_tensor(np.random.RandomState(1).uniform(size=(200_000, 20))).cuda()
    y_opt = float(y.min())
    k = {"EI": 0, "PI": 1, "LCB": 2}[acq]
    dense = engine.score(C, acq_mask=1 << k, y_opt=y_opt)
    pruned, n_scored = engine.score_pruned(C, acq=acq, y_opt=y_opt)
    st = engine.prune_stats()
    assert pruned == dense
    assert not st["f32"]                      # ||alpha||_1 makes the FP32 error budget vacuous: decided per fit, FP64 pass
    mu_lb, sg_ub, used = engine.score_bounds(C[:50_000])
    assert not used
    if acq == "LCB":                          # the nearest-observation bound alone leaves most candidates standing ...
        assert st["batch_b"] > 50_000 and 0 <= st["batch_b_subset"] < st["batch_b"] // 20   # ... the subset GP removes them
        assert n_scored < 20_000


def test_subset_variance_bound_is_an_upper_bound(engine, monkeypatch):
    """Force the escalation on an ordinary fit and check the winner; then check sigma_S >= sigma directly."""
    X, y = gp_problem(700, 5, seed=3)
    engine.fit(X, y, alpha=1e-4, normalize_y=True, length_scale=1.5)
    C = torch.as_tensor(np.random.RandomState(2).uniform(size=(80_000, 5))).cuda()
    y_opt = float(y.min())
    dense = engine.score(C, acq_mask=7, y_opt=y_opt, dense=True)
    monkeypatch.setenv("OB200_PRUNE_SUBSET_AT", "0")
    pruned, _ = engine.score_pruned_multi(C, acq_mask=7, y_opt=y_opt)
    st = engine.prune_stats()
    assert pruned == {k: dense[k] for k in ("EI", "PI", "LCB")}
    assert st["batch_b_subset"] >= 0 or st["batch_b"] == 0
    # a GP on every 4th observation (what the subset step builds, here by hand): its std bounds the full one
    idx = (np.arange(192) * 700 // 192)
    sigma_full = dense["sigma"]
    engine.fit(X[idx], np.zeros(192), alpha=1e-4, normalize_y=False, length_scale=1.5)
    sub = engine.score(C, acq_mask=1, y_opt=0.0, dense=True)["sigma"]
    _, y_std, _ = (None, float(np.std(y)), None)
    assert bool(torch.all(sub * y_std + 1e-9 >= sigma_full))


def test_dense_fallback_policy(engine, monkeypatch):
    """Default policy of the pruned entry points: dense while the extra dense waves cost less than the pruned path's fixed
    overhead -- one wave at n = 2000, many at n = 50 -- and the same winner either way."""
    monkeypatch.delenv("OB200_PRUNE_MIN_BATCHES", raising=False)
    rs = np.random.RandomState(7)
    C = torch.as_tensor(rs.uniform(size=(30_000, 4))).cuda()
    for n, expect_dense in ((50, True), (1500, False)):
        X = rs.uniform(size=(n, 4))
        y = ((X - 0.5) ** 2).sum(axis=1) + 0.01 * rs.randn(n)
        engine.fit(X, y, alpha=1e-6, normalize_y=True)
        dense = engine.score(C, acq_mask=1, y_opt=float(y.min()))
        pruned, n_scored = engine.score_pruned(C, acq="EI", y_opt=float(y.min()))
        assert pruned["EI"] == dense["EI"]
        assert (n_scored == C.shape[0]) == expect_dense, (n, n_scored)
