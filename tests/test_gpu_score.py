"""GPU parity of the fused score kernel (cross-kernel + posterior + acquisition + argmax) against
scikit-learn's GaussianProcessRegressor.predict + restated skopt acquisitions."""
import numpy as np
import pytest
import torch

from metaopt_b200 import compiler
from metaopt_b200.engine import Engine
from oracle import acquisition as oacq
from oracle import gp as ogp

from helpers import cfg3_space, gp_problem, transformed

pytestmark = pytest.mark.gpu

# north_star: mean/std/acquisition within 1e-6 relative in FP64; argmax index bit-exact
RTOL = 1e-6


def _compare(engine, X, y, C, kind=0, amp=1.0, ls=1.0, alpha=1e-10 + 1e-4, normalize_y=True, xi=0.01, kappa=1.96,
             host=False, acq_rtol=None):
    acq_rtol = RTOL if acq_rtol is None else acq_rtol  # (tools/fuzz_parity.py loosens this for EI's far tail)
    d = X.shape[1]
    ls_vec = np.broadcast_to(np.asarray(ls, dtype=float), (d,)).copy()
    engine.fit(X, y, kind=kind, amplitude=amp, length_scale=ls, alpha=alpha, normalize_y=normalize_y)
    ref = ogp.SklearnGP(kind, amp, ls_vec, alpha, normalize_y).fit(X, y)
    mu_r, sd_r = ref.predict(C)
    y_opt = float(y.min())
    if host:
        out = engine.score_host(C, acq_mask=7, y_opt=y_opt, xi=xi, kappa=kappa, dense=True)
        mu, sd, acq = out["mu"], out["sigma"], out["acq"]
    else:
        out = engine.score(torch.as_tensor(C).cuda(), acq_mask=7, y_opt=y_opt, xi=xi, kappa=kappa, dense=True)
        mu, sd, acq = out["mu"].cpu().numpy(), out["sigma"].cpu().numpy(), out["acq"].cpu().numpy()
    np.testing.assert_allclose(mu, mu_r, rtol=RTOL, atol=RTOL * np.abs(mu_r).max())
    # std: relative where sigma is not cancellation-dominated, absolute otherwise
    floor = float(1e-6 * np.sqrt(amp) * np.ravel(ref.gpr._y_train_std)[0])
    np.testing.assert_allclose(sd, sd_r, rtol=RTOL, atol=floor)
    for k, name in ((0, "EI"), (1, "PI"), (2, "LCB")):
        util = oacq.utility(k, mu_r, sd_r, y_opt, xi, kappa)
        np.testing.assert_allclose(acq[k], util, rtol=acq_rtol, atol=acq_rtol * np.abs(util).max())
        pick = oacq.pick(k, mu_r, sd_r, y_opt, xi, kappa)
        got = out[name]
        srt = np.sort(util)
        gap = srt[-1] - srt[-2] if len(srt) > 1 else 1.0
        if gap > 1e-9 * max(abs(srt[-1]), 1e-300):
            assert got["index"] == pick, (name, got, pick, gap)
        else:  # a genuine near-tie: the values must agree
            np.testing.assert_allclose(util[got["index"]], util[pick], rtol=RTOL)
        np.testing.assert_allclose(got["value"], util[got["index"]], rtol=acq_rtol, atol=acq_rtol * np.abs(util).max())
        np.testing.assert_allclose(got["mu"], mu_r[got["index"]], rtol=RTOL, atol=RTOL * np.abs(mu_r).max())
        # the reported winner is exactly the argmax of the dense device values (first occurrence)
        assert got["index"] == int(np.argmax(acq[k]))
    return out


@pytest.mark.parametrize("n,d,M,kind,normalize_y", [
    (5, 2, 1, 0, True), (5, 2, 63, 0, False), (50, 2, 1000, 0, True), (64, 3, 64, 1, True),
    (200, 20, 777, 0, True), (257, 7, 2049, 1, False), (100, 150, 300, 0, True), (193, 9, 129, 0, True), (700, 60, 1500, 0, True), (1030, 100, 513, 0, True)])
def test_posterior_and_acquisition_match_sklearn(engine, n, d, M, kind, normalize_y):
    X, y = gp_problem(n, d, seed=n + d)
    C = np.random.RandomState(1).uniform(size=(M, d))
    _compare(engine, X, y, C, kind=kind, normalize_y=normalize_y, ls=np.linspace(0.8, 1.6, d), amp=1.7)


@pytest.mark.parametrize("n,d,M", [(1, 1, 1), (1, 3, 5), (2, 1, 1), (3, 174, 2)])
def test_degenerate_sizes(engine, n, d, M):
    """One observation, one dimension, one candidate, the widest supported space: same contract as any other size."""
    rs = np.random.RandomState(n * 1000 + d)
    X = rs.uniform(size=(n, d))
    y = rs.uniform(size=n) + 1.0
    C = rs.uniform(size=(M, d))
    # (normalize_y with one observation divides by a zero standard deviation in the reference: sklearn's guard sets it to 1)
    _compare(engine, X, y, C, ls=0.7 if d < 100 else 4.0, normalize_y=True)
    _compare(engine, X, y, C, kind=1, ls=0.7 if d < 100 else 4.0, normalize_y=False)


def test_zero_candidates_and_empty_sample(engine):
    from metaopt_b200 import _lib
    from helpers import real_space
    X, y = gp_problem(20, 3, seed=1)
    engine.fit(X, y, alpha=1e-6)
    with pytest.raises(_lib.Ob200Error) as err:
        engine.score(torch.empty((0, 3), dtype=torch.float64, device="cuda"), acq_mask=1, y_opt=0.0)
    assert err.value.code == _lib.EINVAL
    comp = compiler.compile_space(transformed(real_space(3)))
    compiler.upload(engine, comp)
    assert tuple(engine.sample(1, 0, 0).shape) == (0, 3)


def test_cfg2_shape_n2000_d20(engine):
    """BASELINE config 2 recipe at a CPU-checkable candidate count."""
    X, y = gp_problem(2000, 20, seed=0)
    C = np.random.RandomState(1).uniform(size=(20000, 20))
    _compare(engine, X, y, C)


def test_cfg3_shape_n4000_d60(engine):
    """BASELINE config 3 at its real size (n = 4000, D = 60 after one-hot) at a CPU-checkable candidate count:
    observations AND candidates are points of the mixed space (10 Integer, 10 Categorical k=4, 10 Real),
    normalised with the bounds of `interval()` (one-hot columns live in (-0.1, 1.1))."""
    comp = compiler.compile_space(transformed(cfg3_space()))
    assert comp.d_out == 60
    compiler.upload(engine, comp)
    X = engine.sample(3, 0, 4000).cpu().numpy()
    y = ((X - 0.5) ** 2).sum(axis=1) + 0.01 * np.random.RandomState(0).randn(4000)
    C = engine.sample(11, 0, 5000).cpu().numpy()
    _compare(engine, X, y, C)


def test_largest_supported_fit_n8000_d100(engine):
    """BASELINE config 4 shape at a CPU-checkable candidate count: the n = 8000 Cholesky / inverse / alpha path."""
    X, y = gp_problem(8000, 100, seed=5)
    C = np.random.RandomState(2).uniform(size=(257, 100))
    _compare(engine, X, y, C)


def test_ill_conditioned_gram_matrix(engine):
    """No noise term, alpha = 1e-10 and clustered observations: cond(K) ~ 1e9; the posterior still has to
    agree with scikit-learn where scikit-learn itself is stable (mean relative to its scale, std above 1e-3)."""
    rs = np.random.RandomState(3)
    centres = rs.uniform(size=(20, 3))
    X = np.concatenate([c + 1e-3 * rs.randn(10, 3) for c in centres])
    y = np.sin(3 * X).sum(axis=1)
    C = rs.uniform(size=(500, 3))
    engine.fit(X, y, alpha=1e-10, normalize_y=True, length_scale=0.5)
    ref = ogp.SklearnGP(0, 1.0, 0.5 * np.ones(3), 1e-10, True).fit(X, y)
    mu_r, sd_r = ref.predict(C)
    out = engine.score(torch.as_tensor(C).cuda(), acq_mask=1, y_opt=float(y.min()), dense=True)
    mu, sd = out["mu"].cpu().numpy(), out["sigma"].cpu().numpy()
    np.testing.assert_allclose(mu, mu_r, rtol=1e-4, atol=1e-4 * np.abs(mu_r).max())
    # sigma^2 = c - ||L^-1 k*||^2 cancels ~9 digits here: nobody gets more than cond-limited
    # accuracy.  Against a long-double reference (tools/illcond_probe.py) scikit-learn is off by 7e-9 in the
    # variance and the device path by 8e-9 (substitution-based Cholesky panels + one Newton-Schulz step on
    # inv(L); it was 2e-6 with an inverted-diagonal-block panel), so the two agree to a few 1e-8.
    y_std = float(np.ravel(ref.gpr._y_train_std)[0])
    cond = np.linalg.cond(ogp.cross_kernel(0, X, X, 1.0, 0.5 * np.ones(3)) + 1e-10 * np.eye(len(X)))
    assert cond > 1e10
    np.testing.assert_allclose(sd ** 2, sd_r ** 2, rtol=0, atol=2e-7 * y_std ** 2)


def test_host_entry_point(engine):
    X, y = gp_problem(300, 8, seed=4)
    C = np.random.RandomState(2).uniform(size=(3001, 8))
    _compare(engine, X, y, C, host=True)


def test_candidates_on_observed_points_have_small_sigma(engine):
    X, y = gp_problem(128, 4, seed=9)
    C = np.concatenate([X[:32], np.random.RandomState(3).uniform(size=(32, 4))])
    out = _compare(engine, X, y, C, alpha=1e-10)
    assert np.all(out["sigma"].cpu().numpy()[:32] < 1e-3)


def test_ties_pick_lowest_index(engine):
    X, y = gp_problem(40, 3, seed=5)
    row = np.random.RandomState(6).uniform(size=(1, 3))
    C = np.repeat(row, 200, axis=0)  # all candidates identical -> every acquisition ties
    engine.fit(X, y, alpha=1e-6, normalize_y=True)
    out = engine.score(torch.as_tensor(C).cuda(), acq_mask=7, y_opt=float(y.min()), first_index=1000)
    assert out["EI"]["index"] == out["PI"]["index"] == out["LCB"]["index"] == 1000


def test_strided_candidates_and_mask(engine):
    X, y = gp_problem(100, 5, seed=7)
    engine.fit(X, y, alpha=1e-6)
    C = torch.rand((500, 9), dtype=torch.float64, device="cuda")
    a = engine.score(C[:, :5].contiguous(), acq_mask=4, y_opt=0.0)
    b = engine.score(C[:, :5], acq_mask=4, y_opt=0.0)  # ld = 9
    assert set(a) == {"LCB"} and a == b


def test_generated_candidates_equal_explicit_ones(engine):
    comp = compiler.compile_space(transformed(cfg3_space()))
    compiler.upload(engine, comp)
    X = engine.sample(99, 0, 150).cpu().numpy()
    y = ((X - 0.5) ** 2).sum(axis=1)
    engine.fit(X, y, alpha=1e-6, normalize_y=True)
    C = engine.sample(5, 10_000, 5000)
    a = engine.score(C, acq_mask=7, y_opt=float(y.min()), first_index=10_000)
    b = engine.score(None, M=5000, seed=5, first_index=10_000, acq_mask=7, y_opt=float(y.min()))
    assert a == b
    # sharded: two halves and a manual merge give the same winner as the whole range
    h1 = engine.score(None, M=2500, seed=5, first_index=10_000, acq_mask=1, y_opt=float(y.min()))["EI"]
    h2 = engine.score(None, M=2500, seed=5, first_index=12_500, acq_mask=1, y_opt=float(y.min()))["EI"]
    best = max((h1, h2), key=lambda r: (r["value"], -r["index"]))
    assert best == a["EI"]


@pytest.mark.parametrize("acq", ["EI", "PI", "LCB"])
def test_pruned_scoring_is_bit_identical_to_dense(engine, acq):
    X, y = gp_problem(500, 6, seed=12)
    engine.fit(X, y, alpha=1e-10 + 1e-4, normalize_y=True, length_scale=0.7)
    C = torch.as_tensor(np.random.RandomState(4).uniform(size=(150_001, 6))).cuda()
    k = {"EI": 0, "PI": 1, "LCB": 2}[acq]
    dense = engine.score(C, acq_mask=1 << k, y_opt=float(y.min()), first_index=77)
    pruned, n_scored = engine.score_pruned(C, acq=acq, y_opt=float(y.min()), first_index=77)
    assert pruned == dense
    assert 0 < n_scored <= C.shape[0]
    if acq != "PI":
        assert n_scored < C.shape[0] // 2  # the bound really prunes on this problem
    # small call falls back to the dense path
    small, n_small = engine.score_pruned(C[:1000], acq=acq, y_opt=float(y.min()), first_index=77)
    assert small == engine.score(C[:1000], acq_mask=1 << k, y_opt=float(y.min()), first_index=77) and n_small == 1000


def test_pruned_scoring_with_generated_candidates(engine):
    comp = compiler.compile_space(transformed(cfg3_space()))
    compiler.upload(engine, comp)
    X = engine.sample(99, 0, 300).cpu().numpy()
    y = ((X - 0.5) ** 2).sum(axis=1)
    engine.fit(X, y, alpha=1e-6, normalize_y=True)
    dense = engine.score(None, M=120_000, seed=5, first_index=10_000, acq_mask=1, y_opt=float(y.min()))
    pruned, n_scored = engine.score_pruned(None, M=120_000, seed=5, first_index=10_000, acq="EI", y_opt=float(y.min()))
    assert pruned == dense and n_scored <= 120_000


def test_chunked_generation_equals_single_pass(engine, monkeypatch):
    """Candidate sets larger than 2 GiB are generated and scored chunk by chunk with a running argmax;
    OB200_CHUNK_ROWS shrinks the chunk so the path is exercised at test size (ragged last chunk included)."""
    comp = compiler.compile_space(transformed(cfg3_space()))
    compiler.upload(engine, comp)
    X = engine.sample(99, 0, 200).cpu().numpy()
    y = ((X - 0.5) ** 2).sum(axis=1)
    engine.fit(X, y, alpha=1e-6, normalize_y=True)
    whole = engine.score(None, M=50_003, seed=8, first_index=123, acq_mask=7, y_opt=float(y.min()), dense=True)
    monkeypatch.setenv("OB200_CHUNK_ROWS", "12800")
    parts = engine.score(None, M=50_003, seed=8, first_index=123, acq_mask=7, y_opt=float(y.min()))
    for k in ("EI", "PI", "LCB"):
        assert parts[k] == whole[k]
    mu_only = engine.score(None, M=50_003, seed=8, first_index=123, acq_mask=1, y_opt=float(y.min()), dense=False)
    assert mu_only["EI"] == whole["EI"]


def test_export_import_roundtrip(engine):
    X, y = gp_problem(260, 6, seed=8)
    engine.fit(X, y, alpha=1e-6, normalize_y=True)
    C = torch.rand((1000, 6), dtype=torch.float64, device="cuda")
    a = engine.score(C, acq_mask=7, y_opt=float(y.min()), dense=True)
    blob = engine.export_state()
    other = Engine(0)
    try:
        other.import_state(blob.clone(), 260, 6)
        b = other.score(C, acq_mask=7, y_opt=float(y.min()), dense=True)
        assert other.ystats() == engine.ystats()
    finally:
        other.close()
    for k in ("EI", "PI", "LCB"):
        assert a[k] == b[k]
    assert torch.equal(a["mu"], b["mu"]) and torch.equal(a["sigma"], b["sigma"])


def test_acq_argmax_standalone(engine):
    rs = np.random.RandomState(0)
    mu, sd = rs.randn(100003), np.abs(rs.randn(100003))
    sd[::7] = 0.0
    out = engine.acq_argmax(torch.as_tensor(mu).cuda(), torch.as_tensor(sd).cuda(), first_index=5, acq_mask=7,
                            y_opt=-0.3, dense=True)
    for k, name in ((0, "EI"), (1, "PI"), (2, "LCB")):
        util = oacq.utility(k, mu, sd, -0.3)
        np.testing.assert_allclose(out["acq"][k].cpu().numpy(), util, rtol=1e-9, atol=1e-300)
        assert out[name]["index"] == 5 + int(np.argmax(out["acq"][k].cpu().numpy()))
        assert out[name]["index"] == 5 + oacq.pick(k, mu, sd, -0.3)


def test_score_before_fit_is_an_error():
    from metaopt_b200 import _lib
    e = Engine(0)
    try:
        with pytest.raises(_lib.Ob200Error, match="before"):
            e.score(torch.zeros((4, 2), dtype=torch.float64, device="cuda"))
    finally:
        e.close()


def test_host_entry_point_overlapped_copy_equals_device_path(engine):
    """Large host calls copy the candidate tail on a second stream while the head is scored."""
    X, y = gp_problem(64, 6, seed=14)
    engine.fit(X, y, alpha=1e-6, normalize_y=True)
    M = 148 * 32 * 16 * 4 + 77  # just above the overlap threshold, ragged
    C = np.random.RandomState(9).uniform(size=(M, 6))
    pinned = torch.as_tensor(C).pin_memory()
    a = engine.score_host(pinned.numpy(), acq_mask=7, y_opt=float(y.min()))
    b = engine.score(torch.as_tensor(C).cuda(), acq_mask=7, y_opt=float(y.min()))
    assert a == b
    assert engine.score_host(C, acq_mask=7, y_opt=float(y.min())) == b  # pageable host memory too


def test_pruned_portfolio_shares_the_mean_pass_and_matches_the_dense_winners(engine):
    X, y = gp_problem(300, 6, seed=9)
    engine.fit(X, y, alpha=1e-4, normalize_y=True, length_scale=0.8)
    C = torch.as_tensor(np.random.RandomState(4).rand(40000, 6)).cuda()
    y_opt = float(y.min())
    dense = engine.score(C, acq_mask=7, y_opt=y_opt)
    before = engine.launches
    multi, scored = engine.score_pruned_multi(C, acq_mask=7, y_opt=y_opt)
    launches_multi = engine.launches - before
    assert multi == {k: dense[k] for k in ("EI", "PI", "LCB")}
    assert set(scored) == {"EI", "PI", "LCB"} and all(0 < v <= 40000 for v in scored.values())
    before = engine.launches
    for name in ("EI", "PI", "LCB"):
        one, n_one = engine.score_pruned(C, acq=name, y_opt=y_opt)
        assert one[name] == dense[name] and 0 < n_one <= 40000
    assert launches_multi < engine.launches - before          # two mean-only passes and two first batches saved
    two, scored2 = engine.score_pruned_multi(C, acq_mask=5, y_opt=y_opt)
    assert set(two) == {"EI", "LCB"} and two["EI"] == dense["EI"] and two["LCB"] == dense["LCB"]
