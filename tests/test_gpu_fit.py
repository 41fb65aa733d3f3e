"""GPU parity of the GP fit (Gram + blocked Cholesky + alpha + LML) against sklearn / the numpy oracle."""
import numpy as np
import pytest
import torch

from metaopt_b200 import _lib
from oracle import gp as ogp

from helpers import gp_problem

pytestmark = pytest.mark.gpu

# FP64 tolerances: different (blocked, DMMA) summation order than LAPACK dpotrf/dpotrs.
L_RTOL, ALPHA_RTOL = 1e-9, 1e-7


@pytest.mark.parametrize("n,d,kind,normalize_y", [
    (1, 3, 0, False), (7, 2, 0, True), (64, 5, 1, True), (65, 5, 0, False), (255, 4, 0, True),
    (257, 20, 0, True), (700, 11, 1, False), (1030, 20, 0, True)])
def test_factor_and_alpha_match_oracle(engine, n, d, kind, normalize_y):
    X, y = gp_problem(n, d, seed=n)
    ls = np.linspace(0.7, 1.4, d)
    amp, alpha = 1.3, 1e-10 + 1e-4
    engine.fit(X, y, kind=kind, amplitude=amp, length_scale=ls, alpha=alpha, normalize_y=normalize_y)
    ref = ogp.SklearnGP(kind, amp, ls, alpha, normalize_y).fit(X, y)
    L, al = engine.get_factor()
    assert np.all(np.triu(L, 1) == 0.0)
    np.testing.assert_allclose(L, ref.L, rtol=L_RTOL, atol=1e-12)
    np.testing.assert_allclose(al, ref.alpha_, rtol=ALPHA_RTOL, atol=ALPHA_RTOL * np.abs(ref.alpha_).max())
    y_mean, y_std, y_min = engine.ystats()
    if normalize_y:
        np.testing.assert_allclose([y_mean, y_std], [y.mean(), y.std() if n > 1 else 1.0], rtol=1e-13)
    else:
        assert (y_mean, y_std) == (0.0, 1.0)
    assert y_min == y.min()
    np.testing.assert_allclose(engine.lml(), ref.gpr.log_marginal_likelihood_value_, rtol=1e-9)


def test_device_and_host_inputs_agree(engine):
    X, y = gp_problem(300, 6, seed=3)
    engine.fit(X, y, alpha=1e-6, normalize_y=True)
    L0, a0 = engine.get_factor()
    engine.fit(torch.as_tensor(X).cuda(), torch.as_tensor(y).cuda(), alpha=1e-6, normalize_y=True)
    L1, a1 = engine.get_factor()
    assert np.array_equal(L0, L1) and np.array_equal(a0, a1)


def test_isotropic_length_scale(engine):
    X, y = gp_problem(200, 4, seed=1)
    engine.fit(X, y, length_scale=0.5, alpha=1e-8)
    ref = ogp.SklearnGP(0, 1.0, 0.5 * np.ones(4), 1e-8, False).fit(X, y)
    L, al = engine.get_factor()
    np.testing.assert_allclose(L, ref.L, rtol=L_RTOL, atol=1e-12)


def test_not_positive_definite_is_reported(engine):
    X, y = gp_problem(50, 3, seed=2)
    X[10] = X[3]
    X[20] = X[3]
    with pytest.raises(_lib.NotPositiveDefinite, match="positive definite"):
        engine.fit(X, y, alpha=0.0)
    # context still usable
    engine.fit(X, y, alpha=1e-6)


@pytest.mark.parametrize("schedule", ["OB200_FIT_FUSED_STEP", "OB200_FIT_SPLIT_STEP"])
def test_not_positive_definite_in_a_late_block(engine, monkeypatch, schedule):
    """The failing pivot lies in the fifth 64-column block: under the split schedule the diagonal kernel of a later step
    (on its own stream) has to stop the steps after it; the error names a pivot at or after the duplicated row."""
    monkeypatch.setenv(schedule, "1")
    X, y = gp_problem(400, 4, seed=3)
    X[300] = X[17]
    with pytest.raises(_lib.NotPositiveDefinite, match="positive definite") as err:
        engine.fit(X, y, alpha=0.0)
    pivot = int(str(err.value).split("pivot ")[1].split(" ")[0])
    assert 300 < pivot <= 400
    engine.fit(X, y, alpha=1e-6)
    L, _ = engine.get_factor()
    assert np.isfinite(L).all()


@pytest.mark.parametrize("schedule", ["OB200_FIT_FUSED_STEP", "OB200_FIT_SPLIT_STEP", "OB200_FIT_NO_LOOKAHEAD"])
def test_step_schedules_agree(engine, monkeypatch, schedule):
    """The fit picks its Cholesky schedule by size (fused step below ~3000 observations, split step above): each of them,
    forced at one size, gives the factor of the reference."""
    X, y = gp_problem(700, 6, seed=4)
    monkeypatch.setenv(schedule, "1")
    if schedule == "OB200_FIT_NO_LOOKAHEAD":
        monkeypatch.setenv("OB200_FIT_SPLIT_STEP", "1")
    engine.fit(X, y, alpha=1e-6, normalize_y=True)
    L, al = engine.get_factor()
    ref = ogp.SklearnGP(0, 1.0, np.ones(6), 1e-6, True).fit(X, y)
    np.testing.assert_allclose(L, ref.L, rtol=L_RTOL, atol=1e-12)
    np.testing.assert_allclose(al, ref.alpha_, rtol=ALPHA_RTOL, atol=ALPHA_RTOL * np.abs(ref.alpha_).max())


def test_bad_arguments(engine):
    X, y = gp_problem(10, 3)
    with pytest.raises(_lib.Ob200Error):
        engine.fit(X, y, length_scale=[1.0, 2.0])  # neither 1 nor D entries
    with pytest.raises(_lib.Ob200Error):
        engine.fit(X, y, amplitude=-1.0)
    with pytest.raises(_lib.Ob200Error):
        engine.fit(np.zeros((9000, 2)), np.zeros(9000))  # n > 8192


@pytest.mark.parametrize("kind,n,d,normalize_y", [(0, 120, 3, True), (1, 90, 5, False), (0, 300, 20, True)])
def test_lml_gradient_matches_sklearn(engine, kind, n, d, normalize_y):
    from sklearn.gaussian_process import GaussianProcessRegressor
    from sklearn.gaussian_process.kernels import RBF, ConstantKernel, Matern, WhiteKernel
    X, y = gp_problem(n, d, seed=20 + n)
    ls = np.linspace(0.5, 1.7, d)
    amp, noise, alpha = 1.8, 3e-3, 1e-10
    base = Matern(length_scale=ls, nu=2.5) if kind == 0 else RBF(length_scale=ls)
    kernel = ConstantKernel(amp) * base + WhiteKernel(noise)
    gpr = GaussianProcessRegressor(kernel=kernel, alpha=alpha, optimizer=None, normalize_y=normalize_y).fit(X, y)
    lml_ref, grad_ref = gpr.log_marginal_likelihood(gpr.kernel_.theta, eval_gradient=True)
    engine.fit(X, y, kind=kind, amplitude=amp, length_scale=ls, alpha=alpha + noise, normalize_y=normalize_y)
    lml, g = engine.lml_grad()
    ours = np.concatenate([[g[0]], g[1:1 + d], [noise * g[1 + d]]])  # theta = log c, log ls, log noise
    np.testing.assert_allclose(lml, lml_ref, rtol=1e-9)
    np.testing.assert_allclose(ours, grad_ref, rtol=1e-6, atol=1e-6 * np.abs(grad_ref).max())


def test_hyperparameter_optimisation_reaches_sklearn_optimum(engine):
    import warnings
    from sklearn.gaussian_process import GaussianProcessRegressor
    from sklearn.gaussian_process.kernels import ConstantKernel, Matern, WhiteKernel
    from metaopt_b200 import hyperopt
    X, y = gp_problem(150, 4, seed=77, noise=0.05)
    kernel = ConstantKernel(1.0, (1e-2, 1e3)) * Matern(length_scale=np.ones(4), length_scale_bounds=(1e-2, 1e2), nu=2.5) + \
        WhiteKernel(1.0, (1e-5, 1e5))
    with warnings.catch_warnings():
        warnings.simplefilter("ignore")
        gpr = GaussianProcessRegressor(kernel=kernel, alpha=1e-10, normalize_y=True, n_restarts_optimizer=0).fit(X, y)
    res = hyperopt.optimise(engine, torch.as_tensor(X).cuda(), torch.as_tensor(y).cuda(), kind=0, amplitude=1.0,
                            length_scale=np.ones(4), noise=1.0, alpha=1e-10, normalize_y=True)
    assert res.lml >= gpr.log_marginal_likelihood_value_ - 1e-3 * abs(gpr.log_marginal_likelihood_value_) - 1e-3
    assert res.n_evals < 200 and 1e-5 <= res.noise <= 1e5
    # the engine is left fitted at the optimum
    np.testing.assert_allclose(engine.lml(), res.lml, rtol=1e-12)


@pytest.mark.parametrize("n0,steps,d,kind,normalize_y", [(1, 40, 2, 0, True), (150, 60, 5, 1, False), (500, 30, 20, 0, True)])
def test_appending_observations_equals_refitting(engine, n0, steps, d, kind, normalize_y):
    """`ob200_gp_append` (one row of L / inv(L) per new observation) against a from-scratch fit and against sklearn,
    across a padded-factor boundary (n_pad grows in steps of 192 -> the engine falls back to a refit there)."""
    X, y = gp_problem(n0 + steps, d, seed=7)
    ls = np.linspace(0.8, 1.3, d)
    kw = dict(kind=kind, amplitude=1.2, length_scale=ls, alpha=1e-6, normalize_y=normalize_y)
    engine.fit(X[:n0], y[:n0], **kw)
    before = engine.n_appends
    C = np.random.RandomState(1).rand(300, d)
    for n in range(n0 + 1, n0 + steps + 1):
        engine.fit(X[:n], y[:n], **kw)
    crossed = (n0 + steps - 1) // 192 - (n0 - 1) // 192      # refits forced by a full padded factor
    assert engine.n_appends - before == steps - crossed
    L1, a1 = engine.get_factor()
    out1 = engine.score(torch.as_tensor(C).cuda(), acq_mask=7, y_opt=float(y.min()), dense=True)
    lml1 = engine.lml()
    engine.fit(X, y, incremental=False, **kw)
    L2, a2 = engine.get_factor()
    out2 = engine.score(torch.as_tensor(C).cuda(), acq_mask=7, y_opt=float(y.min()), dense=True)
    # the appended rows come from inv(L) k (error ~ cond(L) eps) instead of a substitution: absolute 1e-10 on entries of O(1)
    np.testing.assert_allclose(L1, L2, rtol=1e-7, atol=1e-10)
    np.testing.assert_allclose(a1, a2, rtol=1e-6, atol=1e-6 * np.abs(a2).max())
    np.testing.assert_allclose(lml1, engine.lml(), rtol=1e-7)  # y^T alpha inherits alpha's conditioning (RBF case: cond ~1e9)
    np.testing.assert_allclose(out1["mu"].cpu().numpy(), out2["mu"].cpu().numpy(), rtol=1e-8, atol=1e-9)
    np.testing.assert_allclose(out1["sigma"].cpu().numpy(), out2["sigma"].cpu().numpy(), rtol=1e-6, atol=1e-8)
    ref = ogp.SklearnGP(kind, 1.2, ls, 1e-6, normalize_y).fit(X, y)
    mu_r, sd_r = ref.predict(C)
    np.testing.assert_allclose(out1["mu"].cpu().numpy(), mu_r, rtol=1e-6, atol=1e-6 * np.abs(mu_r).max())
    np.testing.assert_allclose(out1["sigma"].cpu().numpy(), sd_r, rtol=1e-6, atol=1e-6)
    assert out1["EI"]["index"] == out2["EI"]["index"]


def test_append_is_not_used_when_history_or_kernel_changes(engine):
    X, y = gp_problem(40, 3, seed=2)
    engine.fit(X[:30], y[:30], alpha=1e-6)
    k = engine.n_appends
    engine.fit(X[:31], y[:31], alpha=1e-6)
    assert engine.n_appends == k + 1
    engine.fit(X[:32], y[:32], alpha=1e-5)                 # different diagonal term -> refit
    assert engine.n_appends == k + 1
    X2 = X.copy()
    X2[3, 0] += 0.25
    engine.fit(X2[:33], y[:33], alpha=1e-5)                # an old row changed -> refit
    assert engine.n_appends == k + 1
    engine.fit(X2[:34], y[:34], alpha=1e-5)
    assert engine.n_appends == k + 2
    engine.fit(X2[:36], y[:36], alpha=1e-5)                # two new rows -> refit
    assert engine.n_appends == k + 2
    ref = ogp.SklearnGP(0, 1.0, np.ones(3), 1e-5, False).fit(X2[:36], y[:36])
    L, al = engine.get_factor()
    np.testing.assert_allclose(L, ref.L, rtol=1e-9, atol=1e-12)


def test_append_reports_a_duplicate_point(engine):
    """alpha = 0 and a repeated point: the new pivot is 0 up to rounding, so (deterministically per data set) it comes
    out <= 0 and is reported like numpy.linalg.LinAlgError, or as a tiny positive number; never as NaN."""
    raised = 0
    for seed in range(8):
        X, y = gp_problem(20, 3, seed=seed)
        engine.fit(X, y, alpha=0.0)
        k = engine.n_appends
        try:
            engine.fit(np.vstack([X, X[4]]), np.append(y, y[4]), alpha=0.0)
        except _lib.NotPositiveDefinite as exc:
            assert "positive definite" in str(exc)
            raised += 1
        else:
            assert engine.n_appends == k + 1
            L, _ = engine.get_factor()
            assert np.all(np.isfinite(L)) and 0.0 < L[20, 20] < 1e-6
    assert raised >= 1
    engine.fit(X, y, alpha=1e-8)
