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


def test_bad_arguments(engine):
    X, y = gp_problem(10, 3)
    with pytest.raises(_lib.Ob200Error):
        engine.fit(X, y, length_scale=[1.0, 2.0])  # neither 1 nor D entries
    with pytest.raises(_lib.Ob200Error):
        engine.fit(X, y, amplitude=-1.0)
    with pytest.raises(_lib.Ob200Error):
        engine.fit(np.zeros((9000, 2)), np.zeros(9000))  # n > 8192
