"""GPU parity of the posterior gradients (`ob200_gp_predict_grad_host`) against the numpy oracle and against
central differences of the on-box scikit-learn estimator, and the plugin's `acq_optimizer="lbfgs"` mode."""
import numpy as np
import pytest

from metaopt_b200 import _lib, acq_opt
from metaopt_b200.algo import BayesianOptimizer
from metaopt_b200.space import Categorical, Integer, Real, Space, build_required_space
from oracle import acquisition as oacq
from oracle import gp as ogp

from helpers import gp_problem

pytestmark = pytest.mark.gpu

# FP64; the device sums the n terms in a different (lane-strided, tree) order than numpy's BLAS
MU_RTOL, GRAD_RTOL = 1e-11, 1e-8


@pytest.mark.parametrize("n,d,kind,normalize_y,m", [
    (1, 3, 0, False, 1), (37, 2, 0, True, 5), (200, 7, 1, True, 3), (513, 20, 0, True, 8), (1030, 11, 0, False, 64)])
def test_posterior_gradients_match_oracle(engine, n, d, kind, normalize_y, m):
    X, y = gp_problem(n, d, seed=n + 1)
    ls = np.linspace(0.6, 1.3, d)
    amp, alpha = 1.4, 1e-10 + 1e-4
    engine.fit(X, y, kind=kind, amplitude=amp, length_scale=ls, alpha=alpha, normalize_y=normalize_y)
    ref = ogp.NumpyGP(kind, amp, ls, alpha, normalize_y).fit(X, y)
    P = np.random.RandomState(n).rand(m, d)
    P[0] = X[0]  # a point on an observation: r = 0 must not produce NaN (Matern's 1/r cancels analytically)
    mu, sd, dmu, dsd = engine.predict_grad(P)
    mu0, sd0, dmu0, dsd0 = ref.predict_grad(P)
    assert np.all(np.isfinite(dmu)) and np.all(np.isfinite(dsd))
    np.testing.assert_allclose(mu, mu0, rtol=MU_RTOL, atol=MU_RTOL * np.abs(mu0).max())
    np.testing.assert_allclose(dmu, dmu0, rtol=GRAD_RTOL, atol=GRAD_RTOL * np.abs(dmu0).max())
    ok = sd0 > 1e-3 * sd0.max()  # sigma at an observation is sqrt of a cancellation (~1e-2 here): compare the rest tightly
    np.testing.assert_allclose(sd[ok], sd0[ok], rtol=1e-8)
    np.testing.assert_allclose(dsd[ok], dsd0[ok], rtol=1e-6, atol=1e-6 * np.abs(dsd0[ok]).max())
    # the same mu / sigma as the scoring kernel reports for these rows
    import torch
    dense = engine.score(torch.as_tensor(P).cuda(), acq_mask=1, y_opt=float(y.min()), dense=True)
    np.testing.assert_allclose(mu, dense["mu"].cpu().numpy(), rtol=1e-9, atol=1e-9 * np.abs(mu0).max())
    np.testing.assert_allclose(sd[ok], dense["sigma"].cpu().numpy()[ok], rtol=1e-7)


def test_gradients_match_sklearn_central_differences(engine):
    X, y = gp_problem(150, 4, seed=9)
    ls = np.array([0.5, 0.9, 0.7, 1.2])
    engine.fit(X, y, kind=0, amplitude=1.2, length_scale=ls, alpha=1e-5, normalize_y=True)
    sk = ogp.SklearnGP(0, 1.2, ls, 1e-5, True).fit(X, y)
    P = np.random.RandomState(3).rand(6, 4)
    _, _, dmu, dsd = engine.predict_grad(P)
    h = 1e-6
    for d in range(4):
        e = np.zeros(4)
        e[d] = h
        mp, sp = sk.predict(P + e)
        mm, sm = sk.predict(P - e)
        np.testing.assert_allclose(dmu[:, d], (mp - mm) / (2 * h), rtol=1e-5, atol=1e-6)
        np.testing.assert_allclose(dsd[:, d], (sp - sm) / (2 * h), rtol=1e-4, atol=1e-6)


def test_argument_checks(engine):
    X, y = gp_problem(20, 3, seed=1)
    engine.fit(X, y)
    with pytest.raises(ValueError):
        engine.predict_grad(np.zeros((1, 4)))
    with pytest.raises(_lib.Ob200Error, match="outside"):
        engine.predict_grad(np.zeros((65, 3)))
    blob = engine.export_state().clone()
    engine.import_state(blob, 20, 3)          # an imported fit has no dense inv(L) on this context
    with pytest.raises(_lib.Ob200Error, match="fit made on this context"):
        engine.predict_grad(np.zeros((1, 3)))
    engine.fit(X, y)


def test_refine_climbs_and_stays_in_the_box(engine):
    X, y = gp_problem(120, 5, seed=4)
    engine.fit(X, y, kind=0, amplitude=1.0, length_scale=0.6, alpha=1e-6, normalize_y=True)
    ref = ogp.NumpyGP(0, 1.0, 0.6 * np.ones(5), 1e-6, True).fit(X, y)
    starts = np.random.RandomState(0).rand(4, 5)
    y_opt = float(y.min())
    for k, name in ((0, "EI"), (1, "PI"), (2, "LCB")):
        best = acq_opt.refine(engine, name, starts, y_opt)
        assert np.all(best.x >= 0.0) and np.all(best.x <= 1.0)
        mu, sd = ref.predict(np.vstack([best.x[None], starts]))
        util = oacq.utility(k, mu, sd, y_opt)
        assert util[0] >= util[1:].max() - 1e-12          # never below the best start, judged by the oracle
        assert best.value == pytest.approx(util[0], rel=1e-7, abs=1e-12)
        assert best.n_evals > len(starts)


def test_plugin_lbfgs_mode_is_deterministic_and_valid(engine):
    s = Space()
    s.register(Real("lr", "reciprocal", 1e-5, 1.0))
    s.register(Integer("layers", "uniform", 1, 8))
    s.register(Categorical("act", ["relu", "tanh", "gelu"]))
    s.register(Real("w", "uniform", 0, 1, shape=2))
    t = build_required_space("real", s)

    def f(p):
        act, layers, lr, w = p
        return (np.log10(lr) + 3) ** 2 + 0.1 * (layers - 4) ** 2 + float(np.sum((w - 0.5) ** 2)) + (0.1 if act == "tanh" else 0)

    def run(opt, acq):
        a = BayesianOptimizer(t, seed=3, n_initial_points=6, acq_func=acq, n_points=4096, normalize_y=True, acq_optimizer=opt)
        pts = []
        for _ in range(10):
            (tp,) = a.suggest(1)
            assert tp in t
            pts.append(tp)
            a.observe([tp], [{"objective": f(t.reverse(tp))}])
        return a, pts
    for acq in ("EI", "gp_hedge"):
        a1, p1 = run("lbfgs", acq)
        a2, p2 = run("lbfgs", acq)
        assert repr(p1) == repr(p2)
        a3, p3 = run("sampling", acq)
        assert repr(p1[:6]) == repr(p3[:6]) and repr(p1[6:]) != repr(p3[6:])
        assert a1._offset == a3._offset
