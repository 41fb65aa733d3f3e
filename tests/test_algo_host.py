"""Plugin host logic (no GPU): configuration keys, state, deepcopy, observe bookkeeping, the suggest
flow.  The device is replaced by the oracle-backed `OracleEngine` stand-in (tests only)."""
import copy

import numpy as np
import pytest

from metaopt_b200 import engine as engine_mod
from metaopt_b200.algo import BayesianOptimizer
from metaopt_b200.space import Categorical, Integer, Real, Space, build_required_space

from fake_engine import OracleEngine


@pytest.fixture()
def fake(monkeypatch):
    eng = OracleEngine()
    monkeypatch.setattr(engine_mod, "get_engine", lambda device=None: eng)
    return eng


def small_space():
    s = Space()
    s.register(Real("x", "uniform", -2, 4))
    s.register(Integer("n", "uniform", 1, 8))
    s.register(Categorical("act", ["relu", "tanh", "gelu"]))
    return s


def objective(point):
    act, n, x = point  # dimensions come in sorted-name order: act, n, x
    return (x - 0.5) ** 2 + 0.1 * (n - 3) ** 2 + {"relu": 0.0, "tanh": 0.3, "gelu": 0.1}[act]


def test_reference_config_keys_and_defaults():
    # docs/src/user/algorithms.rst:157-167 and tests/unittests/core/cli/test_info.py:89-98
    a = BayesianOptimizer(None)
    conf = a.configuration["bayesianoptimizer"]
    for key, default in dict(n_initial_points=10, acq_func="gp_hedge", alpha=1e-10, n_restarts_optimizer=0,
                             noise="gaussian", normalize_y=False).items():
        assert conf[key] == default
    assert "seed" in conf and conf["n_points"] == 10000
    assert BayesianOptimizer.requires == "real"
    with pytest.raises(ValueError):
        BayesianOptimizer(None, acq_func="EIps")


def test_no_device_work_in_constructor_and_cheap_deepcopy(fake):
    a = BayesianOptimizer(build_required_space("real", small_space()), seed=3)
    assert fake.calls == {"sample": 0, "fit": 0, "score": 0}
    b = copy.deepcopy(a)
    assert b is not a and b._space is a._space and b.state_dict == a.state_dict
    b._X.append(np.zeros(3))
    assert len(a._X) == 0
    assert "engine" not in " ".join(a.__dict__).lower()  # no device handle inside the instance


def test_seeded_suggestions_are_reproducible_and_state_follows(fake):
    t = build_required_space("real", small_space())
    a, b = BayesianOptimizer(t, seed=7), BayesianOptimizer(t, seed=7)
    pa, pb = a.suggest(3), b.suggest(3)
    assert repr(pa) == repr(pb)
    for p in pa:
        assert p in t
    c = BayesianOptimizer(t, seed=1)
    c.set_state(a.state_dict)
    assert repr(c.suggest(1)) == repr(a.suggest(1))
    a.seed_rng(7)
    assert repr(a.suggest(3)) == repr(pa)


@pytest.mark.parametrize("acq", ["EI", "PI", "LCB", "gp_hedge"])
def test_observe_then_gp_suggest(fake, acq):
    space = small_space()
    t = build_required_space("real", space)
    a = BayesianOptimizer(t, seed=11, n_initial_points=5, acq_func=acq, n_points=500, normalize_y=True,
                          fit_hyperparameters=False)
    for _ in range(8):
        (tp,) = a.suggest(1)
        assert tp in t
        a.observe([tp], [{"objective": objective(t.reverse(tp)), "constraint": [], "gradient": None}])
    # the hedge portfolio scores EI, LCB and PI in ONE pruned call (shared mean-only pass)
    assert len(a._y) == 8 and fake.calls["fit"] == 3 and fake.calls["score"] == 3
    # lies without an objective are ignored, whole-history observe is accepted
    a.observe([tp], [{"objective": None}])
    assert len(a._y) == 8
    d = copy.deepcopy(a)
    d.observe([tp], [{"objective": 1.0}])
    assert len(a._y) == 8 and len(d._y) == 9
    # the engine's fit is keyed by the data: the copy refits, the original reuses
    n_fit = fake.calls["fit"]
    d.suggest(1)
    assert fake.calls["fit"] == n_fit + 1
    a.suggest(1)
    assert fake.calls["fit"] == n_fit + 2  # engine now holds the copy's fit -> original refits once
    a.suggest(1)
    assert fake.calls["fit"] == n_fit + 2


def test_unbounded_space_is_rejected(fake):
    s = Space()
    s.register(Real("x", "norm", 0, 1))
    a = BayesianOptimizer(build_required_space("real", s), seed=1)
    with pytest.raises(ValueError, match="finite interval"):
        a.suggest(1)


def test_mirror_base_matches_reference_base_behaviour():
    a = BayesianOptimizer(None, seed=5, n_points=123)
    assert a.n_points == 123 and a.is_done is False and a.score(None) == 0 and a.judge(None, None) is None
    assert a.should_suspend is False
    assert list(a.configuration) == ["bayesianoptimizer"]


def test_hyperparameters_are_fitted_and_warm_started(fake):
    space = small_space()
    t = build_required_space("real", space)
    a = BayesianOptimizer(t, seed=2, n_initial_points=6, acq_func="EI", n_points=300, normalize_y=True)
    assert a.configuration["bayesianoptimizer"]["fit_hyperparameters"] is True
    for _ in range(9):
        (tp,) = a.suggest(1)
        a.observe([tp], [{"objective": objective(t.reverse(tp))}])
    amp, ls, noise = a._theta
    assert len(ls) == 5 and 1e-2 <= amp <= 1e3 and np.all((ls >= 1e-2 * (1 - 1e-9)) & (ls <= 1e2 * (1 + 1e-9))) and 1e-5 * (1 - 1e-9) <= noise <= 1e5 * (1 + 1e-9)
    assert fake.calls["fit"] > 3  # several likelihood evaluations per refit
    b = copy.deepcopy(a)
    assert b._theta[0] == a._theta[0] and b._theta[1] is not a._theta[1]


def test_batch_suggest_uses_constant_liar_and_leaves_no_lies(fake):
    space = small_space()
    t = build_required_space("real", space)
    a = BayesianOptimizer(t, seed=3, n_initial_points=5, acq_func="EI", n_points=400, normalize_y=True)
    for _ in range(6):
        (tp,) = a.suggest(1)
        a.observe([tp], [{"objective": objective(t.reverse(tp))}])
    b = copy.deepcopy(a)
    b.suggest(1)                                        # what a plain suggest fits on the real data
    theta = copy.deepcopy(b._theta)
    batch = a.suggest(4)
    assert len(batch) == 4 and all(p in t for p in batch)
    assert len({repr(p) for p in batch}) == 4          # the lies push the picks apart
    assert len(a._y) == 6 and len(a._X) == 6           # lies are gone
    assert a._theta[0] == theta[0] and a.fit_hyperparameters is True  # hyper-parameters untouched by lies


@pytest.mark.parametrize("name", ["EI", "PI", "LCB"])
def test_utility_gradients_match_central_differences(name):
    from metaopt_b200 import acq_opt
    from oracle import acquisition as oacq
    from oracle import gp as ogp
    rng = np.random.RandomState(2)
    gp = ogp.NumpyGP(0, 1.3, np.array([0.5, 0.8, 0.6]), 1e-6, False).fit(rng.rand(30, 3), rng.randn(30))
    k = {"EI": 0, "PI": 1, "LCB": 2}[name]
    x = rng.rand(3)

    def util(p):
        mu, sd = gp.predict(p[None])
        return float(oacq.utility(k, mu, sd, -0.3, 0.01, 1.96)[0])
    mu, sd, dmu, dsd = gp.predict_grad(x[None])
    val, grad = acq_opt.utility_and_grad(name, mu[0], sd[0], dmu[0], dsd[0], -0.3, 0.01, 1.96)
    assert val == pytest.approx(util(x), rel=1e-10)
    for d in range(3):
        e = np.zeros(3)
        e[d] = 1e-6
        assert grad[d] == pytest.approx((util(x + e) - util(x - e)) / 2e-6, rel=1e-4, abs=1e-8)
    assert acq_opt.utility_and_grad("EI", 0.0, 0.0, dmu[0], dsd[0], 0.0)[0] == 0.0  # sigma == 0 is masked


@pytest.mark.parametrize("acq", ["EI", "gp_hedge"])
def test_lbfgs_refinement_never_loses_to_the_best_sample(fake, acq):
    s = Space()
    s.register(Real("x", "uniform", -2, 4))
    s.register(Real("z", "uniform", 0, 1))
    t = build_required_space("real", s)

    def run(opt):
        a = BayesianOptimizer(t, seed=5, n_initial_points=6, acq_func=acq, n_points=300, fit_hyperparameters=False,
                              acq_optimizer=opt, acq_n_restarts=3)
        for _ in range(6):
            (p,) = a.suggest(1)
            a.observe([p], [{"objective": (p[0] - 0.5) ** 2 + (p[1] - 0.2) ** 2}])
        return a, a.suggest(1)[0]
    a_s, p_s = run("sampling")
    a_l, p_l = run("lbfgs")
    assert p_l in t and p_s in t
    assert a_s._offset == a_l._offset            # the refinement consumes no stream positions
    if acq == "EI":
        comp = a_l._get_compiled()
        gp = fake.model

        def ei(p):
            mu, sd = gp.predict(comp.normalise(np.asarray(p, float))[None])
            from oracle import acquisition as oacq
            return float(oacq.gaussian_ei(mu, sd, min(a_l._y))[0])
        assert ei(p_l) >= ei(p_s) - 1e-15
        assert ei(p_l) > ei(p_s)                  # 300 samples in 2-D leave room for the ascent
    with pytest.raises(ValueError):
        BayesianOptimizer(t, acq_optimizer="newton")


def test_observation_cache_is_path_independent_and_survives_copies_and_lies(fake):
    t = build_required_space("real", small_space())
    a = BayesianOptimizer(t, seed=3, n_initial_points=4, acq_func="EI", n_points=200, fit_hyperparameters=False)
    pts = a.suggest(4)
    ys = [objective(t.reverse(p)) for p in pts]
    for p, v in zip(pts, ys):                                   # one by one
        a.observe([p], [{"objective": v}])
        a._data()
    Xa, ya, da = a._data()
    b = BayesianOptimizer(t, seed=3, n_initial_points=4, acq_func="EI", n_points=200, fit_hyperparameters=False)
    b.observe(pts, [{"objective": v} for v in ys])              # all at once (restart)
    Xb, yb, db = b._data()
    assert np.array_equal(Xa, Xb) and np.array_equal(ya, yb) and da == db
    c = copy.deepcopy(a)
    c.observe([pts[0]], [{"objective": 9.0}])
    assert c._data()[0].shape[0] == 5 and a._data()[0].shape[0] == 4 and a._data()[2] == da
    n_fit = fake.calls["fit"]
    a.suggest(3)                                                # two constant-liar refits on top of the real fit
    assert fake.calls["fit"] == n_fit + 3
    assert len(a._y) == 4 and a._data()[2] == da                # lies and their cache entries are gone
    a.suggest(1)
    assert fake.calls["fit"] == n_fit + 4                       # the engine holds the last liar fit: one refit, then cached
    a.suggest(1)
    assert fake.calls["fit"] == n_fit + 4


def test_placement_and_speed_knobs_stay_out_of_the_stored_configuration(monkeypatch):
    """Orion stores `configuration` as `experiment.algorithms` and raises AlgorithmConflict when two workers differ
    there (`src/orion/core/evc/conflicts.py:1034-1038`): which GPU a worker uses must not be part of it."""
    a = BayesianOptimizer(None, device=3, prune=False)
    b = BayesianOptimizer(None, device=[0, 1])
    assert a.configuration == b.configuration
    conf = a.configuration["bayesianoptimizer"]
    assert "device" not in conf and "prune" not in conf and not any(k.startswith("_") for k in conf)
    assert a.device == 3 and a.prune is False and b.device == [0, 1] and b.prune is True
    monkeypatch.setenv("OB200_DEVICE", "2,5")
    monkeypatch.setenv("OB200_PRUNE", "0")
    c = BayesianOptimizer(None)
    assert c.device == [2, 5] and c.prune is False and c.configuration == a.configuration
    monkeypatch.setenv("OB200_DEVICE", "1")
    assert BayesianOptimizer(None).device == 1
    assert copy.deepcopy(c).device == [2, 5]


def test_hedge_gains_are_settled_with_the_refitted_model(fake):
    """skopt: `gains_ -= est.predict(next_xs_)` with `est` refitted on the new observation."""
    t = build_required_space("real", small_space())
    a = BayesianOptimizer(t, seed=2, n_initial_points=4, acq_func="gp_hedge", n_points=200, fit_hyperparameters=False)
    for p in a.suggest(4):
        a.observe([p], [{"objective": objective(t.reverse(p))}])
    (p,) = a.suggest(1)
    pending = np.asarray(a._hedge_pending[1:])
    assert a._hedge_pending[0] == 4 and pending.shape == (3, 5) and a._hedge_gains == [0.0, 0.0, 0.0]
    a.observe([p], [{"objective": objective(t.reverse(p))}])
    assert a._hedge_gains == [0.0, 0.0, 0.0]            # nothing happens before the refit
    b = copy.deepcopy(a)                                 # what the producer does: the copy suggests ...
    b.suggest(1)
    mu_new = fake.model.predict(pending)[0]              # ... after refitting on 5 observations
    assert len(fake.model.X) == 5
    np.testing.assert_allclose(b._hedge_gains, -mu_new, rtol=1e-12)
    a.set_state(b.state_dict)                            # ... and its state comes back (`producer.py:84`)
    np.testing.assert_allclose(a._hedge_gains, -mu_new, rtol=1e-12)
    # a constant-liar batch leaves the portfolio as the first (real-data) pick left it
    gains = list(a._hedge_gains)
    a.observe([p], [{"objective": 0.123}])
    a.suggest(3)
    settled = list(a._hedge_gains)
    assert settled != gains and len(a._hedge_pending) == 4 and a._hedge_pending[0] == 6
    c = copy.deepcopy(a)
    c.suggest(1)                                         # same data, no refit of new data -> pending replaced, gains kept
    assert c._hedge_gains == settled
