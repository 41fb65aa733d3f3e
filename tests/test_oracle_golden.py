"""Pin the oracle: against the committed golden vectors (produced by the unmodified reference and by
scikit-learn, see oracle/make_golden.py), against the reference's own known-answer vectors, and --
when /root/reference is present -- against the reference run live."""
import json
import os

import numpy as np
import pytest

from metaopt_b200 import compiler
from oracle import acquisition as oacq
from oracle import gp as ogp
from oracle import philox, ref_loader, sampling

from helpers import space_from_spec, transformed

GOLD = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden")


def _cases():
    meta = json.load(open(os.path.join(GOLD, "sampling_ref.json")))
    arrays = np.load(os.path.join(GOLD, "sampling_ref.npz"))
    return [(c, arrays[c["name"] + "_variates"], arrays[c["name"] + "_flat"]) for c in meta["cases"]]


def test_philox_known_answer_vectors():
    # Random123 kat_vectors for philox4x32-10
    kat = [((0, 0, 0, 0), (0, 0), (0x6627e8d5, 0xe169c58d, 0xbc57ac4c, 0x9b00dbd8)),
           ((0xffffffff,) * 4, (0xffffffff,) * 2, (0x408f276d, 0x41c83b0e, 0xa20bc7c6, 0x6d5451fd)),
           ((0x243f6a88, 0x85a308d3, 0x13198a2e, 0x03707344), (0xa4093822, 0x299f31d0),
            (0xd16cfe09, 0x94fdcceb, 0x5001e420, 0x24126ea1))]
    for ctr, key, want in kat:
        got = philox.philox4x32_10(*ctr, *key)
        assert tuple(int(x) for x in got) == want


@pytest.mark.parametrize("case,variates,flat", _cases(), ids=lambda x: x["name"] if isinstance(x, dict) else "")
def test_oracle_reproduces_reference_samples_from_its_own_variates(case, variates, flat):
    """Reference `Space.sample(n, seed=5)` + `TransformedSpace.transform`, replayed through the oracle."""
    space = space_from_spec(case["spec"])
    assert space.keys() == case["order"]
    comp = compiler.compile_space(transformed(space))
    got, status = sampling.sample_transform(comp.slots, comp.cdf, 0, 0, case["n"], normalise=False, variates=variates)
    assert status == 0
    assert np.array_equal(got, flat)  # bit-exact, including the loguniform column (same numpy exp)


@pytest.mark.skipif(not ref_loader.available(), reason="/root/reference is only in the build container")
@pytest.mark.parametrize("case,variates,flat", _cases(), ids=lambda x: x["name"] if isinstance(x, dict) else "")
def test_golden_fixtures_are_what_the_reference_produces_now(case, variates, flat):
    import warnings
    ref_loader.load()
    from oracle import make_golden
    ref_space = make_golden.build_reference_space([tuple(s) for s in case["spec"]])
    with warnings.catch_warnings():
        warnings.simplefilter("ignore")
        points = ref_space.sample(case["n"], seed=case["seed"])
        live, _ = make_golden.flatten_reference_points(ref_space, points)
    assert np.array_equal(live, flat)
    assert np.array_equal(make_golden.replay_variates(ref_space, case["n"], case["seed"]), variates)


def test_reference_known_answers_for_transforms():
    # tests/unittests/core/test_transformer.py:224-236 (Quantize: floor), :699-712 (floor not round)
    assert np.floor(np.asarray([8.6, 5.3])).astype(int).tolist() == [8, 5]
    # :331-359 OneHotEncode.transform / reverse, k = 3, 2, 1
    assert sampling.one_hot(np.asarray(2), 3).tolist() == [0.0, 0.0, 1.0]
    assert sampling.one_hot(np.asarray([[2, 0], [1, 1]]), 3).shape == (2, 2, 3)
    assert sampling.one_hot(np.asarray(1), 2) == 1.0 and sampling.one_hot(np.asarray(0), 1) == 0.0
    assert sampling.one_hot_reverse(np.asarray([0.1, 0.7, 0.2]), 3) == 1
    assert sampling.one_hot_reverse(np.asarray(0.6), 2) == 1 and sampling.one_hot_reverse(np.asarray(0.4), 2) == 0
    assert sampling.one_hot_reverse(np.asarray(0.9), 1) == 0
    with pytest.raises(AssertionError):
        sampling.one_hot(np.asarray(3), 3)


def test_categorical_sampling_is_numpy_choice():
    """`Categorical.sample` = rng.choice(categories, p) (space.py:542-552): same indices from the same U."""
    probs = np.array([0.1, 0.2, 0.3, 0.4])
    rng = np.random.RandomState(5)
    u = np.random.RandomState(5).random_sample(1000)
    picks = rng.choice(4, p=probs, size=1000)
    slot = dict(kind=2, prior=3, out_col=0, width=4, ncat=4, cdf_off=0, p0=0, p1=0, low=0, high=4, nlo=-0.1, nhi=1.1)
    cdf = (probs.cumsum() / probs.cumsum()[-1]).tolist()
    out, _ = sampling.sample_transform([slot], cdf, 0, 0, 1000, normalise=False, variates=u[:, None])
    assert np.array_equal(out.argmax(axis=1), picks)


def test_numpy_gp_matches_sklearn_goldens():
    g = np.load(os.path.join(GOLD, "gp_sklearn.npz"))
    for tag in ("m52", "rbf"):
        n, d, kind, norm = (int(v) for v in g[tag + "_meta"])
        X, y, C, ls = g[tag + "_X"], g[tag + "_y"], g[tag + "_C"], g[tag + "_ls"]
        model = ogp.NumpyGP(kind, 1.4, ls, 1e-10 + 1e-4, bool(norm)).fit(X, y)
        np.testing.assert_allclose(model.L, g[tag + "_L"], rtol=1e-12, atol=1e-14)
        np.testing.assert_allclose(model.alpha_, g[tag + "_alpha"], rtol=1e-9)
        mu, sd = model.predict(C)
        np.testing.assert_allclose(mu, g[tag + "_mu"], rtol=1e-10)
        np.testing.assert_allclose(sd, g[tag + "_sd"], rtol=1e-9)
        y_opt = float(y.min())
        np.testing.assert_allclose(oacq.gaussian_ei(mu, sd, y_opt), g[tag + "_ei"], rtol=1e-8, atol=1e-300)
        np.testing.assert_allclose(oacq.gaussian_pi(mu, sd, y_opt), g[tag + "_pi"], rtol=1e-8, atol=1e-300)
        np.testing.assert_allclose(oacq.gaussian_lcb(mu, sd), g[tag + "_lcb"], rtol=1e-10)
        assert [oacq.pick(k, mu, sd, y_opt) for k in range(3)] == g[tag + "_pick"].tolist()
        live = ogp.SklearnGP(kind, 1.4, ls, 1e-10 + 1e-4, bool(norm)).fit(X, y)
        mu2, sd2 = live.predict(C)
        np.testing.assert_allclose(mu2, g[tag + "_mu"], rtol=1e-12)
        np.testing.assert_allclose(sd2, g[tag + "_sd"], rtol=1e-12)


def test_acquisition_formulas():
    mu, sd = np.array([0.0, 1.0, -1.0, 0.3]), np.array([1.0, 0.0, 2.0, 0.5])
    ei = oacq.gaussian_ei(mu, sd, y_opt=0.5, xi=0.01)
    assert ei[1] == 0.0  # sigma == 0 is masked
    from scipy.stats import norm
    z = (0.5 - 0.01 - 0.0) / 1.0
    assert ei[0] == pytest.approx(0.49 * norm.cdf(z) + norm.pdf(z))
    assert oacq.gaussian_lcb(mu, sd, 1.96)[2] == pytest.approx(-1 - 3.92)
    assert oacq.pick(oacq.ACQ_LCB, mu, sd, 0.5) == 2


@pytest.mark.parametrize("kind", [ogp.KERNEL_MATERN52, ogp.KERNEL_RBF])
def test_posterior_gradients_match_central_differences_of_predict(kind):
    rng = np.random.RandomState(5)
    X, y = rng.rand(40, 4), rng.randn(40) * 3 + 2
    ls = np.array([0.4, 0.7, 1.1, 0.3])
    gp = ogp.NumpyGP(kind, 1.7, ls, 1e-6, True).fit(X, y)
    P = rng.rand(5, 4)
    mu, sd, dmu, dsd = gp.predict_grad(P)
    mu0, sd0 = gp.predict(P)
    np.testing.assert_allclose(mu, mu0, rtol=1e-12)
    np.testing.assert_allclose(sd, sd0, rtol=1e-10)
    h = 1e-6
    for d in range(4):
        e = np.zeros(4)
        e[d] = h
        mp, sp = gp.predict(P + e)
        mm, sm = gp.predict(P - e)
        np.testing.assert_allclose(dmu[:, d], (mp - mm) / (2 * h), rtol=2e-6, atol=1e-7)
        np.testing.assert_allclose(dsd[:, d], (sp - sm) / (2 * h), rtol=2e-5, atol=1e-7)
