"""GPU parity of the sampler/transform kernel against the oracle (bit-exact where the reference is)."""
import numpy as np
import pytest
import torch

from metaopt_b200 import compiler
from oracle import sampling

from helpers import cfg3_space, cfg5_space, mixed_space, transformed

pytestmark = pytest.mark.gpu

# loguniform (exp) and normal (log/sqrt/cos Box-Muller) real columns go through device libm vs glibc:
# a few ulp; every other column (uniform, integer, categorical/one-hot, normalised or not) is bit-exact.
ULP_COLS_RTOL = 2e-15
ULP_COLS_ATOL = 2e-15


def _exact_and_ulp_cols(comp):
    exact, ulp = [], []
    for s in comp.slots:
        cols = list(range(s["out_col"], s["out_col"] + s["width"]))
        (ulp if s["prior"] in (1, 2) and s["kind"] == 0 else exact).extend(cols)
    return exact, ulp


def _check(dev, ref, comp):
    exact, ulp = _exact_and_ulp_cols(comp)
    assert np.array_equal(dev[:, exact], ref[:, exact])
    if ulp:
        np.testing.assert_allclose(dev[:, ulp], ref[:, ulp], rtol=ULP_COLS_RTOL, atol=ULP_COLS_ATOL)


@pytest.mark.parametrize("space_fn", [mixed_space, cfg3_space, cfg5_space])
@pytest.mark.parametrize("normalise", [False, True])
def test_philox_sampling_matches_oracle(engine, space_fn, normalise):
    comp = compiler.compile_space(transformed(space_fn()))
    compiler.upload(engine, comp)
    seed, first, count = 0x1234567890ABCDEF, 987654321012, 4099
    dev = engine.sample(seed, first, count, normalise=normalise).cpu().numpy()
    ref, status = sampling.sample_transform(comp.slots, comp.cdf, seed, first, count, normalise=normalise)
    assert status == 0
    assert dev.shape == (count, comp.d_out)
    _check(dev, ref, comp)


def test_integer_columns_exact_even_with_normal_prior(engine):
    # An integer drawn from a normal prior floors a libm-dependent value (Box-Muller through log / cos / sqrt): device
    # and numpy can differ in the last ulp of z, which moves floor() only if loc + scale z lands within ~1e-15 of an
    # integer -- probability ~1e-14 per draw.  These 1e5 draws are identical.
    comp = compiler.compile_space(transformed(mixed_space()))
    compiler.upload(engine, comp)
    dev = engine.sample(7, 0, 100000, normalise=False).cpu().numpy()
    ref, _ = sampling.sample_transform(comp.slots, comp.cdf, 7, 0, 100000, normalise=False)
    col = [s["out_col"] for s in comp.slots if s["kind"] == 1 and s["prior"] == 2][0]
    assert np.all(dev[:, col] == np.floor(dev[:, col]))
    assert dev[:, col].min() >= 0 and dev[:, col].max() < 20
    assert np.array_equal(dev[:, col], ref[:, col])


def test_loguniform_integer_columns_are_exact(engine):
    """Integer dimension with a log-uniform prior: floor(exp(.)) with the device's table-based exp (<= 2 ulp) against
    numpy's.  A flip needs exp(.) within 2 ulp of an integer: < 1e-12 per draw for values below 1000;
    one million draws, all identical."""
    from metaopt_b200.space import Integer, Real, Space
    s = Space()
    s.register(Integer("k", "reciprocal", 1, 1000))
    s.register(Integer("m", "reciprocal", 2, 64, shape=2))
    s.register(Real("x", "reciprocal", 1e-5, 1.0))
    comp = compiler.compile_space(transformed(s))
    compiler.upload(engine, comp)
    dev = engine.sample(21, 5, 1_000_000, normalise=False).cpu().numpy()
    ref, status = sampling.sample_transform(comp.slots, comp.cdf, 21, 5, 1_000_000, normalise=False)
    assert status == 0
    ints = [s_["out_col"] for s_ in comp.slots if s_["kind"] == 1]
    assert len(ints) == 3
    assert np.array_equal(dev[:, ints], ref[:, ints])
    assert dev[:, ints].min() >= 1 and dev[:, ints[0]].max() < 1000


def test_sharding_reproduces_the_same_points(engine):
    comp = compiler.compile_space(transformed(cfg5_space()))
    compiler.upload(engine, comp)
    whole = engine.sample(11, 0, 3000).cpu().numpy()
    parts = [engine.sample(11, a, b - a).cpu().numpy() for a, b in ((0, 1), (1, 700), (700, 2048), (2048, 3000))]
    assert np.array_equal(whole, np.concatenate(parts))


def test_strided_output_equals_dense(engine):
    comp = compiler.compile_space(transformed(mixed_space()))
    compiler.upload(engine, comp)
    dense = engine.sample(5, 100, 1001)
    wide = torch.full((1001, comp.d_out + 3), -7.0, dtype=torch.float64, device=dense.device)
    engine.sample(5, 100, 1001, out=wide[:, :comp.d_out])
    assert torch.equal(wide[:, :comp.d_out], dense)
    assert torch.all(wide[:, comp.d_out:] == -7.0)


def test_odd_width_and_single_row(engine):
    from metaopt_b200.space import Real, Space
    s = Space()
    for i in range(3):
        s.register(Real("x%d" % i, "uniform", 0, 1))
    comp = compiler.compile_space(transformed(s))
    compiler.upload(engine, comp)
    for count in (1, 2, 63, 64, 65, 129):
        dev = engine.sample(3, 5, count).cpu().numpy()
        ref, _ = sampling.sample_transform(comp.slots, comp.cdf, 3, 5, count)
        assert np.array_equal(dev, ref)


def test_host_variates_mode_matches_oracle(engine):
    comp = compiler.compile_space(transformed(mixed_space()))
    compiler.upload(engine, comp)
    rs = np.random.RandomState(5)
    var = rs.uniform(size=(777, comp.n_slots))
    for i, s in enumerate(comp.slots):
        if s["prior"] == 2:
            var[:, i] = rs.randn(777) * 0.3
    dev = engine.sample(0, 0, 777, normalise=True, variates=var).cpu().numpy()
    ref, status = sampling.sample_transform(comp.slots, comp.cdf, 0, 0, 777, normalise=True, variates=var)
    assert status == 0
    _check(dev, ref, comp)
    host = engine.sample_host(0, 0, 777, normalise=True, variates=var)
    assert np.array_equal(host, dev)


def test_host_entry_point_equals_device(engine):
    comp = compiler.compile_space(transformed(cfg3_space()))
    compiler.upload(engine, comp)
    assert np.array_equal(engine.sample_host(9, 17, 2500), engine.sample(9, 17, 2500).cpu().numpy())


def test_improbable_bounds_raise_value_error(engine):
    from metaopt_b200.space import Real, Space
    s = Space()
    s.register(Real("x", "norm", 0, 1, low=50, high=51))
    comp = compiler.compile_space(transformed(s))
    compiler.upload(engine, comp)
    with pytest.raises(ValueError, match="Improbable bounds"):
        engine.sample(1, 0, 100, normalise=False)
    # the latch is cleared: a sane space samples fine afterwards
    comp = compiler.compile_space(transformed(cfg3_space()))
    compiler.upload(engine, comp)
    engine.sample(1, 0, 100)


def test_distribution_of_generated_samples(engine):
    from scipy import stats
    comp = compiler.compile_space(transformed(mixed_space()))
    compiler.upload(engine, comp)
    x = engine.sample(2024, 0, 200000, normalise=False).cpu().numpy()
    by_name = {item["name"]: item for item in comp.layout}
    a = x[:, by_name["a_real"]["col"]]
    assert stats.kstest(a, stats.uniform(-3, 9).cdf).pvalue > 1e-3
    b = x[:, by_name["b_log"]["col"]]
    assert stats.kstest(b, stats.reciprocal(1e-5, 1.0).cdf).pvalue > 1e-3
    d = x[:, by_name["d_norm"]["col"]]
    tn = stats.truncnorm((-4 - 0.5) / 2, (5 - 0.5) / 2, loc=0.5, scale=2.0)
    assert stats.kstest(d, tn.cdf).pvalue > 1e-3
    c0 = by_name["e_cat4"]["col"]
    freq = x[:, c0:c0 + 4].mean(axis=0)
    np.testing.assert_allclose(freq, [0.1, 0.2, 0.3, 0.4], atol=0.01)
    assert np.all(x[:, c0:c0 + 4].sum(axis=1) == 1.0)
    ci = x[:, by_name["c_int"]["col"]]
    assert set(np.unique(ci)) == {-3.0, -2.0, -1.0, 0.0, 1.0, 2.0}


def test_normalisation_division_is_ieee_exact_at_scale(engine):
    """The kernel divides by each slot's constant span with a reciprocal + exact-residual correction
    instead of an IEEE division; numpy divides.  Check bit equality on 24M quotients over awkward spans."""
    from metaopt_b200.space import Categorical, Integer, Real, Space
    s = Space()
    spans = [(0.0, 1.0), (-3.0, 9.0), (0.1, 0.30000000000000004), (-1e-3, 7e5), (2.0, 2.0000000000000004 + 1e-9),
             (-0.1, 1.1), (1e-5, 1.0), (123.456, 789.0123), (0.3333333333333333, 0.6666666666666666), (-7.0, 1e-7),
             (5e-324 + 1.0, 3.0), (1.0, 1.9999999999999998)]
    for i, (a, b) in enumerate(spans):
        s.register(Real("r%02d" % i, "uniform", a, b - a))
    for i in range(6):
        s.register(Integer("i%02d" % i, "uniform", -5 * i, 17 + 100 * i))
    for i in range(6):
        s.register(Categorical("c%02d" % i, list(range(2 + i))))
    comp = compiler.compile_space(transformed(s))
    compiler.upload(engine, comp)
    n = 1_000_000
    dev = engine.sample(31337, 5, n, normalise=True).cpu().numpy()
    ref, status = sampling.sample_transform(comp.slots, comp.cdf, 31337, 5, n, normalise=True)
    assert np.array_equal(dev, ref)
