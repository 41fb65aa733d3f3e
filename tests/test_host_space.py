"""Host-side mirror of the reference's Space/Dimension/transformed-space interface, and the
SpaceCompiler.  Known answers are the reference's own (`tests/unittests/algo/test_space.py`,
`tests/unittests/core/test_transformer.py`); when /root/reference is present every mirror object is
also compared with the real Orion object built from the same arguments."""
import warnings

import numpy as np
import pytest

from metaopt_b200 import _lib, compiler
from metaopt_b200.space import (Categorical, Integer, Real, Space, TransformedSpace, build_required_space)
from oracle import ref_loader

from helpers import cfg3_space, cfg5_space, mixed_space, transformed


def test_real_interval_and_contains():
    # test_space.py:211-220: uniform(-3, 4) -> interval (-3, 1); low/high clip
    dim = Real("yolo", "uniform", -3, 4)
    assert dim.interval() == (-3.0, 1.0)
    assert 0.999 in dim and 1.0 not in dim and -3.0 in dim and -3.1 not in dim
    dim = Real("yolo", "norm", 0, 2, low=-3, high=+3)
    assert dim.interval() == (-3, 3)
    assert "a string" not in dim
    with pytest.raises(ValueError, match="Lower bound"):
        Real("yolo", "norm", 0, 2, low=3, high=2)


def test_integer_interval_floor_semantics():
    # test_space.py:252-260: Integer('uniform', -3, 6) -> interval (-3, 3)
    dim = Integer("yolo", "uniform", -3, 6)
    assert dim.interval() == (-3, 3)
    assert 2 in dim and 3 not in dim and 1.5 not in dim and -3 in dim
    # _Discrete.interval (space.py:425-448): a fractional upper bound gains one
    assert Integer("y", "uniform", 0.5, 2.0).interval() == (0, 3)
    assert Integer("y", "norm", 0, 1).interval() == (-np.inf, np.inf)


def test_categorical_basics():
    dim = Categorical("yolo", {"asdfa": 0.1, 2: 0.2, 3: 0.3, 4: 0.4})
    assert dim.categories == ("asdfa", 2, 3, 4) and dim._probs == (0.1, 0.2, 0.3, 0.4)
    assert 2 in dim and "nope" not in dim
    with pytest.raises(RuntimeError, match="Categories have no ``interval``"):
        dim.interval()
    uni = Categorical("u", ["a", "b", "c", "d"])
    assert uni._probs == (0.25, 0.25, 0.25, 0.25)
    assert uni.cast(["a", "d"]) == ["a", "d"]
    with pytest.raises(ValueError, match="Invalid category"):
        uni.cast(["z"])


def test_dimension_argument_checks():
    with pytest.raises(ValueError, match="random_state/seed"):
        Real("x", "uniform", 0, 1, seed=3)
    with pytest.raises(ValueError, match="shape"):
        Real("x", "uniform", 0, 1, size=3)
    with pytest.raises(TypeError):
        Real(3, "uniform", 0, 1)
    with pytest.raises(NotImplementedError):
        Real("x", "gamma", 1.0)
    with pytest.raises(ValueError, match="default value"):
        Real("x", "uniform", 0, 1, default_value=5)


def test_space_is_sorted_and_validates():
    s = Space()
    s.register(Real("b", "uniform", 0, 1))
    s.register(Integer("a", "uniform", 0, 10))
    assert s.keys() == ["a", "b"] and [d.name for d in s.values()] == ["a", "b"] and s[0].name == "a"
    with pytest.raises(ValueError, match="already"):
        s.register(Real("a", "uniform", 0, 1))
    with pytest.raises(TypeError):
        s["c"] = 3
    assert (5, 0.5) in s and (5.5, 0.5) not in s and "a" in s
    assert s.interval() == [(0, 10), (0.0, 1.0)]


def test_transformed_space_known_answers():
    # test_transformer.py:477-491 intervals; :135-159 categorical -> one-hot; :699-712 floor not round
    space = Space()
    space.register(Categorical("c", ["asdfa", 2, 3, 4]))
    space.register(Integer("i", "uniform", -3, 6))
    space.register(Real("r", "uniform", -3, 4))
    space.register(Categorical("b", ["x", "y"]))
    t = build_required_space("real", space)
    assert isinstance(t, TransformedSpace) and t.keys() == ["b", "c", "i", "r"]
    assert t["c"].interval() == (-0.1, 1.1) and t["b"].interval() == (-0.1, 1.1)
    assert t["c"].shape == (4,) and t["b"].shape == () and t["i"].type == "real"
    lo, hi = t["i"].interval()
    assert (float(lo), float(hi)) == (-3.0, 3.0)
    point = ("y", 3, 2, 0.5)
    tp = t.transform(point)
    assert tp[0] == 1.0 and tp[1].tolist() == [0, 0, 1, 0] and tp[2] == 2.0 and tp[3] == 0.5
    assert tp in t
    back = t.reverse((0.7, np.array([0.1, 0.2, 0.9, 0.3]), 2.9, 0.5))
    assert back[0] == "y" and back[1] == 3 and back[2] == 2 and back[3] == 0.5
    assert (0.7, np.array([0.1, 0.2, 0.9, 0.3]), 3.5, 0.5) not in t  # floor(3.5) = 3 is outside [-3, 3)
    with pytest.raises(TypeError):
        build_required_space("integer", space)


def test_compile_space_layout():
    comp = compiler.compile_space(transformed(mixed_space()))
    names = [item["name"] for item in comp.layout]
    assert names == sorted(names)
    assert comp.d_out == 1 + 1 + 1 + 1 + 4 + 1 + 1 + 3 + 1 + 3 and comp.n_slots == 12
    cat4 = [s for s in comp.slots if s["ncat"] == 4][0]
    assert cat4["width"] == 4 and (cat4["nlo"], cat4["nhi"]) == (-0.1, 1.1)
    np.testing.assert_allclose(comp.cdf[cat4["cdf_off"]:cat4["cdf_off"] + 4], [0.1, 0.3, 0.6, 1.0])
    log = [s for s in comp.slots if s["prior"] == _lib.PRIOR_LOGUNIFORM][0]
    assert log["p0"] == np.log(1e-5) and log["p1"] == np.log(1.0) - np.log(1e-5)
    integer = [s for s in comp.slots if s["kind"] == _lib.KIND_INT and s["prior"] == 0][0]
    assert (integer["low"], integer["high"], integer["nlo"], integer["nhi"]) == (-3.0, 3.0, -3.0, 3.0)
    assert comp.normalisable
    assert compiler.compile_space(transformed(cfg3_space())).d_out == 60
    assert compiler.compile_space(transformed(cfg5_space())).d_out == 98


def test_unbounded_prior_is_flagged_not_normalisable():
    s = Space()
    s.register(Real("x", "norm", 0, 1))
    assert not compiler.compile_space(transformed(s)).normalisable


def test_regroup_and_flatten_roundtrip():
    space = mixed_space()
    t = transformed(space)
    comp = compiler.compile_space(t)
    row = np.zeros(comp.d_out)
    by = {item["name"]: item for item in comp.layout}
    row[by["a_real"]["col"]] = 1.25
    row[by["c_int"]["col"]] = -2.0
    row[by["e_cat4"]["col"] + 2] = 1.0
    row[by["f_cat2"]["col"]] = 1.0
    row[by["h_vec"]["col"]:by["h_vec"]["col"] + 3] = [0.1, 0.2, 0.3]
    row[by["i_intnorm"]["col"]] = 7.0
    row[by["j_cat3"]["col"]] = 1.0
    row[by["b_log"]["col"]] = 1e-3
    tp = compiler.regroup_row(comp, row, transformed=True)
    assert tp in t
    assert compiler.flatten_point(comp, tp) == row.tolist()
    op = compiler.regroup_row(comp, row, transformed=False)
    assert op in space
    assert op[by_index(space, "e_cat4")] == "c" and op[by_index(space, "f_cat2")] == "y"
    assert op == tuple(t.reverse(tp)[i] if not isinstance(t.reverse(tp)[i], np.ndarray) else op[i] for i in range(len(op)))
    flat = np.asarray(compiler.flatten_point(comp, tp))
    np.testing.assert_allclose(comp.denormalise(comp.normalise(flat)), flat, rtol=1e-15, atol=1e-15)


def by_index(space, name):
    return space.keys().index(name)


def test_fold_seed():
    assert compiler.fold_seed(5) == 5
    assert compiler.fold_seed((1, 2, 3)) == compiler.fold_seed([1, 2, 3]) != compiler.fold_seed((3, 2, 1))
    assert 0 <= compiler.fold_seed(None) < 2 ** 64


# ---------------------------------------------------------------------------------------------
# live comparison with the unmodified reference
# ---------------------------------------------------------------------------------------------
needs_ref = pytest.mark.skipif(not ref_loader.available(), reason="/root/reference is only in the build container")


def _ref_and_mirror():
    ref_loader.load()
    from orion.algo.space import Categorical as RC, Integer as RI, Real as RR, Space as RS
    specs = [("real", "a", "uniform", (-3, 9), {}), ("real", "b", "reciprocal", (1e-5, 1.0), {}),
             ("integer", "c", "uniform", (-3, 6), {}), ("real", "d", "norm", (0.5, 2.0), {"low": -4.0, "high": 5.0}),
             ("integer", "e", "norm", (10, 3), {"low": 0, "high": 20}), ("real", "f", "uniform", (0, 1), {"shape": 3}),
             ("integer", "g", "uniform", (0.5, 2.0), {})]
    cats = [("h", {"p": 0.1, "q": 0.2, "r": 0.7}), ("i", ["x", "y"]), ("j", ["only"])]
    rs, ms = RS(), Space()
    for kind, name, prior, args, kw in specs:
        rs.register((RR if kind == "real" else RI)(name, prior, *args, **kw))
        ms.register((Real if kind == "real" else Integer)(name, prior, *args, **kw))
    for name, c in cats:
        rs.register(RC(name, c))
        ms.register(Categorical(name, c))
    return rs, ms


@needs_ref
def test_mirror_matches_reference_objects():
    rs, ms = _ref_and_mirror()
    assert rs.keys() == ms.keys()
    rng = np.random.RandomState(0)
    for rd, md in zip(rs.values(), ms.values()):
        assert rd.type == md.type and tuple(rd.shape or ()) == md.shape, rd.name
        if rd.type != "categorical":
            ri, mi = rd.interval(), md.interval()
            assert (float(ri[0]), float(ri[1])) == (float(mi[0]), float(mi[1])), rd.name
            for v in list(rng.uniform(-6, 25, size=40)) + [ri[0], ri[1], 0, 1, 2.0, -3]:
                v = np.full(md.shape, v) if md.shape else v
                assert bool(v in rd) == bool(v in md), (rd.name, v)
        else:
            assert rd.categories == md.categories and tuple(rd._probs) == tuple(md._probs)
            for v in list(rd.categories) + ["zzz", 3]:
                assert bool(v in rd) == bool(v in md)


@needs_ref
def test_compiled_table_is_identical_for_reference_and_mirror_spaces():
    from orion.core.worker.transformer import build_required_space as ref_build
    rs, ms = _ref_and_mirror()
    with warnings.catch_warnings():
        warnings.simplefilter("ignore")
        a = compiler.compile_space(ref_build("real", rs))
    b = compiler.compile_space(build_required_space("real", ms))
    assert a.key == b.key and a.d_out == b.d_out
    assert [(i["name"], i["kind"], i["col"], i["m"], i["k"]) for i in a.layout] == \
           [(i["name"], i["kind"], i["col"], i["m"], i["k"]) for i in b.layout]


@needs_ref
def test_transformed_mirror_matches_reference_transformed_space():
    from orion.core.worker.transformer import build_required_space as ref_build
    rs, ms = _ref_and_mirror()
    rt, mt = ref_build("real", rs), build_required_space("real", ms)
    comp = compiler.compile_space(mt)
    rng = np.random.RandomState(3)
    with warnings.catch_warnings():
        warnings.simplefilter("ignore")
        for _ in range(50):
            row = np.zeros(comp.d_out)
            for item in comp.layout:
                c0, m, k = item["col"], item["m"], item["k"]
                if item["kind"] == "categorical":
                    if k > 2:
                        row[c0 + rng.randint(k)] = 1.0
                    else:
                        row[c0] = float(rng.randint(max(k, 1)))
                else:
                    s = [s for s in comp.slots if s["out_col"] == c0][0]
                    v = rng.uniform(max(s["low"], -50), min(s["high"], 50), size=m)
                    row[c0:c0 + m] = np.floor(v) if item["kind"] == "integer" else v
            tp = compiler.regroup_row(comp, row, transformed=True)
            assert (tp in rt) and (tp in mt)
            r_back, m_back = rt.reverse(tp), mt.reverse(tp)
            for x, y in zip(r_back, m_back):
                assert np.array_equal(np.asarray(x, dtype=object), np.asarray(y, dtype=object))
            # forward transform agrees wherever the reference's runs on numpy 2 (one-hot k>2 does not)
            for i, (rd, md) in enumerate(zip(rt.values(), mt.values())):
                if rd.original_dimension.type == "categorical" and len(rd.original_dimension.categories) > 2:
                    continue
                assert np.array_equal(np.asarray(rd.transform(r_back[i]), dtype=float),
                                      np.asarray(md.transform(m_back[i]), dtype=float))
        for rd, md in zip(rt.values(), mt.values()):
            ri, mi = rd.interval(), md.interval()
            assert (float(np.ravel(ri[0])[0]), float(np.ravel(ri[1])[0])) == (float(np.ravel(mi[0])[0]), float(np.ravel(mi[1])[0]))
            assert tuple(rd.shape or ()) == md.shape


def _random_rows(comp, num, rng):
    """Rows of the transformed space that straddle every bound: inside, on the bounds, just outside, NaN."""
    rows = np.zeros((num, comp.d_out))
    for s in comp.slots:
        c0, w = s["out_col"], s["width"]
        if s["kind"] == 2:  # categorical: anything goes (soft one-hot vectors, out-of-range scalars)
            rows[:, c0:c0 + w] = rng.uniform(-0.5, 1.5, size=(num, w))
            continue
        lo, hi = max(s["low"], -60.0), min(s["high"], 60.0)
        v = rng.uniform(lo - 0.2 * (hi - lo), hi + 0.2 * (hi - lo), size=num)
        special = np.array([lo, hi, np.nextafter(hi, lo), np.nextafter(lo, lo - 1), lo + 0.5, hi - 0.5, hi + 0.5, np.floor(hi)])
        where = rng.uniform(size=num) < 0.25
        v[where] = rng.choice(special, size=int(where.sum()))
        inside = rng.uniform(size=num) < 0.8   # most entries inside, so that whole rows are accepted often enough
        v[inside] = rng.uniform(lo, hi, size=int(inside.sum()))
        v[rng.uniform(size=num) < 0.1] = rng.choice(special, size=1)[0]
        rows[:, c0] = v
    rows[rng.randint(num, size=3), rng.randint(comp.d_out, size=3)] = np.nan
    return rows


@needs_ref
def test_vectorised_contains_and_reverse_match_the_reference_point_by_point():
    """`compiler.contains_rows` / `reverse_rows` against the reference's own per-point path
    (`primary_algo.py:74-81`: `point in transformed_space`, `transformed_space.reverse(point)`)."""
    rs, ms = _ref_and_mirror()
    from orion.core.worker.transformer import build_required_space as ref_build
    with warnings.catch_warnings():
        warnings.simplefilter("ignore")
        rt = ref_build("real", rs)
        comp = compiler.compile_space(rt)
        rng = np.random.RandomState(11)
        rows = _random_rows(comp, 10000, rng)
        ok = compiler.contains_rows(comp, rows)
        assert 500 < ok.sum() < 9500                   # both outcomes are well represented
        points = [compiler.regroup_row(comp, row, transformed=True) for row in rows]
        want = np.array([bool(p in rt) for p in points])
        assert np.array_equal(ok, want), np.flatnonzero(ok != want)[:10]
        back = compiler.reverse_rows(comp, rows[ok][:2000])
        for p, mine in zip([q for q, o in zip(points, ok) if o][:2000], back):
            theirs = rt.reverse(p)
            assert mine in rs
            for a, b in zip(theirs, mine):
                assert np.array_equal(np.asarray(a, dtype=object), np.asarray(b, dtype=object)), (theirs, mine)
                if np.asarray(a).dtype.kind in "if":   # ints stay ints, floats stay floats (categories: 0-d object arrays there)
                    assert np.asarray(a).dtype.kind == np.asarray(b).dtype.kind
    # the exclusive upper bound (`space.py:356`): x == high is rejected, the float just below is accepted
    row = rows[ok][0].copy()
    s = [s for s in comp.slots if s["kind"] == 0][0]
    row[s["out_col"]] = s["high"]
    assert not compiler.contains_rows(comp, row[None])[0]
    row[s["out_col"]] = np.nextafter(s["high"], s["low"])
    assert compiler.contains_rows(comp, row[None])[0]
    with pytest.raises(ValueError):
        compiler.contains_rows(comp, rows[:, :-1])
