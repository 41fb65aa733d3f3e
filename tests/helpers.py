"""Shared builders for the tests: example spaces (mirror classes) and synthetic GP problems."""
import numpy as np

from metaopt_b200.space import Categorical, Integer, Real, Space, build_required_space


def mixed_space():
    s = Space()
    s.register(Real("a_real", "uniform", -3, 9))
    s.register(Real("b_log", "reciprocal", 1e-5, 1.0))
    s.register(Integer("c_int", "uniform", -3, 6))
    s.register(Real("d_norm", "norm", 0.5, 2.0, low=-4.0, high=5.0))
    s.register(Categorical("e_cat4", {"a": 0.1, "b": 0.2, "c": 0.3, "d": 0.4}))
    s.register(Categorical("f_cat2", ["x", "y"]))
    s.register(Categorical("g_cat1", ["only"]))
    s.register(Real("h_vec", "uniform", 0, 1, shape=3))
    s.register(Integer("i_intnorm", "norm", 10, 3, low=0, high=20))
    s.register(Categorical("j_cat3", ["u", "v", "w"]))
    return s


def cfg3_space():
    """BASELINE config 3: 10 Integer uniform(0,100), 10 Categorical k=4, 10 Real uniform(0,1) -> D=60."""
    s = Space()
    for i in range(10):
        s.register(Integer("i%02d" % i, "uniform", 0, 100))
        s.register(Categorical("c%02d" % i, ["a", "b", "c", "d"]))
        s.register(Real("r%02d" % i, "uniform", 0, 1))
    return s


def cfg5_space():
    """BASELINE config 5: 9 uniform + 8 loguniform reals, 17 integers, 16 categoricals k=4 -> D_out=98."""
    s = Space()
    for i in range(9):
        s.register(Real("ru%02d" % i, "uniform", -1.0 - i, 2.0 + i))
    for i in range(8):
        s.register(Real("rl%02d" % i, "reciprocal", 1e-5, 1.0))
    for i in range(17):
        s.register(Integer("in%02d" % i, "uniform", 0, 10 + i))
    for i in range(16):
        probs = [0.25] * 4 if i % 2 == 0 else [0.1, 0.2, 0.3, 0.4]
        s.register(Categorical("ca%02d" % i, dict(zip("abcd", probs))))
    return s


def real_space(d):
    s = Space()
    for i in range(d):
        s.register(Real("x%03d" % i, "uniform", 0, 1))
    return s


def transformed(space):
    return build_required_space("real", space)


def gp_problem(n, d, seed=0, noise=0.01):
    """SURVEY 8(d) recipe: X ~ U[0,1]^d, y = sum (x-0.5)^2 + noise * randn."""
    rs = np.random.RandomState(seed)
    X = rs.uniform(size=(n, d))
    y = ((X - 0.5) ** 2).sum(axis=1) + noise * rs.randn(n)
    return X, y


def space_from_spec(spec):
    """Rebuild a golden-fixture space (tests/golden/sampling_ref.json) with the mirror classes."""
    s = Space()
    for item in spec:
        if item[0] == "categorical":
            _, name, cats, probs = item
            s.register(Categorical(name, dict(zip(cats, probs)) if probs else cats))
        else:
            kind, name, prior, a, b = item
            s.register((Real if kind == "real" else Integer)(name, prior, a, b))
    return s
