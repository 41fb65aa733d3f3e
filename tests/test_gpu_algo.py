"""The plugin end to end on the real device (mirror Space classes; the GPU box has no Orion):
suggest/observe loop, determinism, state, parity of each GP suggestion against the CPU oracle, the
BASELINE config-1 problem (2-D Rosenbrock, 50 trials, 10k candidates), and a single-rank NCCL run of
the multi-GPU helpers."""
import copy
import os

import numpy as np
import pytest
import torch

from metaopt_b200 import compiler
from metaopt_b200.algo import BayesianOptimizer
from metaopt_b200.space import Categorical, Integer, Real, Space, build_required_space
from oracle import acquisition as oacq
from oracle import gp as ogp
from oracle import sampling as osamp

pytestmark = pytest.mark.gpu


def rosen_space():
    s = Space()
    s.register(Real("/x", "uniform", -2, 4))
    s.register(Real("/y", "uniform", -1, 4))
    return s


def rosen(p):
    x, y = p
    return (1 - x) ** 2 + 100 * (y - x ** 2) ** 2


def test_cfg1_rosenbrock_50_trials_with_per_suggest_parity(engine):
    space = rosen_space()
    t = build_required_space("real", space)
    a = BayesianOptimizer(t, seed=1, n_initial_points=10, acq_func="EI", n_points=10000, normalize_y=True,
                          length_scale=0.3)
    comp = compiler.compile_space(t)
    ys = []
    for trial in range(50):
        offset_before = a._offset
        (tp,) = a.suggest(1)
        assert tp in t
        if trial >= 10:
            # oracle replay of this suggestion on the SAME candidates (Philox restated in numpy)
            C, _ = osamp.sample_transform(comp.slots, comp.cdf, a._seed, offset_before, 10000, True)
            amp, ls, noise = a._theta  # hyper-parameters the plugin fitted for this suggest (device L-BFGS-B)
            model = ogp.SklearnGP(0, amp, ls, 1e-10 + noise, True).fit(np.asarray(a._X), np.asarray(a._y))
            mu, sd = model.predict(C)
            pick = oacq.pick(oacq.ACQ_EI, mu, sd, float(min(a._y)))
            want = comp.denormalise(C[pick])
            np.testing.assert_allclose(np.asarray(tp, dtype=float), want, rtol=1e-12, atol=1e-12)
        y = rosen(t.reverse(tp))
        ys.append(y)
        a.observe([tp], [{"objective": y, "constraint": [], "gradient": None}])
    # random search with the same budget from the same sampler
    rnd = [rosen(p) for p in space.sample(50, seed=1)]
    assert min(ys) < min(rnd)
    assert min(ys[10:]) < min(ys[:10])


def test_fixed_hyperparameters_mode_matches_oracle(engine):
    space = rosen_space()
    t = build_required_space("real", space)
    a = BayesianOptimizer(t, seed=4, n_initial_points=8, acq_func="LCB", n_points=20000, normalize_y=True, length_scale=0.3,
                          fit_hyperparameters=False)
    comp = compiler.compile_space(t)
    for trial in range(14):
        off = a._offset
        (tp,) = a.suggest(1)
        if trial >= 8:
            C, _ = osamp.sample_transform(comp.slots, comp.cdf, a._seed, off, 20000, True)
            model = ogp.SklearnGP(0, 1.0, 0.3 * np.ones(2), 1e-10 + 1e-4, True).fit(np.asarray(a._X), np.asarray(a._y))
            mu, sd = model.predict(C)
            want = comp.denormalise(C[oacq.pick(oacq.ACQ_LCB, mu, sd, float(min(a._y)))])
            np.testing.assert_allclose(np.asarray(tp, dtype=float), want, rtol=1e-12, atol=1e-12)
        a.observe([tp], [{"objective": rosen(t.reverse(tp))}])


def test_mixed_space_suggestions_are_valid_and_reproducible(engine):
    s = Space()
    s.register(Real("lr", "reciprocal", 1e-5, 1.0))
    s.register(Integer("layers", "uniform", 1, 8))
    s.register(Categorical("act", ["relu", "tanh", "gelu", "silu"]))
    s.register(Categorical("bn", [True, False]))
    s.register(Real("w", "uniform", 0, 1, shape=3))
    t = build_required_space("real", s)

    def f(p):
        act, bn, layers, lr, w = p
        return (np.log10(lr) + 3) ** 2 + 0.1 * (layers - 4) ** 2 + (0.2 if bn else 0.0) + float(np.sum((w - 0.5) ** 2))

    def run(seed):
        a = BayesianOptimizer(t, seed=seed, n_initial_points=6, acq_func="gp_hedge", n_points=4096, normalize_y=True)
        pts = []
        for _ in range(12):
            naive = copy.deepcopy(a)            # producer.py:161
            (tp,) = naive.suggest(1)
            a.set_state(naive.state_dict)       # producer.py:84
            assert tp in t
            pts.append(tp)
            a.observe([tp], [{"objective": f(t.reverse(tp))}])
        return pts
    p1, p2, p3 = run(5), run(5), run(6)
    assert repr(p1) == repr(p2) and repr(p1) != repr(p3)
    assert len({repr(p) for p in p1}) == 12


def test_space_sample_on_device(engine):
    s = Space()
    s.register(Integer("i", "uniform", -3, 6))
    s.register(Categorical("c", {"asdfa": 0.1, 2: 0.2, 3: 0.3, 4: 0.4}))
    s.register(Real("r", "norm", 0.9, 0.5, low=-1, high=3))
    pts = s.sample(2000, seed=5)
    assert len(pts) == 2000 and all(p in s for p in pts)
    assert s.sample(10, seed=5) == pts[:10] and s.sample(10, seed=6) != pts[:10]
    freq = np.mean([p[0] == 4 for p in pts])
    assert abs(freq - 0.4) < 0.05
    assert isinstance(pts[0][1], int) or isinstance(pts[0][1], np.integer)
    one = Integer("k", "uniform", 0, 5).sample(3, seed=1)
    assert len(one) == 3 and all(0 <= v < 5 for v in one)


def test_multi_gpu_helpers_on_one_rank(engine):
    import torch.distributed as dist
    from metaopt_b200 import dist as obdist
    os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
    os.environ.setdefault("MASTER_PORT", "29617")
    dist.init_process_group("nccl", rank=0, world_size=1, device_id=torch.device("cuda", 0))
    try:
        t = build_required_space("real", rosen_space())
        comp = compiler.compile_space(t)
        compiler.upload(engine, comp)
        X = engine.sample(3, 0, 64).cpu().numpy()
        y = ((X - 0.3) ** 2).sum(axis=1)
        engine.fit(X, y, alpha=1e-6, normalize_y=True)
        assert obdist.broadcast_fit(engine, 64, 2, src=0) == engine.state_bytes(64, 2)
        whole = engine.score(None, M=20000, seed=9, first_index=100, acq_mask=7, y_opt=float(y.min()))
        merged = obdist.sharded_score(engine, 20000, seed=9, first_index=100, acq_mask=7, y_opt=float(y.min()))
        assert merged == whole
        # emulate 4 shards on this one GPU and merge: same winners
        parts = []
        for r in range(4):
            s0, c = obdist.shard_range(20000, r, 4)
            parts.append(engine.score(None, M=c, seed=9, first_index=100 + s0, acq_mask=7, y_opt=float(y.min())))
        assert obdist.merge_best(parts) == whole
        assert obdist.max_over_ranks(3.5, device=engine.tdev) == 3.5
    finally:
        dist.destroy_process_group()
