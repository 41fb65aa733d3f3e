"""Single-process multi-GPU engine (needs >= 2 GPUs; skipped on a one-GPU box): the plugin shards the candidates
of one suggest over several devices and returns exactly what one device returns."""
import numpy as np
import pytest
import torch

from metaopt_b200 import compiler
from metaopt_b200.algo import BayesianOptimizer
from metaopt_b200.space import build_required_space

from helpers import cfg3_space, transformed

pytestmark = pytest.mark.gpu


def _need(n):
    if torch.cuda.device_count() < n:
        pytest.skip("needs %d GPUs" % n)


def test_multi_engine_matches_single_engine(engine):
    _need(2)
    from metaopt_b200.multi import get_multi_engine
    multi = get_multi_engine(list(range(min(torch.cuda.device_count(), 4))))
    comp = compiler.compile_space(transformed(cfg3_space()))
    compiler.upload(engine, comp)
    compiler.upload(multi, comp)
    X = engine.sample(9, 0, 400).cpu().numpy()
    y = ((X - 0.5) ** 2).sum(axis=1)
    engine.fit(X, y, alpha=1e-6, normalize_y=True)
    multi.fit(X, y, alpha=1e-6, normalize_y=True)
    y_opt = float(y.min())
    whole = engine.score(None, M=300_001, seed=4, first_index=55, acq_mask=7, y_opt=y_opt)
    assert multi.score(None, M=300_001, seed=4, first_index=55, acq_mask=7, y_opt=y_opt) == whole
    for name in ("EI", "LCB"):
        a, _ = engine.score_pruned(None, M=300_001, seed=4, first_index=55, acq=name, y_opt=y_opt)
        b, n_scored = multi.score_pruned(None, M=300_001, seed=4, first_index=55, acq=name, y_opt=y_opt)
        assert a == b and a[name] == whole[name] and 0 < n_scored <= 300_001
    # a refit on the first device invalidates the replicas
    multi.fit(X, 2.0 * y, alpha=1e-6, normalize_y=True)
    engine.fit(X, 2.0 * y, alpha=1e-6, normalize_y=True)
    assert multi.score(None, M=100_000, seed=4, acq_mask=1, y_opt=2 * y_opt) == engine.score(None, M=100_000, seed=4, acq_mask=1,
                                                                                            y_opt=2 * y_opt)


def test_plugin_with_a_device_list(engine):
    _need(2)
    t = build_required_space("real", cfg3_space())

    def run(device):
        a = BayesianOptimizer(t, seed=3, n_initial_points=6, acq_func="EI", n_points=200_000, normalize_y=True, device=device)
        pts = []
        for _ in range(9):
            (tp,) = a.suggest(1)
            assert tp in t
            pts.append(tp)
            flat = np.asarray(compiler.flatten_point(a._get_compiled(), tp))
            a.observe([tp], [{"objective": float(((a._get_compiled().normalise(flat) - 0.4) ** 2).sum())}])
        return pts
    assert repr(run([0, 1])) == repr(run(0))
