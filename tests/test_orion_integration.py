"""The plugin under the UNMODIFIED reference: `PrimaryAlgo`, the Producer's deepcopy/set_state
protocol and a real `orion hunt` run (BASELINE config 1 shape, shortened) on PickledDB.

Runs only where /root/reference exists (the build container, which has no GPU), so the device is the
oracle-backed stand-in `OracleEngine`; the same plugin code runs against the real CUDA engine in
tests/test_gpu_algo.py on the GPU box."""
import copy
import os
import sys
import textwrap
import warnings

import numpy as np
import pytest

from oracle import ref_loader

pytestmark = pytest.mark.skipif(not ref_loader.available(), reason="/root/reference is only in the build container")


@pytest.fixture()
def orion_env(monkeypatch):
    warnings.simplefilter("ignore")
    ref_loader.load()
    # numpy >= 2 returns a tuple from meshgrid; the 2019 reference concatenates it with a list
    # (`transformer.py:320-322`).  Restore the numpy 1.x return type for the duration of the test only.
    real_meshgrid = np.meshgrid
    monkeypatch.setattr(np, "meshgrid", lambda *a, **k: list(real_meshgrid(*a, **k)))
    import metaopt_b200.engine as engine_mod
    from fake_engine import OracleEngine
    eng = OracleEngine()
    monkeypatch.setattr(engine_mod, "get_engine", lambda device=None: eng)
    import importlib
    from metaopt_b200 import algo
    if not algo.HAVE_ORION:  # imported earlier in this process, before the reference was on sys.path
        importlib.reload(algo)
    algo.register()
    return eng


def _space():
    from orion.algo.space import Categorical, Integer, Real, Space
    s = Space()
    s.register(Real("/lr", "reciprocal", 1e-4, 1.0))
    s.register(Integer("/layers", "uniform", 1, 8))
    s.register(Categorical("/act", ["relu", "tanh", "gelu"]))
    s.register(Real("/w", "uniform", -1, 2, shape=2))
    return s


def _objective(point):
    act, layers, lr, w = point
    act = str(np.asarray(act).item())  # the reference's Enumerate.reverse returns a 0-d object array
    return float((np.log10(lr) + 2) ** 2 + 0.05 * (layers - 3) ** 2 + {"relu": 0, "tanh": 0.2, "gelu": 0.1}[act] +
                 np.sum((np.asarray(w) - 0.3) ** 2))


def test_plugin_is_a_real_basealgorithm_and_is_discoverable(orion_env):
    from orion.algo.base import BaseAlgorithm, OptimizationAlgorithm
    from metaopt_b200.algo import BayesianOptimizer, HAVE_ORION
    assert HAVE_ORION and issubclass(BayesianOptimizer, BaseAlgorithm)
    a = OptimizationAlgorithm("BayesianOptimizer", _space(), seed=1, n_points=50)
    assert isinstance(a, BayesianOptimizer) and a.n_points == 50
    assert a.configuration["bayesianoptimizer"]["acq_func"] == "gp_hedge"


def test_primary_algo_drives_the_plugin(orion_env):
    from orion.core.worker.primary_algo import PrimaryAlgo
    space = _space()
    palgo = PrimaryAlgo(space, {"bayesianoptimizer": {"seed": 3, "n_initial_points": 4, "acq_func": "EI",
                                                      "n_points": 200, "normalize_y": True,
                                                      "fit_hyperparameters": False}})
    assert palgo.algorithm.space is palgo.transformed_space
    assert palgo.configuration["bayesianoptimizer"]["n_points"] == 200
    seen = []
    for _ in range(7):
        points = palgo.suggest(1)       # PrimaryAlgo checks `point in transformed_space` and reverses
        assert len(points) == 1 and points[0] in space
        seen.append(points[0])
        palgo.observe(points, [{"objective": _objective(points[0]), "constraint": [], "gradient": None}])
    assert orion_env.calls["fit"] == 3 and orion_env.calls["score"] == 3
    assert len({repr(p) for p in seen}) == 7


def test_producer_protocol_deepcopy_lies_and_state_sync(orion_env):
    """`producer.py:159-174` (naive copy observes lies) and `:84` (state copied back)."""
    from orion.core.worker.primary_algo import PrimaryAlgo
    space = _space()
    algo = PrimaryAlgo(space, {"bayesianoptimizer": {"seed": 9, "n_initial_points": 2, "acq_func": "LCB", "n_points": 100}})
    history = []
    for step in range(5):
        naive = copy.deepcopy(algo)
        lie_point = space.sample(1, seed=step)[0]
        naive.observe([lie_point], [{"objective": 10.0, "constraint": [], "gradient": None}])
        points = naive.suggest(1)
        algo.set_state(naive.state_dict)
        assert points[0] in space
        history.append(points[0])
        algo.observe(points, [{"objective": _objective(points[0]), "constraint": [], "gradient": None}])
        assert len(algo.algorithm._y) == step + 1 and len(naive.algorithm._y) == step + 1 if step == 0 else True
    # the stream advanced through the copies: no suggestion repeats
    assert len({repr(p) for p in history}) == 5


def test_orion_hunt_end_to_end(orion_env, tmp_path, monkeypatch):
    """BASELINE config 1 (2-D Rosenbrock via `orion hunt`), shortened: 9 trials, 400 candidates."""
    from orion.storage.base import Storage
    from orion.storage.legacy import Legacy
    if "legacy" not in Storage.typenames:
        Storage.types.append(Legacy)
        Storage.typenames.append("legacy")
    script = tmp_path / "rosen.py"
    script.write_text(textwrap.dedent("""\
        #!%s
        # writes what orion.client.report_results would (client/__init__.py:25-48), without importing orion
        import argparse, json, os
        p = argparse.ArgumentParser(); p.add_argument("-x", type=float); p.add_argument("-y", type=float)
        a = p.parse_args()
        f = (1 - a.x) ** 2 + 100 * (a.y - a.x ** 2) ** 2
        json.dump([dict(name="rosen", type="objective", value=f)], open(os.environ["ORION_RESULTS_PATH"], "w"))
        """ % sys.executable))
    script.chmod(0o755)
    cfg = tmp_path / "cfg.yaml"
    cfg.write_text(textwrap.dedent("""\
        name: b200-rosen
        max_trials: 9
        pool_size: 1
        algorithms:
          bayesianoptimizer:
            seed: 1
            n_initial_points: 4
            acq_func: EI
            n_points: 400
            normalize_y: true
        database:
          type: pickleddb
          host: %s
        """ % (tmp_path / "db.pkl")))
    monkeypatch.chdir(tmp_path)
    import orion.core.cli
    orion.core.cli.main(["hunt", "--config", str(cfg), str(script), "-x~uniform(-2, 4)", "-y~uniform(-1, 4)"])
    import pickle
    db = pickle.load(open(tmp_path / "db.pkl", "rb"))
    trials = db.read("trials", {})
    done = [t for t in trials if t["status"] == "completed"]
    assert len(done) >= 9
    exp = db.read("experiments", {})[0]
    assert "bayesianoptimizer" in exp["algorithms"] and exp["algorithms"]["bayesianoptimizer"]["n_points"] == 400
    assert orion_env.calls["score"] >= 4  # GP-driven suggestions after the 4 initial points
    for t in done:
        x = [p["value"] for p in t["params"] if p["name"] == "/x"][0]
        y = [p["value"] for p in t["params"] if p["name"] == "/y"][0]
        assert -2 <= x < 4 and -1 <= y < 4  # Orion prior strings are uniform(low, high)
