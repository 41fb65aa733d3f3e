"""The plugin under the UNMODIFIED reference: `PrimaryAlgo`, the Producer's deepcopy/set_state
protocol and a real `orion hunt` run (BASELINE config 1 shape, shortened) on PickledDB.

Every test runs in two flavours: `oracle` (CPU: the device is the oracle-backed stand-in `OracleEngine`; this is
what the build container, which has no GPU, can run) and `cuda` (marked gpu: the SAME unmodified reference --
pip-installed under `baseline/_ref` by `baseline/install_ref.sh`, which travels to the GPU box -- drives the real
CUDA engine through the real `orion.algo.base.BaseAlgorithm`, `PrimaryAlgo` and `orion hunt`).  The `cuda` flavour
also runs BASELINE config 1 as stated (2-D Rosenbrock, 50 trials, 10 000 candidates, default plugin mode) with a
per-suggest parity replay against the oracle on the same candidates."""
import copy
import os
import sys
import textwrap
import warnings

import numpy as np
import pytest

from oracle import ref_loader

pytestmark = pytest.mark.skipif(not ref_loader.available(),
                                reason="needs the reference: /root/reference or baseline/_ref (baseline/install_ref.sh)")


class _Counted:
    """The real engine with call counters on the three entry points the tests look at."""

    def __init__(self, eng, monkeypatch):
        self.eng = eng
        self.calls = {"sample": 0, "fit": 0, "score": 0}
        for name, key in (("sample", "sample"), ("fit", "fit"), ("score", "score"), ("score_pruned", "score"),
                          ("score_pruned_multi", "score")):
            monkeypatch.setattr(eng, name, self._wrap(getattr(eng, name), key))

    def _wrap(self, fn, key):
        def call(*a, **k):
            self.calls[key] += 1
            return fn(*a, **k)
        return call


@pytest.fixture(params=["oracle", pytest.param("cuda", marks=pytest.mark.gpu)])
def orion_env(request, monkeypatch):
    warnings.simplefilter("ignore")
    ref_loader.load()
    # numpy >= 2 returns a tuple from meshgrid; the 2019 reference concatenates it with a list
    # (`transformer.py:320-322`).  Restore the numpy 1.x return type for the duration of the test only.
    real_meshgrid = np.meshgrid
    monkeypatch.setattr(np, "meshgrid", lambda *a, **k: list(real_meshgrid(*a, **k)))
    import metaopt_b200.engine as engine_mod
    if request.param == "oracle":
        from fake_engine import OracleEngine
        eng = OracleEngine()
        monkeypatch.setattr(engine_mod, "get_engine", lambda device=None: eng)
    else:
        eng = _Counted(engine_mod.get_engine(0), monkeypatch)   # libob200.so on cuda:0, nothing replaced
    eng.flavour = request.param
    import importlib
    from metaopt_b200 import algo
    if not algo.HAVE_ORION:  # imported earlier in this process, before the reference was on sys.path
        importlib.reload(algo)
    algo.register()
    # the reference's storage/database are process-wide singletons: every test starts without them
    # (what the reference's own `tests/conftest.py:201-209` fixture does)
    from orion.core.io.database import Database
    from orion.core.io.database.pickleddb import PickledDB
    from orion.storage.base import Storage
    from orion.storage.legacy import Legacy
    for cls in (Storage, Legacy, Database, PickledDB):
        cls.instance = None
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
    # `base.py:87-119`: a `seed` kwarg calls seed_rng() during super().__init__, before the subclass body ends
    assert a._seed is not None and a._offset == 0 and a.configuration["bayesianoptimizer"]["seed"] == 1


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


@pytest.mark.parametrize("orion_env", [pytest.param("cuda", marks=pytest.mark.gpu)], indirect=True)
def test_cfg1_orion_hunt_50_trials_10k_candidates_on_the_device(orion_env, tmp_path, monkeypatch):
    """BASELINE config 1 as stated: `orion hunt` (unmodified reference: CLI -> experiment -> worker -> Producer ->
    PrimaryAlgo) on 2-D Rosenbrock, 50 trials, 10 000 candidates per suggest, default plugin mode for the GP
    (hyper-parameters fitted on the device), EI.  Every GP suggestion is recorded and replayed afterwards against
    the oracle (scikit-learn GP with the SAME fitted hyper-parameters + restated EI) on the same candidates."""
    from oracle import acquisition as oacq
    from oracle import gp as ogp
    from oracle import sampling as osamp
    from metaopt_b200 import algo as algo_mod
    from metaopt_b200 import compiler
    from orion.storage.base import Storage
    from orion.storage.legacy import Legacy
    if "legacy" not in Storage.typenames:
        Storage.types.append(Legacy)
        Storage.typenames.append("legacy")
    records = []
    real_suggest_gp = algo_mod.BayesianOptimizer._suggest_gp

    def recording(self, eng, comp):
        before = self._offset
        row = real_suggest_gp(self, eng, comp)
        records.append(dict(offset=before, seed=self._seed, X=np.array(self._X), y=np.array(self._y),
                            theta=copy.deepcopy(self._theta), row=np.array(row), comp=comp))
        return row
    monkeypatch.setattr(algo_mod.BayesianOptimizer, "_suggest_gp", recording)

    script = tmp_path / "rosen.py"
    script.write_text(textwrap.dedent("""\
        #!%s
        import argparse, json, os
        p = argparse.ArgumentParser(); p.add_argument("-x", type=float); p.add_argument("-y", type=float)
        a = p.parse_args()
        f = (1 - a.x) ** 2 + 100 * (a.y - a.x ** 2) ** 2
        json.dump([dict(name="rosen", type="objective", value=f)], open(os.environ["ORION_RESULTS_PATH"], "w"))
        """ % sys.executable))
    script.chmod(0o755)
    cfg = tmp_path / "cfg.yaml"
    cfg.write_text(textwrap.dedent("""\
        name: b200-rosen-cfg1
        max_trials: 50
        pool_size: 1
        algorithms:
          bayesianoptimizer:
            seed: 1
            n_initial_points: 10
            acq_func: EI
            n_points: 10000
            normalize_y: true
        database:
          type: pickleddb
          host: %s
        """ % (tmp_path / "db.pkl")))
    monkeypatch.chdir(tmp_path)
    import orion.core.cli
    orion.core.cli.main(["hunt", "--config", str(cfg), str(script), "-x~uniform(-2, 2)", "-y~uniform(-1, 3)"])
    import pickle
    db = pickle.load(open(tmp_path / "db.pkl", "rb"))
    done = [t for t in db.read("trials", {}) if t["status"] == "completed"]
    assert len(done) >= 50
    objectives = [r["value"] for t in done for r in t["results"] if r["type"] == "objective"]
    # random search at equal budget from the reference's own sampler
    from orion.algo.space import Real, Space
    space = Space()
    space.register(Real("/x", "uniform", -2, 4))   # scipy notation (loc, scale) == Orion prior uniform(-2, 2)
    space.register(Real("/y", "uniform", -1, 4))
    rnd = [(1 - x) ** 2 + 100 * (y - x ** 2) ** 2 for x, y in space.sample(50, seed=1)]
    assert min(objectives) < min(rnd)
    assert orion_env.calls["score"] >= 40 and len(records) >= 40
    # per-suggest parity replay
    for rec in records:
        comp = rec["comp"]
        C, _ = osamp.sample_transform(comp.slots, comp.cdf, rec["seed"], rec["offset"], 10000, True)
        amp, ls, noise = rec["theta"]
        model = ogp.SklearnGP(0, amp, np.broadcast_to(np.asarray(ls, float), (2,)).copy(), 1e-10 + noise, True).fit(rec["X"], rec["y"])
        mu, sd = model.predict(C)
        pick = oacq.pick(oacq.ACQ_EI, mu, sd, float(rec["y"].min()))
        np.testing.assert_allclose(rec["row"], comp.denormalise(C[pick]), rtol=1e-12, atol=1e-12)
