"""CUDA path against the committed golden vectors: the reference's own samples (replayed from the
variates it consumed) and scikit-learn's fit/predict outputs."""
import json
import os

import numpy as np
import pytest
import torch

from metaopt_b200 import compiler

from helpers import space_from_spec, transformed

pytestmark = pytest.mark.gpu
GOLD = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden")


def _cases():
    meta = json.load(open(os.path.join(GOLD, "sampling_ref.json")))
    arrays = np.load(os.path.join(GOLD, "sampling_ref.npz"))
    return [(c, arrays[c["name"] + "_variates"], arrays[c["name"] + "_flat"]) for c in meta["cases"]]


@pytest.mark.parametrize("case,variates,flat", _cases(), ids=lambda x: x["name"] if isinstance(x, dict) else "")
def test_device_reproduces_reference_samples(engine, case, variates, flat):
    comp = compiler.compile_space(transformed(space_from_spec(case["spec"])))
    compiler.upload(engine, comp)
    got = engine.sample(0, 0, case["n"], normalise=False, variates=variates).cpu().numpy()
    ulp = [s["out_col"] for s in comp.slots if s["prior"] == 1]  # loguniform: device exp vs glibc exp
    exact = [c for c in range(comp.d_out) if c not in ulp]
    assert np.array_equal(got[:, exact], flat[:, exact])
    if ulp:
        np.testing.assert_allclose(got[:, ulp], flat[:, ulp], rtol=4e-16, atol=0)
    if comp.normalisable:
        norm = engine.sample(0, 0, case["n"], normalise=True, variates=variates).cpu().numpy()
        want = comp.normalise(flat)
        assert np.array_equal(norm[:, exact], want[:, exact])


@pytest.mark.parametrize("tag", ["m52", "rbf"])
def test_device_matches_sklearn_goldens(engine, tag):
    g = np.load(os.path.join(GOLD, "gp_sklearn.npz"))
    n, d, kind, norm = (int(v) for v in g[tag + "_meta"])
    X, y, C, ls = g[tag + "_X"], g[tag + "_y"], g[tag + "_C"], g[tag + "_ls"]
    engine.fit(X, y, kind=kind, amplitude=1.4, length_scale=ls, alpha=1e-10 + 1e-4, normalize_y=bool(norm))
    L, al = engine.get_factor()
    np.testing.assert_allclose(L, g[tag + "_L"], rtol=1e-9, atol=1e-12)
    np.testing.assert_allclose(al, g[tag + "_alpha"], rtol=1e-7, atol=1e-7 * np.abs(g[tag + "_alpha"]).max())
    np.testing.assert_allclose(engine.lml(), float(g[tag + "_lml"]), rtol=1e-9)
    out = engine.score(torch.as_tensor(C).cuda(), acq_mask=7, y_opt=float(y.min()), dense=True)
    np.testing.assert_allclose(out["mu"].cpu().numpy(), g[tag + "_mu"], rtol=1e-6)
    np.testing.assert_allclose(out["sigma"].cpu().numpy(), g[tag + "_sd"], rtol=1e-6, atol=1e-9)
    acq = out["acq"].cpu().numpy()
    np.testing.assert_allclose(acq[0], g[tag + "_ei"], rtol=1e-6, atol=1e-12)
    np.testing.assert_allclose(acq[1], g[tag + "_pi"], rtol=1e-6, atol=1e-12)
    np.testing.assert_allclose(-acq[2], g[tag + "_lcb"], rtol=1e-6)
    assert [out["EI"]["index"], out["PI"]["index"], out["LCB"]["index"]] == g[tag + "_pick"].tolist()
