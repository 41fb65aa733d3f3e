"""Generate tests/golden/* from the UNMODIFIED reference (run in the build container only).

    python -m oracle.make_golden

* `sampling_ref.{json,npz}`: for several spaces built with the reference's own classes, the points
  returned by the reference's `Space.sample(n, seed)` (`src/orion/algo/space.py:745-773`), flattened
  through the reference's `TransformedSpace.transform` (`src/orion/core/worker/transformer.py:473`)
  -- except the one-hot step for k>2, which raises under numpy 2 (`transformer.py:319-323`) and is
  taken from the restatement `oracle.sampling.one_hot` (pinned separately by the reference's
  known-answer vectors in tests/test_oracle_golden.py) -- together with the uniform/normal variates
  the reference consumed, recovered by replaying the same `RandomState(seed)` stream in the
  reference's dimension-major order (`tests/unittests/algo/test_space.py:602-635`).
* `gp_sklearn.npz`: scikit-learn 1.9.0 `GaussianProcessRegressor` fit/predict outputs (the real
  third-party code behind the reference's plugin) + restated skopt acquisitions on fixed inputs.
"""
import json
import os
import warnings

import numpy as np

from . import acquisition as oacq
from . import gp as ogp
from . import ref_loader, sampling

HERE = os.path.dirname(os.path.abspath(__file__))
GOLD = os.path.join(os.path.dirname(HERE), "tests", "golden")

# (name, [dimension specs]) -- spec: (kind, name, args...) mirrored 1:1 by tests/helpers.space_from_spec
CASES = [
    ("reals20", [("real", "x%02d" % i, "uniform", 0.0, 1.0) for i in range(20)]),
    ("mixed", [("real", "a", "uniform", -3.0, 9.0), ("real", "b", "reciprocal", 1e-5, 1.0),
               ("integer", "c", "uniform", -3.0, 6.0), ("real", "d", "norm", 0.5, 2.0),
               ("categorical", "e", ["p", "q", "r", "s"], [0.1, 0.2, 0.3, 0.4]),
               ("categorical", "f", ["x", "y"], None), ("categorical", "g", ["only"], None),
               ("integer", "h", "uniform", 0.0, 100.0), ("categorical", "i", ["u", "v", "w"], None)]),
    ("cfg3_small", [("integer", "i%d" % i, "uniform", 0.0, 100.0) for i in range(3)] +
     [("categorical", "c%d" % i, ["a", "b", "c", "d"], None) for i in range(3)] +
     [("real", "r%d" % i, "uniform", 0.0, 1.0) for i in range(3)]),
]


def build_reference_space(spec):
    from orion.algo.space import Categorical, Integer, Real, Space
    space = Space()
    for item in spec:
        if item[0] == "categorical":
            _, name, cats, probs = item
            space.register(Categorical(name, dict(zip(cats, probs)) if probs else cats))
        else:
            kind, name, prior, a, b = item
            cls = Real if kind == "real" else Integer
            space.register(cls(name, prior, a, b))
    return space


def replay_variates(space, n, seed):
    """The U/Z draws `Space.sample(n, seed)` consumes, [n x n_dims], dimension-major order."""
    rng = np.random.RandomState(seed)
    cols = []
    for dim in space.values():
        if dim.type == "categorical":
            col = [rng.random_sample() for _ in range(n)]
        elif dim._prior_name == "norm":
            col = [rng.standard_normal() for _ in range(n)]
        else:
            col = [rng.random_sample() for _ in range(n)]
        cols.append(col)
    return np.asarray(cols).T.copy()


def flatten_reference_points(space, points):
    from orion.core.worker.transformer import build_required_space
    tspace = build_required_space("real", space)
    rows = []
    for point in points:
        flat = []
        for value, dim, tdim in zip(point, space.values(), tspace.values()):
            if dim.type == "categorical" and len(dim.categories) > 2:
                idx = dim.categories.index(value)
                flat.extend(sampling.one_hot(np.asarray(idx), len(dim.categories)).tolist())
            else:
                flat.append(float(tdim.transform(value)))
        rows.append(flat)
    return np.asarray(rows, dtype=np.float64), [tuple(map(float, np.ravel(iv))) if not isinstance(iv, tuple) or True
                                                else iv for iv in tspace.interval()]


def main():
    assert ref_loader.load() is not None, "needs /root/reference"
    os.makedirs(GOLD, exist_ok=True)
    meta, arrays = {"cases": []}, {}
    for name, spec in CASES:
        space = build_reference_space(spec)
        n, seed = 64, 5
        with warnings.catch_warnings():
            warnings.simplefilter("ignore")
            points = space.sample(n, seed=seed)
            flat, intervals = flatten_reference_points(space, points)
        variates = replay_variates(space, n, seed)
        arrays[name + "_variates"] = variates
        arrays[name + "_flat"] = flat
        # dims in the reference's sorted order, so fixtures are self-describing
        meta["cases"].append({"name": name, "spec": spec, "order": list(space.keys()), "n": n, "seed": seed,
                              "intervals": [[float(np.ravel(lo)[0]), float(np.ravel(hi)[0])] for lo, hi in intervals]})
    np.savez_compressed(os.path.join(GOLD, "sampling_ref.npz"), **arrays)
    with open(os.path.join(GOLD, "sampling_ref.json"), "w") as f:
        json.dump(meta, f, indent=1)

    # ---- GP goldens ------------------------------------------------------------------------------
    rs = np.random.RandomState(0)
    out = {}
    for tag, (n, d, kind, norm) in {"m52": (60, 4, ogp.KERNEL_MATERN52, True), "rbf": (45, 3, ogp.KERNEL_RBF, False)}.items():
        X = rs.uniform(size=(n, d))
        y = ((X - 0.5) ** 2).sum(axis=1) + 0.01 * rs.randn(n)
        C = rs.uniform(size=(128, d))
        ls = np.linspace(0.6, 1.5, d)
        model = ogp.SklearnGP(kind, 1.4, ls, 1e-10 + 1e-4, norm).fit(X, y)
        mu, sd = model.predict(C)
        y_opt = float(y.min())
        out.update({tag + "_X": X, tag + "_y": y, tag + "_C": C, tag + "_ls": ls, tag + "_L": model.L,
                    tag + "_alpha": model.alpha_, tag + "_mu": mu, tag + "_sd": sd,
                    tag + "_lml": np.float64(model.gpr.log_marginal_likelihood_value_),
                    tag + "_ei": oacq.gaussian_ei(mu, sd, y_opt), tag + "_pi": oacq.gaussian_pi(mu, sd, y_opt),
                    tag + "_lcb": oacq.gaussian_lcb(mu, sd),
                    tag + "_pick": np.asarray([oacq.pick(k, mu, sd, y_opt) for k in range(3)]),
                    tag + "_meta": np.asarray([n, d, kind, int(norm)])})
    np.savez_compressed(os.path.join(GOLD, "gp_sklearn.npz"), **out)
    print("wrote", os.listdir(GOLD))


if __name__ == "__main__":
    main()
