"""SpaceCompiler: turn a search space into the flat slot table of `ob200_space_set`.

Accepts the mirror classes of `metaopt_b200.space` or the reference's own objects
(`orion.algo.space.{Real,Integer,Categorical,Space}`, `orion.core.worker.transformer.
{TransformedDimension,TransformedSpace}`) by duck typing.  Attributes read per dimension:
`type`, `name`, `shape`, `interval()`, and on the original dimension `_prior_name`, `_args`,
`_kwargs`, `categories`, `_probs` (`/root/reference/src/orion/algo/space.py:111-127,515-540`).

One *slot* = one scalar source variate; a dimension of shape (m,) gives m slots.  Flat column
order = dimensions in sorted-name order (`space.py:848-858`), each flattened like
`orion.core.utils.points.flatten_dims` (`points.py:49-74`, 1-D shapes only).
"""
import math
import os

import numpy as np

from . import _lib

_PRIOR_ALIASES = {"uniform": "uniform", "reciprocal": "reciprocal", "loguniform": "reciprocal",
                  "norm": "norm", "normal": "norm", "gaussian": "norm"}


class CompiledSpace:
    """slots/cdf for the device + the layout needed to regroup flat rows into Orion points."""

    def __init__(self, slots, cdf, layout, d_out, normalisable):
        self.slots, self.cdf, self.layout = slots, cdf, layout
        self.d_out, self.normalisable = d_out, normalisable
        self.n_slots = len(slots)
        self.key = (tuple(tuple(sorted(s.items())) for s in slots), tuple(cdf))
        self.nlo = np.zeros(d_out)
        self.nhi = np.ones(d_out)
        for s in slots:
            self.nlo[s["out_col"]:s["out_col"] + s["width"]] = s["nlo"]
            self.nhi[s["out_col"]:s["out_col"] + s["width"]] = s["nhi"]

    def normalise(self, flat):
        """(x - low) / (high - low), two roundings (skopt Normalize)."""
        return (np.asarray(flat, dtype=float) - self.nlo) / (self.nhi - self.nlo)

    def denormalise(self, flat):
        return np.asarray(flat, dtype=float) * (self.nhi - self.nlo) + self.nlo


def _prior_params(orig):
    name = _PRIOR_ALIASES.get(getattr(orig, "_prior_name", None))
    if name is None:
        raise NotImplementedError("Dimension %r: prior %r has no device implementation"
                                  % (orig.name, getattr(orig, "_prior_name", None)))
    args = list(getattr(orig, "_args", ()))
    kwargs = dict(getattr(orig, "_kwargs", {}) or {})
    if name in ("uniform", "norm"):
        loc = kwargs.get("loc", args[0] if len(args) > 0 else 0.0)
        scale = kwargs.get("scale", args[1] if len(args) > 1 else 1.0)
        return (_lib.PRIOR_UNIFORM if name == "uniform" else _lib.PRIOR_NORMAL), float(loc), float(scale)
    a = kwargs.get("a", args[0] if len(args) > 0 else None)
    b = kwargs.get("b", args[1] if len(args) > 1 else None)
    if a is None or b is None or not (0 < a < b):
        raise ValueError("Dimension %r: reciprocal/loguniform needs 0 < a < b" % orig.name)
    if kwargs.get("loc", 0.0) != 0.0 or kwargs.get("scale", 1.0) != 1.0:
        raise NotImplementedError("Dimension %r: loc/scale on reciprocal is not supported" % orig.name)
    # scipy reciprocal._ppf: exp(log(a) + q*(log(b) - log(a)))
    return _lib.PRIOR_LOGUNIFORM, float(np.log(a)), float(np.log(b) - np.log(a))


def compile_space(space):
    """Return a `CompiledSpace` for a (transformed or original) space."""
    slots, cdf, layout = [], [], []
    col = 0
    normalisable = True
    for dim in space.values():
        orig = getattr(dim, "original_dimension", dim)
        kind = orig.type
        shape = tuple(orig.shape or ())
        if len(shape) > 1:
            raise NotImplementedError("Dimension %r: only scalar or 1-D shapes are supported (shape=%r)"
                                      % (orig.name, shape))
        m = shape[0] if shape else 1
        first_col = col
        if kind == "categorical":
            k = len(orig.categories)
            probs = np.asarray(orig._probs, dtype=float)
            c = probs.cumsum()
            c = c / c[-1]  # numpy rng.choice normalises the cdf the same way
            width = k if k > 2 else 1
            if shape and k > 2:
                raise NotImplementedError("Dimension %r: shaped categoricals with more than 2 categories give a "
                                          "2-D transformed point, which Orion's flatten_dims rejects too" % orig.name)
            for _ in range(m):
                slots.append(dict(kind=_lib.KIND_CAT, prior=_lib.PRIOR_CHOICE, out_col=col, width=width, ncat=k,
                                  cdf_off=len(cdf), p0=0.0, p1=0.0, low=0.0, high=float(k), nlo=-0.1, nhi=1.1))
                col += width
            cdf.extend(c.tolist())
            layout.append(dict(name=orig.name, kind="categorical", shape=shape, col=first_col, m=m, k=k,
                               categories=tuple(orig.categories)))
            continue
        if kind not in ("real", "integer"):
            raise TypeError("Unsupported dimension type ('{}') or requirement ('real')".format(kind))
        prior, p0, p1 = _prior_params(orig)
        low, high = orig.interval()
        low, high = float(low), float(high)
        if not high > low:
            raise ValueError("Lower bound {} has to be less than upper bound {}".format(low, high))
        finite = math.isfinite(low) and math.isfinite(high)
        normalisable &= finite
        nlo, nhi = (low, high) if finite else (0.0, 1.0)
        for _ in range(m):
            slots.append(dict(kind=_lib.KIND_INT if kind == "integer" else _lib.KIND_REAL, prior=prior, out_col=col,
                              width=1, ncat=0, cdf_off=0, p0=p0, p1=p1, low=low, high=high, nlo=nlo, nhi=nhi))
            col += 1
        layout.append(dict(name=orig.name, kind=kind, shape=shape, col=first_col, m=m, k=0))
    if not slots:
        raise ValueError("cannot compile an empty space")
    return CompiledSpace(slots, cdf, layout, col, normalisable)


def regroup_row(compiled, row, transformed=True):
    """Flat (un-normalised) row -> tuple with one entry per dimension.

    transformed=True gives a point of the `requires='real'` space (floats, one-hot 1-D arrays),
    i.e. what `BaseAlgorithm.suggest` must return; transformed=False gives a point of the
    original space (ints, category objects), i.e. what `Space.sample` returns."""
    out = []
    for item in compiled.layout:
        c0, m, k, shape = item["col"], item["m"], item["k"], item["shape"]
        if item["kind"] == "categorical":
            width = k if k > 2 else 1
            block = np.asarray(row[c0:c0 + m * width], dtype=float).reshape(m, width)
            if transformed:
                if k > 2:
                    out.append(block[0].copy())  # 1-D one-hot vector
                else:
                    out.append(block[:, 0].copy() if shape else float(block[0, 0]))
            else:
                idx = block.argmax(axis=1) if k > 2 else ((block[:, 0] > 0.5).astype(int) if k == 2
                                                          else np.zeros(m, dtype=int))
                cats = item["categories"]
                if shape:
                    arr = np.empty(m, dtype=object)
                    for i, j in enumerate(idx):
                        arr[i] = cats[j]
                    out.append(arr)
                else:
                    out.append(cats[int(idx[0])])
            continue
        vals = np.asarray(row[c0:c0 + m], dtype=float)
        if item["kind"] == "integer" and not transformed:
            vals = vals.astype(int)
        out.append(vals.copy() if shape else vals[0].item())
    return tuple(out)


def _slot_columns(compiled):
    """Per-kind flat column indices and bounds, cached on the compiled space."""
    cache = getattr(compiled, "_cols", None)
    if cache is None:
        real = [s for s in compiled.slots if s["kind"] == _lib.KIND_REAL]
        integer = [s for s in compiled.slots if s["kind"] == _lib.KIND_INT]
        col = lambda group: np.asarray([s["out_col"] for s in group], dtype=np.intp)
        bound = lambda group, key: np.asarray([s[key] for s in group], dtype=np.float64)
        cache = dict(real=col(real), real_lo=bound(real, "low"), real_hi=bound(real, "high"),
                     int=col(integer), int_lo=bound(integer, "low"), int_hi=bound(integer, "high"))
        compiled._cols = cache
    return cache


def contains_rows(compiled, rows):
    """Vectorised `point in transformed_space` for a `[num x D_out]` array of flat, un-normalised rows of the
    `requires='real'` space: one boolean per row.

    Replaces the per-point Python check Orion runs on every suggestion (`primary_algo.py:74-81` ->
    `TransformedDimension.__contains__` `transformer.py:394-402` = `reverse`, then the original dimension's
    `__contains__`), with the same outcome:
      * real column: `low <= x < high` with `interval()`'s bounds, upper bound EXCLUSIVE (`space.py:336-356`);
      * integer column (`Reverse(Quantize)`): the reverse is `floor(x)` (`transformer.py:250`), which is integral by
        construction, so the test is `low <= floor(x) < high` with the integer interval (`space.py:425-448,467-483`);
      * categorical columns: `argmax` / `> 0.5` / zeros always name a valid category (`transformer.py:325-343`,
        `space.py:565-576`) -- nothing to check;
      * NaN fails every comparison, as it does in numpy."""
    rows = np.asarray(rows, dtype=np.float64)
    if rows.ndim != 2 or rows.shape[1] != compiled.d_out:
        raise ValueError("rows must be [num x %d]; got %r" % (compiled.d_out, rows.shape))
    c = _slot_columns(compiled)
    ok = np.ones(rows.shape[0], dtype=bool)
    if c["real"].size:
        x = rows[:, c["real"]]
        ok &= np.all((x >= c["real_lo"]) & (x < c["real_hi"]), axis=1)
    if c["int"].size:
        x = np.floor(rows[:, c["int"]])
        ok &= np.all((x >= c["int_lo"]) & (x < c["int_hi"]), axis=1)
    return ok


def reverse_rows(compiled, rows):
    """Vectorised `TransformedSpace.reverse` (`transformer.py:375,478-481`) for a `[num x D_out]` array of flat,
    un-normalised transformed rows: a list of `num` points of the ORIGINAL space (floats, Python ints after `floor`,
    category objects; 1-D arrays for shaped dimensions), equal to what the reference's per-point reverse returns."""
    rows = np.asarray(rows, dtype=np.float64)
    if rows.ndim != 2 or rows.shape[1] != compiled.d_out:
        raise ValueError("rows must be [num x %d]; got %r" % (compiled.d_out, rows.shape))
    num = rows.shape[0]
    columns = []  # per dimension: a [num] or [num x m] array (or object array) of reversed values
    for item in compiled.layout:
        c0, m, k, shape = item["col"], item["m"], item["k"], item["shape"]
        if item["kind"] == "categorical":
            width = k if k > 2 else 1
            block = rows[:, c0:c0 + m * width].reshape(num, m, width)
            if k > 2:
                idx = block.argmax(axis=2)
            elif k == 2:
                idx = (block[:, :, 0] > 0.5).astype(np.intp)
            else:
                idx = np.zeros((num, m), dtype=np.intp)
            cats = np.empty(len(item["categories"]), dtype=object)
            for j, cat in enumerate(item["categories"]):
                cats[j] = cat
            vals = cats[idx]
        elif item["kind"] == "integer":
            vals = np.floor(rows[:, c0:c0 + m]).astype(np.int64)
        else:
            vals = rows[:, c0:c0 + m]
        columns.append(vals if shape else vals[:, 0])
    points = []
    for i in range(num):
        points.append(tuple(col[i].copy() if col.ndim == 2 else (col[i] if col.dtype == object else col[i].item())
                            for col in columns))
    return points


def flatten_point(compiled, point):
    """Inverse of `regroup_row(..., transformed=True)`: tuple of scalars / 1-D arrays -> flat row."""
    flat = []
    for item, sub in zip(compiled.layout, point):
        flat.extend(np.asarray(sub, dtype=float).reshape(-1).tolist())
    if len(flat) != compiled.d_out:
        raise ValueError("point has %d flat entries, space has %d" % (len(flat), compiled.d_out))
    return flat


_MASK64 = (1 << 64) - 1


def fold_seed(seed):
    """int | tuple of ints | None -> 64-bit Philox key (splitmix64 fold; None = fresh entropy)."""
    if seed is None:
        return int.from_bytes(os.urandom(8), "little")
    if isinstance(seed, (int, np.integer)):
        parts = [int(seed)]
    else:
        parts = [int(x) for x in np.asarray(seed).reshape(-1)]
    h = 0x9E3779B97F4A7C15
    for p in parts:
        z = (h + (p & _MASK64) + 0x9E3779B97F4A7C15) & _MASK64
        z = ((z ^ (z >> 30)) * 0xBF58476D1CE4E5B9) & _MASK64
        z = ((z ^ (z >> 27)) * 0x94D049BB133111EB) & _MASK64
        h = z ^ (z >> 31)
    return h if len(parts) > 1 else (parts[0] & _MASK64)


def upload(engine, compiled):
    """Register `compiled` on the engine unless the identical table is already there."""
    if getattr(engine, "_space_key", None) != compiled.key:
        engine.set_space(compiled.slots, compiled.cdf)
        engine._space_key = compiled.key


def sample_points(space, n_samples=1, seed=None, engine=None):
    """Device-side `Space.sample`: returns a list of `n_samples` point tuples."""
    from .engine import get_engine
    engine = engine or get_engine()
    compiled = compile_space(space)
    upload(engine, compiled)
    mat = engine.sample(fold_seed(seed), 0, int(n_samples), normalise=False).cpu().numpy()
    transformed = hasattr(space.values()[0], "original_dimension")
    return [regroup_row(compiled, row, transformed=transformed) for row in mat]
