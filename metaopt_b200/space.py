"""Host-side mirror of the reference's search-space interface for the hot path.

Same public names, argument meaning and error behaviour as `orion.algo.space`
(`/root/reference/src/orion/algo/space.py`: Dimension :69, Real :291, Integer :454, Categorical
:500, Space :732) and of the `requires='real'` view built by
`orion.core.worker.transformer.build_required_space` (`transformer.py:21-77`, TransformedDimension
:356, TransformedSpace :468) -- but restricted to what the B200 path can express as a flat slot
table: closed-form priors (uniform, reciprocal/loguniform, norm) and scalar or 1-D shapes.

It exists because the GPU box has no Orion install: tests and `bench.py` build spaces with these
classes, while under a real Orion the plugin receives Orion's own objects; `compiler.compile_space`
accepts either (duck typing on the attributes listed in `compiler.py`).

Sampling is NOT done here on the CPU: `Space.sample` / `Dimension.sample` go to the device sampler
(`ob200_sample_transform`); without a GPU they raise.
"""
import math
import numbers

import numpy as np

_SUPPORTED_PRIORS = {
    "uniform": "uniform", "reciprocal": "reciprocal", "loguniform": "reciprocal",
    "norm": "norm", "normal": "norm", "gaussian": "norm",
}


def prior_support(prior_name, args):
    """`scipy.stats.<prior>.interval(1.0, *args)` in closed form for the supported priors."""
    if prior_name == "uniform":
        loc, scale = (list(args) + [0.0, 1.0][len(args):])[:2]
        return float(loc), float(loc) + float(scale)
    if prior_name == "reciprocal":
        a, b = args[:2]
        return float(a), float(b)
    if prior_name == "norm":
        return -math.inf, math.inf
    raise NotImplementedError("prior %r has no device implementation" % (prior_name,))


def _numeric(point):
    def one(x):
        return isinstance(x, (numbers.Number, np.ndarray))
    try:
        return all(one(x) for x in point)
    except TypeError:
        return one(point)


class Dimension:
    """One named search-space dimension (mirror of `space.py:69-275`)."""

    NO_DEFAULT_VALUE = None

    def __init__(self, name, prior, *args, **kwargs):
        if not (isinstance(name, str) or name is None):
            raise TypeError("Dimension's name must be either string or None. "
                            "Provided: {}, of type: {}".format(name, type(name)))
        self.name = name
        for banned, msg in (("random_state", "random_state/seed cannot be set in a parameter's definition! "
                                             "Set seed globally!"),
                            ("seed", "random_state/seed cannot be set in a parameter's definition! "
                                     "Set seed globally!"),
                            ("discrete", "Do not use kwarg 'discrete' on `Dimension`, use pure `_Discrete` "
                                         "class instead!"),
                            ("size", "Use 'shape' keyword only instead of 'size'.")):
            if banned in kwargs:
                raise ValueError(msg)
        self._default_value = kwargs.pop("default_value", self.NO_DEFAULT_VALUE)
        shape = kwargs.pop("shape", None)
        self._shape = shape
        if prior is not None:
            if not isinstance(prior, str):
                prior = getattr(prior, "name", str(prior))
            if prior not in _SUPPORTED_PRIORS:
                raise NotImplementedError("prior %r has no device implementation (supported: %s)"
                                          % (prior, sorted(set(_SUPPORTED_PRIORS))))
            # loc/scale may come as keywords like with scipy
            if "loc" in kwargs or "scale" in kwargs:
                args = (kwargs.pop("loc", 0.0), kwargs.pop("scale", 1.0))
            prior = _SUPPORTED_PRIORS[prior]
        self._prior_name = prior if prior is not None else "None"
        self._args = tuple(args)
        self._kwargs = dict(kwargs)
        if self._default_value is not self.NO_DEFAULT_VALUE and self._default_value not in self:
            raise ValueError("{} is not a valid value for this Dimension. "
                             "Can't set default value.".format(self._default_value))

    # -- identity -------------------------------------------------------------------------
    @property
    def type(self):
        return self.__class__.__name__.lower()

    @property
    def default_value(self):
        return self._default_value

    @property
    def shape(self):
        if self._shape is None:
            return ()
        if isinstance(self._shape, numbers.Integral):
            return (int(self._shape),)
        return tuple(int(s) for s in self._shape)

    def _key(self):
        return (self.name, self.shape, self.type, self._args, tuple(self._kwargs.items()),
                self._default_value, self._prior_name)

    def __eq__(self, other):
        return isinstance(other, Dimension) and self._key() == other._key()

    def __hash__(self):
        return hash(self._key())

    def __repr__(self):
        return "{0}(name={1}, prior={{{2}: {3}, {4}}}, shape={5}, default value={6})".format(
            self.__class__.__name__, self.name, self._prior_name, self._args, self._kwargs, self.shape,
            self._default_value)

    # -- behaviour ------------------------------------------------------------------------
    def interval(self, alpha=1.0):
        """(low, high): lower bound inclusive, upper bound exclusive."""
        if alpha != 1.0:
            raise NotImplementedError("only the full support (alpha=1.0) is available")
        return prior_support(self._prior_name, self._args)

    def sample(self, n_samples=1, seed=None):
        """Draw from the prior on the device (one-dimension space)."""
        space = Space()
        space.register(self)
        return [p[0] for p in space.sample(n_samples, seed)]

    def cast(self, point):
        raise NotImplementedError

    def __contains__(self, point):
        raise NotImplementedError


class Real(Dimension):
    """Real-valued dimension; `low` inclusive / `high` exclusive clip the prior's support."""

    def __init__(self, name, prior, *args, **kwargs):
        self._low = kwargs.pop("low", -np.inf)
        self._high = kwargs.pop("high", np.inf)
        if self._high <= self._low:
            raise ValueError("Lower bound {} has to be less than upper bound {}".format(self._low, self._high))
        super().__init__(name, prior, *args, **kwargs)

    def interval(self, alpha=1.0):
        lo, hi = Dimension.interval(self, alpha)
        return max(lo, self._low), min(hi, self._high)

    def __contains__(self, point):
        if not _numeric(point):
            return False
        low, high = self.interval()
        arr = np.asarray(point)
        if arr.shape != self.shape:
            return False
        return bool(np.all(arr < high) and np.all(arr >= low))

    def cast(self, point):
        out = np.asarray(point).astype(float)
        return out if isinstance(point, np.ndarray) else out.tolist()


class Integer(Real):
    """Integer dimension: samples are `floor` of the real prior draw (`space.py:410-423`)."""

    def interval(self, alpha=1.0):
        low, high = Real.interval(self, alpha)
        int_low = -np.inf if math.isinf(low) else int(np.floor(low))
        if math.isinf(high):
            int_high = np.inf
        else:
            int_high = int(np.floor(high))
            if int_high < high:  # exclusive upper bound
                int_high += 1
        return int_low, int_high

    def __contains__(self, point):
        if not _numeric(point):
            return False
        arr = np.asarray(point)
        if not np.all(np.equal(np.mod(arr, 1), 0)):
            return False
        return Real.__contains__(self, point)

    def cast(self, point):
        out = np.asarray(point).astype(int)
        return out if isinstance(point, np.ndarray) else out.tolist()


class Categorical(Dimension):
    """Unordered choices; a dict maps category -> probability, any other iterable is uniform."""

    def __init__(self, name, categories, **kwargs):
        if isinstance(categories, dict):
            self.categories = tuple(categories.keys())
            self._probs = tuple(categories.values())
        else:
            self.categories = tuple(categories)
            self._probs = tuple(np.tile(1.0 / len(self.categories), len(self.categories)))
        super().__init__(name, None, **kwargs)
        self._prior_name = "choices"

    def interval(self, alpha=1.0):
        raise RuntimeError("Categories have no ``interval`` (as they are not ordered).\n"
                           "Use ``self.categories`` instead.")

    def __contains__(self, point):
        arr = np.asarray(point, dtype=object)
        if arr.shape != self.shape:
            return False
        return all(x in self.categories for x in arr.reshape(-1))

    def cast(self, point):
        by_str = {str(c): c for c in self.categories}

        def one(v):
            if str(v) not in by_str:
                raise ValueError("Invalid category: {}".format(v))
            return by_str[str(v)]

        arr = np.asarray(point, dtype=object)
        out = np.empty(arr.shape, dtype=object)
        flat = out.reshape(-1)
        for i, v in enumerate(arr.reshape(-1)):
            flat[i] = one(v)
        return out if isinstance(point, np.ndarray) else out.tolist()

    def __repr__(self):
        prior = "{" + ", ".join("{}: {:.2f}".format(c, p) for c, p in zip(self.categories, self._probs)) + "}"
        return "Categorical(name={0}, prior={1}, shape={2}, default value={3})".format(
            self.name, prior, self.shape, self.default_value)


class Space(dict):
    """Name-sorted dictionary of dimensions (`space.py:732-858`)."""

    contains = Dimension

    def register(self, dimension):
        self[dimension.name] = dimension

    def __setitem__(self, key, value):
        if not isinstance(key, str):
            raise TypeError("Keys registered to {} must be string types. Provided: {}".format(
                self.__class__.__name__, key))
        if not isinstance(value, self.contains):
            raise TypeError("Values registered to {} must be {} types. Provided: {}".format(
                self.__class__.__name__, self.contains.__name__, value))
        if key in self:
            raise ValueError("There is already a Dimension registered with this name. "
                             "Register it with another name. Provided: {}".format(key))
        super().__setitem__(key, value)

    def __getitem__(self, key):
        if isinstance(key, str):
            return super().__getitem__(key)
        return self.values()[key]

    def __iter__(self):
        return iter(sorted(super().keys()))

    def keys(self):
        return list(iter(self))

    def values(self):
        return [super(Space, self).__getitem__(k) for k in self.keys()]

    def items(self):
        return [(k, super(Space, self).__getitem__(k)) for k in self.keys()]

    def interval(self, alpha=1.0):
        return [d.categories if d.type == "categorical" else d.interval(alpha) for d in self.values()]

    def __contains__(self, value):
        if isinstance(value, str):
            return super().__contains__(value)
        try:
            len(value)
        except TypeError as exc:
            raise TypeError("Can check only for dimension names or for tuples with parameter values.") from exc
        if not self:
            return False
        return all(component in dim for component, dim in zip(value, self.values()))

    def __repr__(self):
        return "Space([{}])".format(",\n       ".join(map(str, self.values())))

    def sample(self, n_samples=1, seed=None):
        """`n_samples` points as tuples, one entry per dimension, drawn on the device.

        `seed`: int (or tuple of ints, folded) selecting the Philox stream; None draws a fresh one."""
        from . import compiler  # local import: needs torch + CUDA
        return compiler.sample_points(self, n_samples, seed)


# ---------------------------------------------------------------------------------------------
# requires='real' view (mirror of transformer.py:21-77, :356-481) in closed form
# ---------------------------------------------------------------------------------------------
class TransformedDimension:
    """A dimension as a 'real'-requiring algorithm sees it: reals unchanged, integers as floats,
    categoricals enumerated then one-hot encoded (k>2), 0/1 float (k==2) or 0.0 (k==1)."""

    NO_DEFAULT_VALUE = None

    def __init__(self, original_dimension):
        self.original_dimension = original_dimension
        t = original_dimension.type
        if t not in ("real", "integer", "categorical"):
            raise TypeError("Unsupported dimension type ('{}') or requirement ('real')".format(t))
        self._kind = t
        self._k = len(original_dimension.categories) if t == "categorical" else 0

    name = property(lambda self: self.original_dimension.name)
    type = property(lambda self: "real")

    @property
    def shape(self):
        shape = tuple(self.original_dimension.shape or ())
        return shape + (self._k,) if self._k > 2 else shape

    def transform(self, point):
        if self._kind == "real":
            return point
        if self._kind == "integer":
            return np.asarray(point).astype(float)
        cats = self.original_dimension.categories
        arr = np.asarray(point, dtype=object)
        idx = np.empty(arr.shape, dtype=np.int32)
        flat = idx.reshape(-1)
        for i, v in enumerate(arr.reshape(-1)):
            flat[i] = cats.index(v)
        if self._k <= 2:
            return np.asarray(idx, dtype=float)
        hot = np.zeros(idx.shape + (self._k,))
        np.put_along_axis(hot, idx[..., None].astype(np.int64), 1.0, axis=-1)
        return hot

    def reverse(self, transformed_point):
        if self._kind == "real":
            return transformed_point
        if self._kind == "integer":
            return np.floor(np.asarray(transformed_point)).astype(int)
        x = np.asarray(transformed_point)
        if self._k == 2:
            idx = (x > 0.5).astype(int)
        elif self._k == 1:
            idx = np.zeros_like(x, dtype=int)
        else:
            assert x.shape[-1] == self._k
            idx = x.argmax(axis=-1)
        cats = self.original_dimension.categories
        out = np.empty(idx.shape, dtype=object)
        flat = out.reshape(-1)
        for i, v in enumerate(np.asarray(idx).reshape(-1)):
            flat[i] = cats[v]
        return out if out.shape else out[()]

    def interval(self, alpha=1.0):
        if self._kind == "categorical":
            return (-0.1, 1.1)
        low, high = self.original_dimension.interval(alpha)
        return self.transform(low), self.transform(high)

    def __contains__(self, point):
        try:
            orig = self.reverse(point)
        except AssertionError:
            return False
        return orig in self.original_dimension

    def sample(self, n_samples=1, seed=None):
        return [self.transform(s) for s in self.original_dimension.sample(n_samples, seed)]

    def __repr__(self):
        return "Transformed({!r})".format(self.original_dimension)


class TransformedSpace(Space):
    contains = TransformedDimension

    def transform(self, point):
        return tuple(dim.transform(point[i]) for i, dim in enumerate(self.values()))

    def reverse(self, transformed_point):
        return tuple(dim.reverse(transformed_point[i]) for i, dim in enumerate(self.values()))


def build_required_space(requirements, original_space):
    """Mirror of `transformer.build_required_space` for the one requirement this plugin has."""
    reqs = requirements if isinstance(requirements, list) else [requirements]
    if reqs != ["real"]:
        raise TypeError("Unsupported dimension type or requirement ('{}')".format(requirements))
    space = TransformedSpace()
    for dim in original_space.values():
        space.register(TransformedDimension(dim))
    return space
