"""Sampling + transform + normalise restatement (test infrastructure; see oracle/__init__.py).

Follows, per *slot* (one scalar source variate = one flat entry of an Orion dimension):

* prior closed forms of `Dimension.sample` -> scipy `rvs`
  (`src/orion/algo/space.py:160-186`; `src/orion/core/io/space_builder.py:149-179`):
  uniform(a, b) = loc + scale*U ; loguniform(a, b) = exp(log a + U*(log b - log a)) ;
  normal(mu, s) = mu + s*Z.
* `Real.sample` rejection, at most 4 tries until `low <= x < high`, else "Improbable
  bounds" (`space.py:371-391`, `:336-356`).
* `_Discrete.sample`: `floor(x).astype(int)` *before* the containment check
  (`space.py:410-423`, MRO of `Integer(Real, _Discrete)` `:454`).
* `Categorical.sample` = `rng.choice(categories, p)` == `searchsorted(cumsum(p)/cumsum(p)[-1],
  U, 'right')` (`space.py:542-552`).
* transforms for `requires='real'` (`src/orion/core/worker/transformer.py:51-66`):
  integer -> `astype(float)` (`:219-221,252-254`); categorical -> Enumerate (`:269-282`) ->
  OneHotEncode (`:304-323`): k>2 one-hot of 0.0/1.0 on a new last axis, k==2 float(index),
  k==1 0.0.
* normalise (skopt `Normalize`, not in the reference tree): `(x - low)/(high - low)` with the
  bounds of `TransformedDimension.interval()` (`transformer.py:384-392`).

Slot table columns (same meaning as `ob200_slot` in include/ob200.h).
"""
import numpy as np

from . import philox

KIND_REAL, KIND_INT, KIND_CAT = 0, 1, 2
PRIOR_UNIFORM, PRIOR_LOGUNIFORM, PRIOR_NORMAL, PRIOR_CHOICE = 0, 1, 2, 3
STATUS_OK, STATUS_IMPROBABLE = 0, 1
TWO_PI = 6.283185307179586


def _draw(slot, u, z=None):
    if slot['prior'] == PRIOR_UNIFORM:
        return slot['p0'] + slot['p1'] * u
    if slot['prior'] == PRIOR_LOGUNIFORM:
        return np.exp(slot['p0'] + u * slot['p1'])
    if slot['prior'] == PRIOR_NORMAL:
        return slot['p0'] + slot['p1'] * z
    raise ValueError(slot['prior'])


def box_muller(ua, ub):
    """Standard normal from two [0,1) uniforms (device formula; not a reference algorithm)."""
    return np.sqrt(-2.0 * np.log(1.0 - ua)) * np.cos(TWO_PI * ub)


def one_hot(index, k):
    """`OneHotEncode.transform` restated (`transformer.py:304-323`); index: int array [...]."""
    index = np.asarray(index)
    assert np.all(index < k) and np.all(index >= 0) and np.all(index % 1 == 0)
    if k <= 2:
        return np.asarray(index, dtype=float)
    hot = np.zeros(tuple(index.shape) + (k,))
    np.put_along_axis(hot, index[..., None].astype(np.int64), 1.0, axis=-1)
    return hot


def one_hot_reverse(x, k):
    """`OneHotEncode.reverse` restated (`transformer.py:325-343`)."""
    x = np.asarray(x)
    if k == 2:
        return (x > 0.5).astype(int)
    if k == 1:
        return np.zeros_like(x, dtype=int)
    assert x.shape[-1] == k
    return x.argmax(axis=-1)


def slot_op(s):
    """Operation code of a slot (same numbering as the device kernel)."""
    if s['kind'] == KIND_CAT:
        return 6 if s['ncat'] > 2 else 7
    base = {PRIOR_UNIFORM: 0, PRIOR_LOGUNIFORM: 2, PRIOR_NORMAL: 4}[s['prior']]
    return base + (1 if s['kind'] == KIND_INT else 0)


NARROW_OPS = (1, 6, 7)          # uniform int, one-hot, small categorical: 32 random bits per draw, four slots per call
WIDE_OPS = (0, 2, 3, 4, 5)      # uniform real, loguniform real / int, normal real / int: 53 bits, two slots per call


def rng_positions(slots):
    """(pos, n_narrow): position q of every slot in the device's RNG order (device contract version 2,
    ob200_sample.cu / oracle/philox.py): narrow groups first, each padded to a multiple of 4, then the wide groups,
    each padded to an even size; table order inside a group."""
    pos, q = [0] * len(slots), 0
    for op in NARROW_OPS:
        for i, s in enumerate(slots):
            if slot_op(s) == op:
                pos[i] = q
                q += 1
        q += (-q) % 4
    n_narrow = q
    for op in WIDE_OPS:
        for i, s in enumerate(slots):
            if slot_op(s) == op:
                pos[i] = q
                q += 1
        q += q & 1
    return pos, n_narrow


def sample_transform(slots, cdf, seed, first_index, count, normalise=True, variates=None):
    """Return (out [count x D_out] float64, status).

    slots: list of dicts (kind, prior, out_col, width, ncat, cdf_off, p0, p1, low, high, nlo, nhi)
    variates: optional [count x n_slots] host-supplied U (uniform/loguniform/choice slots) or
    Z (normal slots); single try, no resampling (parity mode against the reference's own draws).
    """
    d_out = max(s['out_col'] + s['width'] for s in slots) if slots else 0
    out = np.zeros((count, d_out))
    status = STATUS_OK
    idx = np.arange(first_index, first_index + count, dtype=np.uint64)
    qpos, n_narrow = rng_positions(slots)
    for si, s in enumerate(slots):
        q = qpos[si]
        value = np.zeros(count)
        done = np.zeros(count, dtype=bool)
        tries = 1 if variates is not None else 4
        for t in range(tries):
            if variates is not None:
                u = z = np.asarray(variates[:, si], dtype=np.float64)
            elif s['prior'] == PRIOR_NORMAL:
                ua, ub = philox.uniform_pair(seed, idx, 0x80000000 | q, t)
                u, z = None, box_muller(ua, ub)
            elif q < n_narrow:
                r = philox.words(seed, idx, q >> 2, t)
                u, z = r[q & 3].astype(np.float64) * 2.0 ** -32, None
            else:
                ua, ub = philox.uniform_pair(seed, idx, n_narrow // 4 + (q - n_narrow) // 2, t)
                u, z = (ua if q % 2 == 0 else ub), None
            if s['kind'] == KIND_CAT:
                c = np.asarray(cdf[s['cdf_off']:s['cdf_off'] + s['ncat']])
                x = np.searchsorted(c, u, side='right').astype(np.float64)
                x = np.minimum(x, s['ncat'] - 1)
                ok = np.ones(count, dtype=bool)
            else:
                x = _draw(s, u, z)
                if s['kind'] == KIND_INT:
                    x = np.floor(x)
                ok = (x >= s['low']) & (x < s['high'])
            take = ok & ~done
            value[take] = x[take]
            done |= ok
        if not done.all():
            status = STATUS_IMPROBABLE
        c0, w = s['out_col'], s['width']
        if s['kind'] == KIND_CAT and s['ncat'] > 2:
            block = one_hot(value.astype(np.int64), s['ncat'])
        elif s['kind'] == KIND_CAT and s['ncat'] == 1:
            block = np.zeros((count, 1))
        else:
            block = value[:, None]
        if normalise:
            block = (block - s['nlo']) / (s['nhi'] - s['nlo'])
        out[:, c0:c0 + w] = block
    return out, status
