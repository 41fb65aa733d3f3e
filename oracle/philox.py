"""Philox4x32-10 counter-based generator, numpy restatement (test infrastructure).

Not a reference algorithm: the reference samples with numpy's MT19937
(`src/orion/algo/space.py:47-60,745-773`), which is inherently sequential.  The B200
path replaces it by Philox (Salmon et al., SC'11: multipliers 0xD2511F53/0xCD9E8D57,
Weyl keys 0x9E3779B9/0xBB67AE85, 10 rounds), keyed so that candidate i is the same
whatever the sharding.  This file restates the device generator bit for bit so the
sampled integers/uniforms can be checked exactly.

Counter layout used by `ob200_sample_transform` (metaopt_b200/csrc/ob200_sample.cu), RNG contract version 2:
  counter = (index_lo, index_hi, c2, try), key = (seed_lo, seed_hi)
  q = position of the slot in the device's grouped order (oracle.sampling.rng_positions): the NARROW groups first
      (uniform int, one-hot, small categorical; each padded to a multiple of 4), then the WIDE groups (uniform real,
      loguniform real, loguniform int, normal real, normal int; each padded to an even size); n_narrow = number of
      narrow positions
  narrow slot:  c2 = q >> 2,                                   u = r_(q & 3) * 2**-32            (32 random bits)
  wide uniform / loguniform slot:  c2 = n_narrow / 4 + (q - n_narrow) // 2,  component (q & 1) of
       u_a = ((r1<<32 | r0) >> 11) * 2**-53 ,  u_b = ((r3<<32 | r2) >> 11) * 2**-53            (53 random bits)
  normal slot:  c2 = 0x80000000 | q, Box-Muller of (u_a, u_b)
(Version 1, round 1, gave every slot 53 bits, two slots per call; a categorical or uniform-integer draw needs no more
than 32, and the Philox rounds were half of the sampler's instructions.)
"""
import numpy as np

M0 = np.uint64(0xD2511F53)
M1 = np.uint64(0xCD9E8D57)
W0 = 0x9E3779B9
W1 = 0xBB67AE85
MASK32 = np.uint64(0xFFFFFFFF)
S32 = np.uint64(32)


def philox4x32_10(c0, c1, c2, c3, k0, k1):
    """All arguments broadcastable uint32-valued arrays; returns 4 uint64 arrays (<2**32)."""
    c0 = np.asarray(c0, dtype=np.uint64) & MASK32
    c1 = np.asarray(c1, dtype=np.uint64) & MASK32
    c2 = np.asarray(c2, dtype=np.uint64) & MASK32
    c3 = np.asarray(c3, dtype=np.uint64) & MASK32
    k0 = int(k0) & 0xFFFFFFFF
    k1 = int(k1) & 0xFFFFFFFF
    c0, c1, c2, c3 = np.broadcast_arrays(c0, c1, c2, c3)
    for _ in range(10):
        p0 = M0 * c0
        p1 = M1 * c2
        hi0, lo0 = p0 >> S32, p0 & MASK32
        hi1, lo1 = p1 >> S32, p1 & MASK32
        c0, c1, c2, c3 = (hi1 ^ c1 ^ np.uint64(k0)), lo1, (hi0 ^ c3 ^ np.uint64(k1)), lo0
        k0 = (k0 + W0) & 0xFFFFFFFF
        k1 = (k1 + W1) & 0xFFFFFFFF
    return c0, c1, c2, c3


def words(seed, index, slot, block):
    """The four raw 32-bit outputs (as uint64 arrays) for (seed, global point index, c2 = slot, try = block)."""
    seed = int(seed) & 0xFFFFFFFFFFFFFFFF
    index = np.asarray(index, dtype=np.uint64)
    return philox4x32_10(index & MASK32, index >> S32, slot, block, seed & 0xFFFFFFFF, seed >> 32)


def uniform_pair(seed, index, slot, block):
    """Two [0,1) doubles for (seed, global point index, slot, block)."""
    seed = int(seed) & 0xFFFFFFFFFFFFFFFF
    index = np.asarray(index, dtype=np.uint64)
    r0, r1, r2, r3 = philox4x32_10(index & MASK32, index >> S32, slot, block,
                                   seed & 0xFFFFFFFF, seed >> 32)
    two53 = 2.0 ** -53
    ua = (((r1 << S32) | r0) >> np.uint64(11)).astype(np.float64) * two53
    ub = (((r3 << S32) | r2) >> np.uint64(11)).astype(np.float64) * two53
    return ua, ub
