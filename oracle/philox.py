"""Philox4x32-10 counter-based generator, numpy restatement (test infrastructure).

Not a reference algorithm: the reference samples with numpy's MT19937
(`src/orion/algo/space.py:47-60,745-773`), which is inherently sequential.  The B200
path replaces it by Philox (Salmon et al., SC'11: multipliers 0xD2511F53/0xCD9E8D57,
Weyl keys 0x9E3779B9/0xBB67AE85, 10 rounds), keyed so that candidate i is the same
whatever the sharding.  This file restates the device generator bit for bit so the
sampled integers/uniforms can be checked exactly.

Counter layout used by `ob200_sample_transform` (metaopt_b200/csrc/ob200_sample.cu):
  counter = (index_lo, index_hi, c2, try), key = (seed_lo, seed_hi)
  q = position of the slot in the device's grouped order (oracle.sampling.rng_positions)
  c2 = q>>1 for uniform/loguniform/choice slots (component q&1 of the pair),
  c2 = 0x80000000|q for normal slots (Box-Muller of the pair)
  -> 4 x uint32 (r0..r3) -> two 53-bit uniforms in [0,1):
       u_a = ((r1<<32 | r0) >> 11) * 2**-53 ,  u_b = ((r3<<32 | r2) >> 11) * 2**-53
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
