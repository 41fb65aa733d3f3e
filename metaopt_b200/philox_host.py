"""Single-call Philox4x32-10 on the host (pure Python ints) for the few scalar draws the plugin
makes itself (gp_hedge's categorical pick).  Same generator as the device sampler
(`csrc/ob200_internal.cuh`), on a reserved counter lane (c2 = 0xFFFFFFFF) so it never collides
with candidate draws."""

_M0, _M1 = 0xD2511F53, 0xCD9E8D57
_W0, _W1 = 0x9E3779B9, 0xBB67AE85
_MASK = 0xFFFFFFFF


def philox4x32_10(c0, c1, c2, c3, k0, k1):
    for _ in range(10):
        p0, p1 = _M0 * c0, _M1 * c2
        c0, c1, c2, c3 = ((p1 >> 32) ^ c1 ^ k0) & _MASK, p1 & _MASK, ((p0 >> 32) ^ c3 ^ k1) & _MASK, p0 & _MASK
        k0, k1 = (k0 + _W0) & _MASK, (k1 + _W1) & _MASK
    return c0, c1, c2, c3


def uniform01(seed, index):
    """[0,1) double for (seed, index) on the host lane."""
    seed &= (1 << 64) - 1
    index &= (1 << 64) - 1
    r0, r1, _, _ = philox4x32_10(index & _MASK, index >> 32, 0xFFFFFFFF, 0, seed & _MASK, seed >> 32)
    return ((r1 << 32 | r0) >> 11) * 2.0 ** -53
