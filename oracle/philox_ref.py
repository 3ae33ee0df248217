"""TEST INFRASTRUCTURE ONLY -- numpy restatement of the device noise generator.

Philox4x32-10 (Salmon, Moraes, Dror, Shaw: "Parallel random numbers: as easy as 1, 2,
3", SC'11) with the counter layout of lmc_atomi_b200/csrc/mix_kernels.cu:
counter = (block index, stream, chain_lo, chain_hi), key = (seed_lo, seed_hi);
Box-Muller on two 32-bit words.  Not a restatement of reference code (the reference
uses numpy's PCG64 stream, which the validation mode consumes directly).
"""
import numpy as np

M0, M1 = np.uint64(0xD2511F53), np.uint64(0xCD9E8D57)
W0, W1 = 0x9E3779B9, 0xBB67AE85
MASK = np.uint64(0xFFFFFFFF)


def philox4x32_10(c0, c1, c2, c3, k0, k1):
    c0, c1, c2, c3 = (np.asarray(x, dtype=np.uint64) for x in (c0, c1, c2, c3))
    k0, k1 = int(k0), int(k1)
    for _ in range(10):
        p0 = M0 * c0
        p1 = M1 * c2
        hi0, lo0 = p0 >> np.uint64(32), p0 & MASK
        hi1, lo1 = p1 >> np.uint64(32), p1 & MASK
        c0, c1, c2, c3 = hi1 ^ c1 ^ np.uint64(k0), lo1, hi0 ^ c3 ^ np.uint64(k1), lo0
        k0 = (k0 + W0) & 0xFFFFFFFF
        k1 = (k1 + W1) & 0xFFFFFFFF
    return c0, c1, c2, c3


def _normals(a, b, bits24=False):
    a = a.astype(np.float64)
    b = b.astype(np.float64)
    if bits24:   # the fp32 kernel keeps the top 24 bits
        u1 = (np.floor(a / 256) + 1.0) * 2.0 ** -24
        u2 = np.floor(b / 256) * 2.0 ** -24
    else:
        u1 = (a + 1.0) * 2.0 ** -32
        u2 = b * 2.0 ** -32
    r = np.sqrt(-2.0 * np.log(u1))
    return r * np.cos(2 * np.pi * u2), r * np.sin(2 * np.pi * u2)


def draws(C, K, seed, chain_offset=0, adjusted=False, bits24=False):
    """theta0 (C,2), normals (C,K,2)[, uniforms (C,K)] as the kernels draw them."""
    k0, k1 = seed & 0xFFFFFFFF, (seed >> 32) & 0xFFFFFFFF
    gc = np.arange(C, dtype=np.uint64) + np.uint64(chain_offset)
    lo, hi = gc & MASK, gc >> np.uint64(32)
    r = philox4x32_10(np.zeros(C), np.ones(C), lo, hi, k0, k1)
    z0, z1 = _normals(r[0], r[1], bits24)
    theta0 = np.stack([z0, z1], axis=1)
    xi = np.empty((C, K, 2))
    un = np.empty((C, K))
    if adjusted:
        # step k takes words 3k, 3k + 1 (normals) and 3k + 2 (uniform) of the chain's stream of blocks 0, 1, 2, ...
        # (mix_kernels.cu run_production: three blocks per four steps)
        nblk = (3 * K + 3) // 4 + 1
        words = np.concatenate([np.stack(philox4x32_10(np.full(C, b), np.zeros(C), lo, hi, k0, k1), axis=1)
                                for b in range(nblk)], axis=1)          # (C, 4 nblk)
        for k in range(K):
            a, b, u = words[:, 3 * k], words[:, 3 * k + 1], words[:, 3 * k + 2]
            xi[:, k, 0], xi[:, k, 1] = _normals(a, b, bits24)
            un[:, k] = u.astype(np.float64) * 2.0 ** -32 if not bits24 else np.floor(u.astype(np.float64) / 256) * 2.0 ** -24
        return theta0, xi, un
    for kb in range((K + 1) // 2):
        r = philox4x32_10(np.full(C, kb), np.zeros(C), lo, hi, k0, k1)
        xi[:, 2 * kb, 0], xi[:, 2 * kb, 1] = _normals(r[0], r[1], bits24)
        if 2 * kb + 1 < K:
            xi[:, 2 * kb + 1, 0], xi[:, 2 * kb + 1, 1] = _normals(r[2], r[3], bits24)
    return theta0, xi
