"""Philox4x32-10 (Salmon et al., SC'11; the Random123 reference vectors pin it) in NumPy, and the
mapping from (seed, generation, pair, gene) to the uniforms that csrc/variation.cu draws.

TEST INFRASTRUCTURE.  The reference draws from jax.random (Threefry, one array per draw); the CUDA
variation kernel draws from counter-based Philox instead, so the two cannot share random numbers.
What CAN be held to the reference is the arithmetic: regenerate the kernel's uniforms here, feed them
to the reference's own `simulated_binary` / `polynomial` source (tests/golden/make_golden_variation.py
patches `jax.random` with these arrays) and compare the kernel's output with what comes out.
"""
import numpy as np

M0, M1 = np.uint64(0xD2511F53), np.uint64(0xCD9E8D57)
W0, W1 = 0x9E3779B9, 0xBB67AE85
MASK = np.uint64(0xFFFFFFFF)
KEY1 = 0x5BD1E995                      # second key word used by k_sbx_pm


def philox4x32_10(c0, c1, c2, c3, k0, k1):
    """Vectorised over the counter words (uint32 arrays or scalars); k0, k1 Python ints.
    Returns four uint32 arrays."""
    c = [np.asarray(v, dtype=np.uint64) & MASK for v in np.broadcast_arrays(c0, c1, c2, c3)]
    k0, k1 = int(k0) & 0xFFFFFFFF, int(k1) & 0xFFFFFFFF
    for _ in range(10):
        p0, p1 = M0 * c[0], M1 * c[2]
        hi0, lo0, hi1, lo1 = p0 >> np.uint64(32), p0 & MASK, p1 >> np.uint64(32), p1 & MASK
        c = [hi1 ^ c[1] ^ np.uint64(k0), lo1, hi0 ^ c[3] ^ np.uint64(k1), lo0]
        k0, k1 = (k0 + W0) & 0xFFFFFFFF, (k1 + W1) & 0xFFFFFFFF
    return tuple(v.astype(np.uint32) for v in c)


def u01(bits):
    """csrc/variation.cu `u01`: the top 24 bits as a float32 in [0, 1)."""
    return ((np.asarray(bits, np.uint32) >> np.uint32(8)).astype(np.float32) * np.float32(1.0 / 16777216.0)).astype(np.float32)


def variation_draws(n, d, seed, generation):
    """Every random quantity k_sbx_pm uses for a population of n rows x d genes, as arrays shaped the
    way the reference's operators draw them (pairs = n // 2):
      sbx_mu [pairs, d], sbx_sign [pairs, d] (0/1), sbx_keep [pairs, d], sbx_pair [pairs, 1],
      pm_site [2*pairs, d], pm_mu [2*pairs, d]   (rows: first children, then second children)."""
    half = n // 2
    t = np.arange(half * d, dtype=np.uint64)
    lo, hi = (t & MASK), (t >> np.uint64(32))
    a = philox4x32_10(lo, hi, 0, generation, seed, KEY1)
    b = philox4x32_10(lo, hi, 1, generation, seed, KEY1)
    p = np.arange(half, dtype=np.uint64)
    pr = philox4x32_10(p & MASK, p >> np.uint64(32), 2, generation, seed, KEY1)
    sh = (half, d)
    return {
        "sbx_mu": u01(a[0]).reshape(sh),
        "sbx_sign": (a[1] & np.uint32(1)).astype(np.int32).reshape(sh),
        "sbx_keep": u01(a[2]).reshape(sh),
        "sbx_pair": u01(pr[0]).reshape(half, 1),
        "pm_site": np.concatenate([u01(a[3]).reshape(sh), u01(b[1]).reshape(sh)], axis=0),
        "pm_mu": np.concatenate([u01(b[0]).reshape(sh), u01(b[2]).reshape(sh)], axis=0),
    }
