"""Restatement of the JAX PRNG pieces the reference's tests use to make inputs.

TEST INFRASTRUCTURE ONLY (see oracle/README.md).  Three reference tests draw
their inputs from `jax.random` (tests/test_lsmop.py:10-14, tests/test_maf.py:10-11,
tests/test_crowding_distance.py:25-26); jaxlib is a third-party dependency that is
absent here (pyproject.toml:33-34, unpinned `jax >= 0.4.16`).  This file restates
the published algorithm - Threefry-2x32 with 20 rounds (Salmon et al., SC'11) as
wired up by jax._src.prng for the default, non-partitionable key implementation -
so those inputs can be regenerated bit-for-bit and the tests' golden outputs used
as known-answer vectors.  The restatement is itself pinned by those goldens: 18
LSMOP spot values, the MaF1 spot value and 90 crowding distances only come out
right if every bit of the input is right.
"""
import numpy as np
from scipy.special import erfinv

_U32 = np.uint32


def _rotl(x, r):
    return (x << _U32(r)) | (x >> _U32(32 - r))


def threefry2x32(k1, k2, x0, x1):
    """20-round Threefry-2x32 block function on uint32 arrays x0, x1."""
    with np.errstate(over="ignore"):
        k1, k2 = _U32(k1), _U32(k2)
        ks = [k1, k2, k1 ^ k2 ^ _U32(0x1BD11BDA)]
        rot = [(13, 15, 26, 6), (17, 29, 16, 24)]
        x0 = x0.astype(_U32) + ks[0]
        x1 = x1.astype(_U32) + ks[1]
        for i in range(5):
            for r in rot[i % 2]:
                x0 = x0 + x1
                x1 = _rotl(x1, r)
                x1 = x1 ^ x0
            x0 = x0 + ks[(i + 1) % 3]
            x1 = x1 + ks[(i + 2) % 3] + _U32(i + 1)
    return x0, x1


def prng_key(seed):
    return _U32((seed >> 32) & 0xFFFFFFFF), _U32(seed & 0xFFFFFFFF)


def random_bits(key, shape):
    """32 random bits per element, counter layout of the original (non-partitionable) threefry path."""
    size = int(np.prod(shape))
    cnt = np.arange(size, dtype=_U32)
    if size % 2:
        cnt = np.concatenate([cnt, np.zeros(1, _U32)])
    half = cnt.size // 2
    a, b = threefry2x32(key[0], key[1], cnt[:half], cnt[half:])
    out = np.concatenate([a, b])[:size]
    return out.reshape(shape)


def uniform(key, shape, minval=0.0, maxval=1.0):
    """jax.random.uniform for float32: mantissa-fill in [1,2), minus 1, scale, clamp."""
    bits = random_bits(key, shape)
    f = ((bits >> _U32(9)) | _U32(0x3F800000)).view(np.float32) - np.float32(1.0)
    lo = np.broadcast_to(np.asarray(minval, np.float32), shape)
    hi = np.broadcast_to(np.asarray(maxval, np.float32), shape)
    return np.maximum(lo, (f * (hi - lo) + lo).astype(np.float32))


def normal(key, shape):
    """jax.random.normal for float32: sqrt(2) * erfinv(U(-1,1)); erfinv evaluated in
    float64 and rounded (XLA uses a float32 polynomial: agreement to ~1 ulp)."""
    lo = np.nextafter(np.float32(-1.0), np.float32(0.0))
    u = uniform(key, shape, lo, 1.0)
    return (np.float32(np.sqrt(2.0)) * erfinv(u.astype(np.float64)).astype(np.float32)).astype(np.float32)


def split(key, num=2):
    """jax.random.split for the default (non-partitionable) threefry keys: `num` new keys, row i =
    words (2i, 2i+1) of the key's bit stream of shape (num, 2)."""
    b = random_bits(key, (num, 2))
    return [(b[i, 0], b[i, 1]) for i in range(num)]


def randint_bits(key):
    """The two 32-bit words jax.random.randint(key, (), lo, hi) consumes: k1, k2 = split(key);
    higher = bits(k1), lower = bits(k2)."""
    k1, k2 = split(key, 2)
    return int(random_bits(k1, ())), int(random_bits(k2, ()))


def randint_from_bits(higher, lower, minval, maxval):
    """jax.random.randint's arithmetic for 32-bit integers (jax._src.random._randint): uint32 wrap-around,
    span = maxval - minval (1 if not positive), multiplier = (2^16 mod span)^2 mod span."""
    span = (maxval - minval) & 0xFFFFFFFF
    if maxval <= minval:
        span = 1
    mult = ((65536 % span) * (65536 % span)) & 0xFFFFFFFF
    mult %= span
    off = (((higher % span) * mult) & 0xFFFFFFFF) + (lower % span)
    off = (off & 0xFFFFFFFF) % span
    return minval + off
