"""CPU restatement of jax.random as the reference scripts use it -- TEST INFRASTRUCTURE (see reference_math.py's header for
who may import this package).

Call sites: jax.random.PRNGKey(0) / jax.random.key(0) and jax.random.normal(key, shape) (case1a.py:16-18;
case4_gspmd_ff.py:26-28; case6_attention.py:147,152), the `{"params": key}` rng of model.init (case6_attention.py:171-180)
through which flax draws nn.initializers.lecun_normal() for every Dense kernel (case6_attention.py:57,86).

The algorithm lives in jax (not vendored under /root/reference, no version pinned; API usage brackets jax 0.4.16-0.4.2x,
where the default PRNG is threefry2x32 with jax_threefry_partitionable = False).  Restated from its published definition:

  threefry2x32      Salmon et al., "Parallel random numbers: as easy as 1, 2, 3" (SC'11), 20 rounds, rotations
                    (13, 15, 26, 6) / (17, 29, 16, 24), key-schedule constant 0x1BD11BDA -- jax/_src/prng.py
  random_bits       counts = iota(n), padded to even length and split in HALVES: x0 = counts[:h], x1 = counts[h:],
                    (y0, y1) = threefry(key, x0, x1), bits = concat(y0, y1)[:n]
  split / fold_in   split(key, num) = random_bits(key, 2 num) reshaped (num, 2); fold_in(key, d) = threefry(key, (0, d))
  uniform           float_bits = bits >> 9 | 0x3F800000; u = bitcast(float_bits) - 1; max(lo, u * (hi - lo) + lo)
  normal            sqrt(2) * erf_inv(uniform(lo = nextafter(-1, 0), hi = 1)), erf_inv = Giles' single-precision polynomial
                    as XLA's ErfInv32 evaluates it
  truncated_normal  sqrt(2) * erf_inv(uniform(erf(-2 / sqrt 2), erf(2 / sqrt 2))), clipped to the open interval
  lecun_normal      variance_scaling(1.0, "fan_in", "truncated_normal"): truncated_normal * sqrt(1 / fan_in) / 0.87962566103423978

PINS (tests/test_prng.py).  The integer stages are bit-exact and anchored on published known answers that are NOT this
repo's own output: the three threefry2x32 vectors of the Random123 distribution (also asserted by jax's own
tests/random_test.py::testThreefry2x32), the documented value of jax.random.split(PRNGKey(0)), and the first documented
draw jax.random.normal(PRNGKey(0), (1,)) = -0.20584226.  The float stages follow the formulas above in f32; XLA may
differ from them by an ulp of its log1p / polynomial evaluation, so floats are compared at 4 ulp, not bit for bit.
"""
from __future__ import annotations

import hashlib
import math

import numpy as np

U32 = np.uint32
_ROT = ((13, 15, 26, 6), (17, 29, 16, 24))


def _rotl(x, r):
    return (x << U32(r)) | (x >> U32(32 - r))


def threefry2x32(key, x0, x1):
    """threefry2x32_p: key = (k0, k1) uint32, x0 / x1 uint32 arrays of equal shape -> (y0, y1)."""
    with np.errstate(over="ignore"):
        k0, k1 = U32(key[0]), U32(key[1])
        ks = (k0, k1, U32(k0 ^ k1 ^ U32(0x1BD11BDA)))
        x0 = (np.asarray(x0, dtype=U32) + ks[0]).astype(U32)
        x1 = (np.asarray(x1, dtype=U32) + ks[1]).astype(U32)
        for block in range(5):
            for r in _ROT[block % 2]:
                x0 = (x0 + x1).astype(U32)
                x1 = _rotl(x1, r).astype(U32)
                x1 = x1 ^ x0
            x0 = (x0 + ks[(block + 1) % 3]).astype(U32)
            x1 = (x1 + ks[(block + 2) % 3] + U32(block + 1)).astype(U32)
    return x0, x1


def PRNGKey(seed: int):
    """jax.random.PRNGKey(seed) / jax.random.key(seed) for a 32-bit-representable seed: (hi, lo) = (0, seed)."""
    return np.array([(seed >> 32) & 0xFFFFFFFF, seed & 0xFFFFFFFF], dtype=U32)


def random_bits(key, n: int):
    """_threefry_random_bits_original for bit_width 32: n uint32 words."""
    if n == 0:
        return np.zeros((0,), dtype=U32)
    counts = np.arange(n, dtype=U32)
    if n % 2:
        counts = np.concatenate([counts, np.zeros((1,), dtype=U32)])
    h = counts.size // 2
    y0, y1 = threefry2x32(key, counts[:h], counts[h:])
    return np.concatenate([y0, y1])[:n]


def split(key, num: int = 2):
    return random_bits(key, 2 * num).reshape(num, 2)


def fold_in(key, data: int):
    y0, y1 = threefry2x32(key, np.array([0], dtype=U32), np.array([data & 0xFFFFFFFF], dtype=U32))
    return np.array([y0[0], y1[0]], dtype=U32)


def _uniform_from_bits(bits, lo, hi):
    lo, hi = np.float32(lo), np.float32(hi)
    f = ((bits >> U32(9)) | U32(0x3F800000)).view(np.float32) - np.float32(1.0)
    return np.maximum(lo, f * np.float32(hi - lo) + lo).astype(np.float32)


def uniform(key, shape, lo=0.0, hi=1.0):
    n = int(np.prod(shape)) if len(shape) else 1
    return _uniform_from_bits(random_bits(key, n), lo, hi).reshape(shape)


def erf_inv_f32(x):
    """XLA ErfInv32 (Giles, "Approximating the erfinv function"): f32 throughout."""
    x = np.asarray(x, dtype=np.float32)
    w = (-np.log1p((-x * x).astype(np.float32))).astype(np.float32)
    lt = w < np.float32(5.0)
    w1 = (w - np.float32(2.5)).astype(np.float32)
    w2 = (np.sqrt(w) - np.float32(3.0)).astype(np.float32)
    ww = np.where(lt, w1, w2).astype(np.float32)
    c_lt = (2.81022636e-08, 3.43273939e-07, -3.5233877e-06, -4.39150654e-06, 0.00021858087, -0.00125372503,
            -0.00417768164, 0.246640727, 1.50140941)
    c_ge = (-0.000200214257, 0.000100950558, 0.00134934322, -0.00367342844, 0.00573950773, -0.0076224613,
            0.00943887047, 1.00167406, 2.83297682)
    p = np.where(lt, np.float32(c_lt[0]), np.float32(c_ge[0])).astype(np.float32)
    for a, b in zip(c_lt[1:], c_ge[1:]):
        p = (np.where(lt, np.float32(a), np.float32(b)) + p * ww).astype(np.float32)
    out = (p * x).astype(np.float32)
    return np.where(np.abs(x) == 1, np.float32(np.inf) * x, out).astype(np.float32)


def normal(key, shape):
    """jax.random.normal(key, shape) in float32 (case1a.py:17-18; case6_attention.py:152)."""
    lo = np.nextafter(np.float32(-1.0), np.float32(0.0))
    u = uniform(key, shape, lo, 1.0)
    return (np.float32(math.sqrt(2.0)) * erf_inv_f32(u)).astype(np.float32)


def truncated_normal(key, shape, lower=-2.0, upper=2.0):
    s2 = np.float32(math.sqrt(2.0))
    a = np.float32(math.erf(np.float32(lower) / s2))
    b = np.float32(math.erf(np.float32(upper) / s2))
    u = uniform(key, shape, a, b)
    out = (s2 * erf_inv_f32(u)).astype(np.float32)
    lo = np.nextafter(np.float32(lower), np.float32(np.inf))
    hi = np.nextafter(np.float32(upper), np.float32(-np.inf))
    return np.clip(out, lo, hi).astype(np.float32)


LECUN_TRUNC_STD = 0.87962566103423978      # std of the unit normal truncated to (-2, 2)


def lecun_normal(key, shape, fan_in=None):
    """nn.initializers.lecun_normal() for a Dense kernel (in_features, out_features): fan_in = shape[0]."""
    fan_in = shape[0] if fan_in is None else fan_in
    stddev = np.float32(math.sqrt(1.0 / fan_in) / LECUN_TRUNC_STD)
    return (truncated_normal(key, shape) * stddev).astype(np.float32)


def path_key(key, path: str):
    """Per-parameter key: the module key folded with the first four bytes of sha1(path), the way flax folds static
    names into an rng (flax.core.scope._fold_in_static).  The exact chain flax walks from `{"params": key}` to a leaf
    depends on its version; this fixes ONE documented derivation so that the library and the oracle agree."""
    digest = hashlib.sha1(path.encode("utf-8")).digest()
    return fold_in(key, int.from_bytes(digest[:4], "big"))
