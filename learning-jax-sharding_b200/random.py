"""jax.random as the reference scripts use it, with values BORN SHARDED: every device fills its own slice of the array
in its own HBM (``b200_random_fill``), so a parameter tree initialised under ``out_shardings`` never exists unsharded.

Reference call sites: ``jax.random.PRNGKey(0)`` / ``jax.random.key(0)`` and ``jax.random.normal(key, shape)``
(case1a.py:16-18; case4_gspmd_ff.py:26-28; case6_attention.py:147,152); the ``{"params": key}`` rng of ``model.init``
under ``jit(..., out_shardings=logical_state_sharding)`` (case6_attention.py:171-196) through which flax draws
``nn.initializers.lecun_normal()`` (case6_attention.py:57,86).

Keys are jax threefry keys (two uint32 words).  ``split`` and ``fold_in`` touch a handful of words and run on the host in
plain Python integers; the arrays themselves are generated on the devices.  The bit stream is jax's (threefry2x32,
``jax_threefry_partitionable = False`` counter layout), checked against published known answers in tests/test_prng.py.
"""
from __future__ import annotations

import hashlib
import math
from typing import Optional, Sequence, Tuple

import torch

from . import runtime as _rt
from .array import ShardedArray
from .sharding import ReplicatedSharding, Sharding

_M32 = 0xFFFFFFFF
_ROTATIONS = ((13, 15, 26, 6), (17, 29, 16, 24))
LECUN_TRUNCATED_STD = 0.87962566103423978     # std of a unit normal truncated to (-2, 2)

Key = Tuple[int, int]


def _threefry(key: Key, x0: int, x1: int) -> Tuple[int, int]:
    """One threefry2x32 block (20 rounds) on Python integers."""
    ks = (key[0], key[1], key[0] ^ key[1] ^ 0x1BD11BDA)
    x0 = (x0 + ks[0]) & _M32
    x1 = (x1 + ks[1]) & _M32
    for block in range(5):
        for r in _ROTATIONS[block % 2]:
            x0 = (x0 + x1) & _M32
            x1 = (((x1 << r) | (x1 >> (32 - r))) & _M32) ^ x0
        x0 = (x0 + ks[(block + 1) % 3]) & _M32
        x1 = (x1 + ks[(block + 2) % 3] + block + 1) & _M32
    return x0, x1


def PRNGKey(seed: int) -> Key:
    """``jax.random.PRNGKey(seed)``: the key words (seed >> 32, seed & 0xFFFFFFFF)."""
    return ((seed >> 32) & _M32, seed & _M32)


key = PRNGKey          # jax.random.key(seed) (case6_attention.py:147): same key data


def _words(k: Key, n: int):
    """The first n words of jax's bit stream for a flat array of n words (host side, small n only)."""
    h = (n + 1) // 2
    lo, hi = [], []
    for j in range(h):
        y0, y1 = _threefry(k, j, h + j if h + j < n else 0)
        lo.append(y0)
        hi.append(y1)
    return (lo + hi)[:n]


def split(k: Key, num: int = 2):
    """``jax.random.split(key, num)``: ``num`` new keys."""
    w = _words(k, 2 * num)
    return [(w[2 * i], w[2 * i + 1]) for i in range(num)]


def fold_in(k: Key, data: int) -> Key:
    """``jax.random.fold_in(key, data)``."""
    return _threefry(k, 0, data & _M32)


def path_key(k: Key, path: str) -> Key:
    """Key of one parameter: the module key folded with the first four bytes of sha1(path) -- how flax folds static names
    into an rng.  (The exact chain flax walks from ``{"params": key}`` to a leaf depends on the flax version; this is the
    one documented derivation this library and its oracle share.)"""
    return fold_in(k, int.from_bytes(hashlib.sha1(path.encode("utf-8")).digest()[:4], "big"))


def _fill(k: Key, shape: Sequence[int], sharding: Optional[Sharding], distribution: int, lo=0.0, hi=0.0, scale=1.0,
          dtype=torch.float32) -> ShardedArray:
    from . import capi
    rt = _rt.get_runtime()
    shape = tuple(int(s) for s in shape)
    if sharding is None:
        sharding = ReplicatedSharding([dv.id for dv in rt.devices])
    index = sharding.devices_indices_map(shape)
    shards = {}
    for d in rt.local_device_ids:
        if d not in index:
            continue
        sl = index[d]
        out = torch.empty(tuple(s.stop - s.start for s in sl), dtype=dtype, device=rt.buffer_device(d))
        capi.random_fill(out, k, shape, tuple(s.start for s in sl), distribution, lo=lo, hi=hi, scale=scale)
        shards[d] = out
    return ShardedArray(shape, dtype, sharding, shards)


def bits(k: Key, shape, sharding: Optional[Sharding] = None) -> ShardedArray:
    """``jax.random.bits(key, shape)`` (uint32), held as int32 bit patterns."""
    from . import capi
    return _fill(k, shape, sharding, capi.RNG_BITS, dtype=torch.int32)


def uniform(k: Key, shape, minval: float = 0.0, maxval: float = 1.0, sharding: Optional[Sharding] = None) -> ShardedArray:
    from . import capi
    return _fill(k, shape, sharding, capi.RNG_UNIFORM, lo=minval, hi=maxval)


def normal(k: Key, shape, sharding: Optional[Sharding] = None) -> ShardedArray:
    """``jax.random.normal(key, shape)`` in float32; with ``sharding`` each device generates only its slice."""
    from . import capi
    return _fill(k, shape, sharding, capi.RNG_NORMAL)


def truncated_normal(k: Key, lower: float, upper: float, shape, sharding: Optional[Sharding] = None,
                     scale: float = 1.0) -> ShardedArray:
    from . import capi
    return _fill(k, shape, sharding, capi.RNG_TRUNCATED_NORMAL, lo=lower, hi=upper, scale=scale)


def lecun_normal(k: Key, shape, sharding: Optional[Sharding] = None, fan_in: Optional[int] = None) -> ShardedArray:
    """``nn.initializers.lecun_normal()(key, shape)`` for a Dense kernel (in_features, out_features):
    variance_scaling(1.0, "fan_in", "truncated_normal")."""
    fan_in = shape[0] if fan_in is None else fan_in
    return truncated_normal(k, -2.0, 2.0, shape, sharding, scale=math.sqrt(1.0 / fan_in) / LECUN_TRUNCATED_STD)
