"""b200 sharded layers: the data-parallel / tensor-parallel hot path of the
entrpn/learning-jax-sharding case scripts (sharded matmul, Dense, FFN, multi-head attention)
as hand-written sm_100a kernels behind a C ABI, with the host-side mesh plumbing the scripts use.

The directory is named ``learning-jax-sharding_b200``; import it as ``learning_jax_sharding_b200``
(the repo-root shim module of that name loads this package).

    import learning_jax_sharding_b200 as ljs
    ljs.init_virtual(8)                                   # or ljs.init_distributed() under torchrun
    sharding = ljs.PositionalSharding(ljs.create_device_mesh((2, 4)))
    A = ljs.device_put(A, sharding.replicate(axis=0, keepdims=True))
    C = ljs.lax.dot(A, B)
"""
from . import capi, debug, lax, nn, random, runtime, train  # noqa: F401
from .array import ShardedArray, device_put, reshard  # noqa: F401
from .runtime import get_runtime, init_distributed, init_virtual  # noqa: F401
from .sharding import (Device, Mesh, NamedSharding, PartitionSpec, PositionalSharding,  # noqa: F401
                       ReplicatedSharding, SliceSharding, create_device_mesh, current_mesh)

__all__ = [
    "capi", "debug", "lax", "nn", "random", "runtime", "train", "ShardedArray", "device_put", "reshard", "get_runtime", "init_distributed",
    "init_virtual", "Device", "Mesh", "NamedSharding", "PartitionSpec", "PositionalSharding", "ReplicatedSharding",
    "SliceSharding", "create_device_mesh", "current_mesh",
]
