"""Sharded arrays: placement (``device_put``), per-device access (``.device_buffers[i]``,
``.addressable_shards[i].data``) and resharding.

Reference call sites: ``jax.device_put(A, sharding...)`` case1a.py:24,30, case6_attention.py:162;
``A.device_buffers[0]`` case1a.py:35-55; ``C.addressable_shards[0].data`` case4_gspmd_ff.py:56.
"""
from __future__ import annotations

from dataclasses import dataclass
from typing import Dict, Optional, Tuple

import numpy as np
import torch

from . import runtime as _rt
from .sharding import NamedSharding, ReplicatedSharding, Sharding, _slices_eq


@dataclass
class Shard:
    device: object
    index: Tuple[slice, ...]
    data: torch.Tensor


class ShardedArray:
    """A global array whose pieces live on the devices of ``sharding``.  This process holds the
    buffers of its addressable devices only (all of them in a virtual runtime, exactly one in a
    one-process-per-GPU job -- the same rule ``jax.Array.addressable_shards`` follows)."""

    def __init__(self, shape, dtype, sharding: Sharding, shards: Dict[int, torch.Tensor]):
        self.shape = tuple(int(s) for s in shape)
        self.dtype = dtype
        self.sharding = sharding
        self._shards = dict(sorted(shards.items()))
        self._index = sharding.devices_indices_map(self.shape)

    @property
    def ndim(self):
        return len(self.shape)

    def index_of(self, device_id: int):
        return self._index[device_id]

    def shard(self, device_id: int) -> torch.Tensor:
        return self._shards[device_id]

    @property
    def local_shards(self) -> Dict[int, torch.Tensor]:
        return self._shards

    @property
    def addressable_shards(self):
        rt = _rt.get_runtime()
        return [Shard(rt.devices[d], self._index[d], t) for d, t in self._shards.items()]

    @property
    def device_buffers(self):
        """Per-device buffers in device order (``A.device_buffers[i]``, case1a.py:35)."""
        return [t for _, t in sorted(self._shards.items())]

    def global_array(self) -> torch.Tensor:
        """Assemble the global value on the host (needs every piece to be addressable, or, in a
        distributed job, gathers over the job)."""
        rt = _rt.get_runtime()
        full = tuple(slice(0, d) for d in self.shape)
        out = rt.exchange(self._shards, self._index, {d: full for d in rt.local_device_ids}, self.shape, self.dtype)
        return out[rt.local_device_ids[0]].cpu()

    def __array__(self, dtype=None):
        a = self.global_array()
        a = a.float().numpy() if a.dtype == torch.bfloat16 else a.numpy()
        return a.astype(dtype) if dtype is not None else a

    def __repr__(self):
        return f"ShardedArray(shape={self.shape}, dtype={self.dtype}, sharding={self.sharding})"


def _as_tensor(x) -> torch.Tensor:
    if isinstance(x, torch.Tensor):
        return x
    return torch.from_numpy(np.asarray(x))


def device_put(x, sharding: Optional[Sharding] = None) -> ShardedArray:
    """Slice a host (or device) array by ``sharding`` and copy each slice to its device(s)
    (host->device copies; outside any timed region, like the reference's placement)."""
    rt = _rt.get_runtime()
    if isinstance(x, ShardedArray):
        return reshard(x, sharding) if sharding is not None else x
    t = _as_tensor(x)
    if sharding is None:
        sharding = ReplicatedSharding([d.id for d in rt.devices])
    index = sharding.devices_indices_map(tuple(t.shape))
    shards = {}
    for d in rt.local_device_ids:
        if d in index:
            shards[d] = t[index[d]].to(rt.buffer_device(d)).contiguous().clone()
    return ShardedArray(t.shape, t.dtype, sharding, shards)


def reshard(x: ShardedArray, sharding: Sharding) -> ShardedArray:
    """Move ``x`` to another sharding.  Recognises the mesh-axis patterns that have a dedicated
    collective (all-gather along one dim over mesh axes, local slicing of replicated data,
    all-to-all between two dims); anything else goes through the generic exchange."""
    rt = _rt.get_runtime()
    if x.sharding.same_layout(sharding, x.shape):
        return ShardedArray(x.shape, x.dtype, sharding, x.local_shards)
    new_index = sharding.devices_indices_map(x.shape)
    # (1) every device already holds a superset of what it needs: local slicing, no communication
    local_only = True
    for d in rt.local_device_ids:
        if d in new_index:
            if d not in x._index or any(n.start < o.start or n.stop > o.stop for n, o in zip(new_index[d], x._index[d])):
                local_only = False
    if rt.kind == "distributed":
        local_only = _all_ranks_agree(rt, local_only)
    if local_only:
        shards = {}
        for d in rt.local_device_ids:
            if d in new_index:
                o = x._index[d]
                sl = tuple(slice(n.start - oo.start, n.stop - oo.start) for n, oo in zip(new_index[d], o))
                shards[d] = x.shard(d)[sl].contiguous()
        return ShardedArray(x.shape, x.dtype, sharding, shards)
    # (2) named shardings on one mesh that differ in exactly one dim by trailing axes -> all-gather
    if isinstance(x.sharding, NamedSharding) and isinstance(sharding, NamedSharding) and x.sharding.mesh is sharding.mesh:
        plan = _named_plan(x, sharding)
        if plan is not None:
            return plan
    # (3) generic
    wanted = {d: new_index[d] for d in rt.local_device_ids if d in new_index}
    shards = rt.exchange(x.local_shards, x._index, wanted, x.shape, x.dtype)
    return ShardedArray(x.shape, x.dtype, sharding, shards)


def _all_ranks_agree(rt, flag: bool) -> bool:
    t = torch.tensor([1 if flag else 0], dtype=torch.int32, device=rt.torch_device)
    rt.dist.all_reduce(t, op=rt.dist.ReduceOp.MIN)
    return bool(t.item())


def _named_plan(x: ShardedArray, target: NamedSharding) -> Optional[ShardedArray]:
    rt = _rt.get_runtime()
    mesh = target.mesh
    src_axes = x.sharding.dim_axes(x.ndim)
    dst_axes = target.dim_axes(x.ndim)
    diff = [i for i in range(x.ndim) if src_axes[i] != dst_axes[i]]
    if len(diff) == 1:
        i = diff[0]
        s, t = src_axes[i], dst_axes[i]
        # all-gather: target drops trailing axes of the source on this dim
        if len(t) < len(s) and s[: len(t)] == t:
            gather_axes = s[len(t):]
            out = rt.all_gather(x.local_shards, mesh.axis_groups(gather_axes), dim=i)
            return ShardedArray(x.shape, x.dtype, target, out)
    if len(diff) == 2:
        i, j = diff
        # all-to-all: one axis moves from dim j to dim i (or i to j), other axes untouched
        for a, b in ((i, j), (j, i)):
            sa, sb, ta, tb = src_axes[a], src_axes[b], dst_axes[a], dst_axes[b]
            if len(sb) == len(tb) + 1 and sb[: len(tb)] == tb and len(ta) == len(sa) + 1 and ta[: len(sa)] == sa \
                    and ta[-1] == sb[-1]:
                axis = sb[-1]
                out = rt.all_to_all(x.local_shards, mesh.axis_groups((axis,)), split_dim=a, concat_dim=b)
                return ShardedArray(x.shape, x.dtype, target, out)
    return None
