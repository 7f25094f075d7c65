"""Sharded matmul primitives: ``dot`` (jax.lax.dot on committed sharded operands -- case1a.py:49,
case1b.py:47, case2.py:44, case3_fully_sharded.py:46, case4_gspmd_ff.py:52) and the batched
``einsum("ABC,ACD->ABD")`` of case4_gspmd_ff.py:30.

The reference leaves the choice of collectives to XLA's GSPMD partitioner.  Here the plan is
derived from the global slices each device holds:

  * the result keeps A's row blocks and B's column blocks: device d ends up with C[rows_A(d), cols_B(d)]
    (replicated C for case1a/1b, row-sharded for case2, fully sharded for case3/case4);
  * "partial-sum" plan when the devices that share an output block also partition K between them
    (A's K blocks): B is re-sliced to A's K blocks (local slicing, or an exchange when the two
    operands' K blocks sit on different devices as in case1a's ``reshape(4,2)``), every device
    multiplies its K block, and the partials are all-reduced inside each K-covering sub-group;
  * "gather" plan otherwise (case3: fully sharded operands): all-gather the contraction dim of
    both operands, then one local full-K product -- the FSDP pattern;
  * nothing to exchange when every device already holds full K on both operands (case4).
"""
from __future__ import annotations

from typing import Dict, List, Optional

import torch

from . import runtime as _rt
from .array import ShardedArray, reshard
from .sharding import SliceSharding, _slices_eq


def _key(sl):
    return tuple((s.start, s.stop) for s in sl)


def _local_matmul(a: torch.Tensor, b: torch.Tensor) -> torch.Tensor:
    comp = _rt.local_compute()
    if a.dtype == torch.float32:
        return comp.gemm_f32(a, b)
    if a.dtype == torch.bfloat16:
        return comp.gemm_bf16(a, b)
    raise TypeError(f"dot: unsupported dtype {a.dtype}")


def plan_dot(a: ShardedArray, b: ShardedArray) -> dict:
    """Pure planning (no data touched): returns the strategy, the per-device slices each operand
    must be brought to, the output index and the all-reduce groups.  Deterministic on every rank."""
    if a.ndim != 2 or b.ndim != 2 or a.shape[1] != b.shape[0]:
        raise ValueError(f"dot: shapes {a.shape} and {b.shape} do not contract")
    K = a.shape[1]
    ai, bi = a._index, b._index
    devices = sorted(set(ai) & set(bi))
    if not devices:
        raise ValueError("dot: operands share no device")
    out_index = {d: (ai[d][0], bi[d][1]) for d in devices}
    full_k = slice(0, K)
    # output replica groups
    groups: Dict[tuple, List[int]] = {}
    for d in devices:
        groups.setdefault(_key(out_index[d]), []).append(d)
    # can every output replica group cover K with A's K blocks, each block the same number of times?
    reduce_groups: List[List[int]] = []
    coverable = True
    for g in groups.values():
        by_block: Dict[tuple, List[int]] = {}
        for d in g:
            by_block.setdefault((ai[d][1].start, ai[d][1].stop), []).append(d)
        blocks = sorted(by_block)
        pos = 0
        for (lo, hi) in blocks:
            if lo != pos:
                coverable = False
            pos = hi
        if pos != K or len({len(v) for v in by_block.values()}) != 1:
            coverable = False
        if not coverable:
            break
        mult = len(next(iter(by_block.values())))
        for j in range(mult):
            reduce_groups.append([by_block[blk][j] for blk in blocks])
    if coverable:
        a_want = {d: ai[d] for d in devices}
        b_want = {d: (ai[d][1], bi[d][1]) for d in devices}
        strategy = "local" if all(len(g) == 1 for g in reduce_groups) else "partial_allreduce"
    else:
        a_want = {d: (ai[d][0], full_k) for d in devices}
        b_want = {d: (full_k, bi[d][1]) for d in devices}
        reduce_groups = [[d] for d in devices]
        strategy = "gather"
    return {"strategy": strategy, "a_want": a_want, "b_want": b_want, "out_index": out_index,
            "reduce_groups": reduce_groups}


def dot(a: ShardedArray, b: ShardedArray) -> ShardedArray:
    rt = _rt.get_runtime()
    plan = plan_dot(a, b)
    a_loc = reshard(a, SliceSharding(plan["a_want"]))
    b_loc = reshard(b, SliceSharding(plan["b_want"]))
    partial = {}
    for d in rt.local_device_ids:
        if d in plan["out_index"]:
            partial[d] = _local_matmul(a_loc.shard(d), b_loc.shard(d))
    if plan["strategy"] == "partial_allreduce":
        partial = rt.all_reduce(partial, plan["reduce_groups"])
    out_shape = (a.shape[0], b.shape[1])
    return ShardedArray(out_shape, a.dtype, SliceSharding(plan["out_index"]), partial)


def einsum(subscripts: str, a: ShardedArray, b: ShardedArray) -> ShardedArray:
    """``einsum("ABC,ACD->ABD", a, b)``: batched matmul on arrays every device holds whole
    (case4_gspmd_ff.py:17-33 runs it unsharded)."""
    spec = subscripts.replace(" ", "")
    lhs, out = spec.split("->")
    x, y = lhs.split(",")
    batched = len(x) == 3 and len(y) == 3 and x[0] == y[0] == out[0] and x[2] == y[1] and out == x[0] + x[1] + y[2]
    plain = len(x) == 2 and len(y) == 2 and x[1] == y[0] and out == x[0] + y[1]
    if not (batched or plain):
        raise NotImplementedError(f"einsum pattern {subscripts!r} is not on the sharded hot path")
    if plain:
        return dot(a, b)
    rt = _rt.get_runtime()
    if not (a.sharding.is_fully_replicated(a.shape) and b.sharding.is_fully_replicated(b.shape)):
        raise NotImplementedError("batched einsum is implemented for replicated operands (as the reference runs it)")
    comp = _rt.local_compute()
    shards = {}
    for d in rt.local_device_ids:
        if d in a.local_shards and d in b.local_shards:
            shards[d] = comp.gemm_f32(a.shard(d), b.shard(d))
    out_shape = (a.shape[0], a.shape[1], b.shape[2])
    from .sharding import ReplicatedSharding
    return ShardedArray(out_shape, a.dtype, ReplicatedSharding(sorted(a._index)), shards)
