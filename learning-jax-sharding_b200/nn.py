"""Sharded layers behind the Flax surface the reference scripts use: logical-axis rules,
``Dense``, softmax attention and the feed-forward block.

Reference call sites:
  nn.Dense(.., dtype=bf16)                         case6_attention.py:61-90,96-99,139; case5_attention_dense.py:59-67
  nn.with_logical_partitioning / with_logical_constraint / axis_rules / logical_to_mesh_sharding
                                                   case6_attention.py:56-59,85-88,105-107,114-116,137,141,183-191
  attention core (einsum, softmax, einsum)         case6_attention.py:121-135
  FFN  y = act(x.Win).Wout (comment only)          case6_attention.py:36-40

Every layer picks its collectives from the operands' PartitionSpecs (what GSPMD does for the
reference) and runs the per-shard compute through the C ABI (``runtime.local_compute()``).
"""
from __future__ import annotations

import contextlib
import math
import threading
from typing import Dict, Optional, Sequence, Tuple

import torch

from . import runtime as _rt
from .array import ShardedArray, device_put, reshard
from .sharding import Mesh, NamedSharding, PartitionSpec, current_mesh

# ------------------------------------------------------------------ logical axis rules
_rules = threading.local()


@contextlib.contextmanager
def axis_rules(rules: Sequence[Tuple[str, Optional[str]]]):
    """``with nn_partitioning.axis_rules(rules):`` (case6_attention.py:234)."""
    stack = getattr(_rules, "stack", None)
    if stack is None:
        stack = _rules.stack = []
    stack.append(tuple(rules))
    try:
        yield
    finally:
        stack.pop()


def current_rules():
    stack = getattr(_rules, "stack", None)
    if not stack:
        raise RuntimeError("no logical axis rules: use `with axis_rules(rules):`")
    return stack[-1]


def logical_to_mesh_axes(names: Sequence[Optional[str]], rules=None) -> PartitionSpec:
    """Flax ``logical_to_mesh_axes``: rules are tried in order, a mesh axis is used at most once
    per array, a logical name with no applicable rule is left unsharded, ``None`` is replicated."""
    rules = current_rules() if rules is None else rules
    used = set()
    out = []
    for name in names:
        axis = None
        if name is not None:
            for lname, maxis in rules:
                if lname == name and maxis is not None and maxis not in used:
                    axis = maxis
                    break
        if axis is not None:
            used.add(axis)
        out.append(axis)
    return PartitionSpec(*out)


def logical_to_mesh_sharding(names, mesh: Mesh, rules):
    """``nn.logical_to_mesh_sharding``: a tuple of logical names -> NamedSharding; a tree (dict / tuple of such tuples,
    as ``get_partition_spec`` returns for a state) -> the same tree of NamedShardings (case6_attention.py:189-191)."""
    if isinstance(names, dict):
        return {k: logical_to_mesh_sharding(v, mesh, rules) for k, v in names.items()}
    if isinstance(names, (tuple, list)) and names and isinstance(names[0], (tuple, list)):
        return type(names)(logical_to_mesh_sharding(v, mesh, rules) for v in names)
    return NamedSharding(mesh, logical_to_mesh_axes(names, rules))


def get_partition_spec(model):
    """``nn.get_partition_spec(abstract_state)`` for this library's modules: the tree of logical axis names the module
    attached to its parameters with ``with_logical_partitioning`` (case6_attention.py:56-59,85-88,189).  Unannotated
    leaves (the bias) are replicated: an empty tuple."""
    tree = {}
    for name, axes in model.param_axes.items():
        tree[name] = (tuple(axes), ()) if name in getattr(model, "biased", ()) else tuple(axes)
    return tree


def with_logical_constraint(x: ShardedArray, names: Sequence[Optional[str]]) -> ShardedArray:
    """Reshard ``x`` to the mesh sharding its logical names translate to (the resharding points of
    case6_attention.py:105-107,114-116,137,141)."""
    mesh = current_mesh()
    if mesh is None:
        raise RuntimeError("with_logical_constraint needs an active mesh (`with mesh:`)")
    return reshard(x, NamedSharding(mesh, logical_to_mesh_axes(names)))


# ----------------------------------------------------------------------------- helpers
def _axes(x: ShardedArray):
    if not isinstance(x.sharding, NamedSharding):
        raise TypeError("sharded layers take NamedSharding operands")
    return x.sharding.dim_axes(x.ndim)


def _spec(axes_per_dim) -> PartitionSpec:
    return PartitionSpec(*[None if not a else (a[0] if len(a) == 1 else tuple(a)) for a in axes_per_dim])


def _to_bf16(t: torch.Tensor) -> torch.Tensor:
    if t.dtype == torch.bfloat16:
        return t
    return _rt.local_compute().cast_f32_to_bf16(t)


def _param_bf16(p: ShardedArray) -> ShardedArray:
    """The bf16 image of an f32 PARAMETER array, shard by shard, made once and kept on the shard tensors themselves (flax casts
    its f32 parameters on every nn.Dense call; a resident layer need not).  Casting BEFORE any resharding also halves the bytes
    an FSDP-style weight gather moves.  `invalidate_param_casts` drops the copies of parameters that were updated in place
    (TrainState.apply_gradients calls it)."""
    if p.dtype == torch.bfloat16:
        return p
    shards = {}
    for d, t in p.local_shards.items():
        c = t.__dict__.get("_b200_bf16")
        if c is None:
            c = _to_bf16(t)
            t.__dict__["_b200_bf16"] = c
        shards[d] = c
    return ShardedArray(p.shape, torch.bfloat16, p.sharding, shards)


def invalidate_param_casts(*params):
    for p in params:
        if isinstance(p, ShardedArray):
            for t in p.local_shards.values():
                t.__dict__.pop("_b200_bf16", None)


# ------------------------------------------------------------------------------- Dense
def dense(x: ShardedArray, kernel: ShardedArray, bias: Optional[ShardedArray] = None, *, epilogue: str = "none",
          scatter_dim: Optional[int] = None) -> ShardedArray:
    """``y = bf16(x) @ bf16(kernel) (+ bf16(bias))`` on sharded operands.  ``epilogue`` in
    {"none", "relu", "gelu_tanh"} is fused into the GEMM.  Plans (m = a mesh axis):

      kernel dim0 over m, x features unsharded   -> all-gather kernel over m   (FSDP; case5, case6 q/k/v)
      x features and kernel dim0 both over m     -> local partial products (f32) + all-reduce over m
                                                    (row-parallel; FFN W2, head-sharded out-proj);
                                                    with ``scatter_dim`` a reduce-scatter that leaves
                                                    that row dim of y sharded over m instead
      kernel dim1 over m                         -> column-parallel, y features sharded over m
                                                    (FFN W1; case6 to_out_0), no communication
    """
    rt = _rt.get_runtime()
    comp = _rt.local_compute()
    mesh = x.sharding.mesh
    kernel = _param_bf16(kernel)                  # cast once per parameter (and before any gather: half the bytes)
    if bias is not None:
        bias = _param_bf16(bias)
    xa = _axes(x)
    kw, nw = _axes(kernel)
    rows, kx = xa[:-1], xa[-1]
    row_axes = {a for r in rows for a in r}
    # y's feature dim cannot reuse a mesh axis the rows of x already use: gather the kernel there
    # ... nor the axis x's own feature dim is split over (a PartitionSpec may name a mesh axis once)
    nw_eff = tuple(a for a in nw if a not in row_axes and a not in kx)
    if nw_eff != nw:
        kernel = reshard(kernel, NamedSharding(mesh, _spec([kw, nw_eff])))
    if kx and kw and kx != kw:
        x = reshard(x, NamedSharding(mesh, _spec(list(rows) + [()])))
        kx = ()
    if kw and not kx:
        kernel = reshard(kernel, NamedSharding(mesh, _spec([(), nw_eff])))      # FSDP all-gather
        kw = ()
    if kx and not kw:
        kernel = reshard(kernel, NamedSharding(mesh, _spec([kx, nw_eff])))      # local row slicing
        kw = kx
    if bias is not None and nw_eff:
        bias = reshard(bias, NamedSharding(mesh, _spec([nw_eff])))
    elif bias is not None and not bias.sharding.is_fully_replicated(bias.shape):
        bias = reshard(bias, NamedSharding(mesh, PartitionSpec()))

    reduce_axes = kx                      # == kw here; () when the contraction is local
    epi_code = {"none": 0, "relu": 1, "gelu_tanh": 2}[epilogue]
    from . import capi as C
    out_local = {}
    for d in rt.local_device_ids:
        xs = x.shard(d)
        lead = xs.shape[:-1]
        x2 = _to_bf16(xs.reshape(-1, xs.shape[-1]))
        w = kernel.shard(d)
        b = bias.shard(d) if bias is not None else None
        if reduce_axes:
            y2 = comp.gemm_bf16(x2, w, out_dtype=torch.float32)          # raw f32 partials
        else:
            if b is not None:
                epi = (C.EPI_BIAS, C.EPI_BIAS_RELU, C.EPI_BIAS_GELU_TANH)[epi_code]
            else:
                epi = (C.EPI_NONE, C.EPI_RELU, C.EPI_GELU_TANH)[epi_code]
            y2 = comp.gemm_bf16(x2, w, bias=b, epilogue=epi)
        out_local[d] = y2.reshape(*lead, y2.shape[-1])

    out_rows = [tuple(r) for r in rows]
    if reduce_axes:
        if epi_code != 0:
            raise NotImplementedError("an activation cannot be fused into a row-parallel partial product")
        groups = mesh.axis_groups(reduce_axes)
        if scatter_dim is not None:
            out_local = rt.reduce_scatter(out_local, groups, dim=scatter_dim)
            out_rows[scatter_dim] = out_rows[scatter_dim] + tuple(reduce_axes)
        else:
            out_local = rt.all_reduce(out_local, groups)
        for d in list(out_local):
            t = out_local[d]
            b = _to_bf16(bias.shard(d)) if bias is not None else None
            y2 = comp.sum_partials_to_bf16(t.reshape(1, -1, t.shape[-1]).contiguous(), bias=b)
            out_local[d] = y2.reshape(t.shape)
    out_shape = x.shape[:-1] + (kernel.shape[1],)
    return ShardedArray(out_shape, torch.bfloat16, NamedSharding(mesh, _spec(out_rows + [nw_eff])), out_local)


# ------------------------------------------------------------------------------ attention
def split_heads(x: ShardedArray, heads: int) -> ShardedArray:
    """(b, S, N*D) -> (b, S, N, D); a view on every shard (case6_attention.py:109-112).  A mesh axis on
    the feature dim moves to the heads dim."""
    xa = _axes(x)
    d_head = x.shape[-1] // heads
    nparts = 1
    for a in xa[-1]:
        nparts *= x.sharding.mesh.axis_size(a)
    if heads % nparts:
        raise ValueError(f"{heads} heads cannot be split over {nparts} devices")
    shards = {d: t.reshape(*t.shape[:-1], heads // nparts, d_head) for d, t in x.local_shards.items()}
    spec = _spec(list(xa[:-1]) + [xa[-1], ()])
    return ShardedArray(x.shape[:-1] + (heads, d_head), x.dtype, NamedSharding(x.sharding.mesh, spec), shards)


def merge_heads(x: ShardedArray) -> ShardedArray:
    """(b, S, N, D) -> (b, S, N*D) (case6_attention.py:135)."""
    xa = _axes(x)
    if xa[-1]:
        raise ValueError("the head dim (D) must not be sharded")
    shards = {d: t.reshape(*t.shape[:-2], t.shape[-2] * t.shape[-1]) for d, t in x.local_shards.items()}
    spec = _spec(list(xa[:-2]) + [xa[-2]])
    return ShardedArray(x.shape[:-2] + (x.shape[-2] * x.shape[-1],), x.dtype, NamedSharding(x.sharding.mesh, spec),
                        shards)


def _ring_attention(q: ShardedArray, k: ShardedArray, v: ShardedArray, scale, causal: bool) -> ShardedArray:
    """Blockwise attention for q, k, v all split along S over the same mesh axes (the literal case6 layout,
    case6_attention.py:105-107,114-116): K / V are never gathered.  Every device attends its query block to one key/value
    block at a time -- its own first, then the blocks that travel round the ring of the S-sharding group, the transfer of
    the next block running under the attention of the current one -- and merges the per-block results through their
    log-sum-exp (b200_attention_fwd_lse + b200_attention_combine).  With `causal`, blocks that lie entirely in the future
    of the query block are skipped and only the diagonal block is masked."""
    rt = _rt.get_runtime()
    comp = _rt.local_compute()
    mesh = q.sharding.mesh
    groups = mesh.axis_groups(_axes(k)[1])
    pos = {d: i for g in groups for i, d in enumerate(g)}
    W = len(groups[0])
    cur = {d: torch.stack([k.shard(d), v.shard(d)]) for d in rt.local_device_ids}
    acc, lse, started = {}, {}, {}
    for t in range(W):
        nxt, pending = rt.ring_shift_start(cur, groups) if t + 1 < W else (None, [])
        for d in rt.local_device_ids:
            src = (pos[d] - t) % W                      # whose block this device holds at step t
            if causal and src > pos[d]:
                continue                                # every key of that block is later than every query here
            ql = q.shard(d)
            o_b, lse_b = comp.attention_fwd_lse(ql, cur[d][0], cur[d][1], scale=scale, causal=causal and src == pos[d])
            if d not in acc:
                acc[d] = torch.empty(ql.shape, dtype=torch.float32, device=ql.device)
                lse[d] = torch.empty_like(lse_b)
            comp.attention_combine(acc[d], lse[d], o_b.contiguous(), lse_b, first=d not in started)
            started[d] = True
        if nxt is not None:
            rt.ring_shift_finish(pending)
            cur = nxt
    out = {d: comp.cast_f32_to_bf16(acc[d]) for d in rt.local_device_ids}
    return ShardedArray(q.shape, torch.bfloat16, q.sharding, out)


def dot_product_attention(q: ShardedArray, k: ShardedArray, v: ShardedArray, *, scale: Optional[float] = None,
                          causal: bool = False, ring: Optional[bool] = None) -> ShardedArray:
    """softmax(scale q k^T) v on (b, S, N, D) arrays (case6_attention.py:121-135).  Batch and head
    dims may be sharded (no communication: head-sharded attention).  A sequence-sharded q keeps its query block; when k
    and v are split along S the same way (the literal case6 layout) the key/value blocks go round a ring and are never
    gathered (``ring``: None = that choice, False = all-gather k and v along S instead, what XLA would do)."""
    rt = _rt.get_runtime()
    comp = _rt.local_compute()
    mesh = q.sharding.mesh
    qa, ka = _axes(q), _axes(k)
    same_blocks = bool(ka[1]) and tuple(qa[1]) == tuple(ka[1]) and tuple(_axes(v)[1]) == tuple(ka[1]) and \
        q.shape[1] == k.shape[1] and tuple(qa[0]) == tuple(ka[0]) and tuple(qa[2]) == tuple(ka[2])
    if ring is None:
        ring = same_blocks
    if ring:
        if not same_blocks:
            raise ValueError("ring attention needs q, k and v split along S over the same mesh axes, with Sq == Sk")
        return _ring_attention(q, k, v, scale, causal)
    if ka[1]:
        full_kv = NamedSharding(mesh, _spec([ka[0], (), ka[2], ka[3]]))
        k = reshard(k, full_kv)
        v = reshard(v, full_kv)
    S = q.shape[1]
    out = {}
    for d in rt.local_device_ids:
        ql, kl, vl = q.shard(d), k.shard(d), v.shard(d)
        if causal and qa[1]:
            # a query block of a sequence-sharded q attends to keys up to its own global positions
            f0, f1 = q.index_of(d)[1].start, q.index_of(d)[1].stop
            kl, vl = kl[:, :f1], vl[:, :f1]
        out[d] = comp.attention_fwd(ql, kl, vl, scale=scale, causal=causal)
    return ShardedArray(q.shape, torch.bfloat16, q.sharding, out)


# ------------------------------------------------------------------- fused fast paths (one process per GPU)
# The generic plans above are what runs on a virtual mesh and under gloo.  On real GPUs (DistributedRuntime over NCCL) the
# same calls, with the same operand shardings, are served by the resident fused layers of layers.py -- peer-memory
# reduce-scatter fused into the row-parallel GEMM, bf16 weight copies made once, output buffers reused -- so that the
# reference-shaped API IS the throughput path.  A layer object is cached per (parameter arrays, activation shape); the
# cache keeps the parameter arrays alive, and `invalidate_fast_layers()` drops it (train steps that update f32 master
# parameters in place call it, because the layers hold bf16 copies of those).
_fast_layers: Dict[tuple, object] = {}


def invalidate_fast_layers():
    _fast_layers.clear()


def _model_group(rt, mesh, axis):
    """(ranks, process group) of the ranks that differ from this one only along `axis`; collective over the job."""
    ranks, pg = rt._my_group(mesh.axis_groups((axis,)))
    if pg is None and len(ranks) > 1:      # the axis spans the whole job: the default group, as an explicit handle
        pg = rt.dist.group.WORLD
    return ranks, pg


def _fast_ffn(x, w1, b1, w2, b2, act, scatter_dim):
    rt = _rt.get_runtime()
    if getattr(rt, "kind", "") != "distributed" or not rt.on_cuda or scatter_dim is None or _rt._compute_override is not None:
        return None
    if not all(isinstance(t.sharding, NamedSharding) for t in (x, w1, w2)):
        return None
    mesh = x.sharding.mesh
    xa = _axes(x)
    (k1, n1), (k2, n2) = _axes(w1), _axes(w2)
    if k1 or n2 or len(n1) != 1 or tuple(k2) != tuple(n1) or xa[-1]:
        return None                                   # not the column-parallel -> row-parallel pattern
    m = n1[0]
    W = mesh.axis_size(m)
    if W == 1 or any(m in r for r in xa) or x.ndim not in (2, 3) or scatter_dim >= x.ndim - 1:
        return None                                   # x must be replicated over the model axis
    xs = x.shard(rt.rank)
    lead = tuple(xs.shape[:-1])
    T = 1
    for v in lead:
        T *= v
    d = xs.shape[-1]
    if x.ndim == 3 and scatter_dim == 1:
        batch = lead[0]
        if lead[1] % W or (lead[1] // W) % 128:
            return None
        out_local = (lead[0], lead[1] // W, d)
    else:                                             # leading dim: contiguous blocks of the flattened rows
        batch = 1
        if lead[0] % W:
            return None
        out_local = (lead[0] // W,) + lead[1:] + (d,)
    if T % (W * 256):
        return None
    key = ("ffn", id(w1), id(b1), id(w2), id(b2), lead, act, scatter_dim)
    entry = _fast_layers.get(key)
    if entry is None:
        from . import layers
        ranks, pg = _model_group(rt, mesh, m)
        if b1 is not None:
            b1 = reshard(b1, NamedSharding(mesh, PartitionSpec(m)))
        if b2 is not None and not b2.sharding.is_fully_replicated(b2.shape):
            b2 = reshard(b2, NamedSharding(mesh, PartitionSpec()))
        bf = lambda t: None if t is None else _to_bf16(t.shard(rt.rank)).contiguous()
        try:
            layer = layers.FusedTensorParallelFFN(bf(w1), bf(b1), bf(w2), bf(b2), act, T, pg, batch=batch)
        except Exception:       # noqa: BLE001 - e.g. CUDA IPC not permitted: the generic plan still works
            _fast_layers[key] = False
            return None
        entry = _fast_layers[key] = (layer, (w1, b1, w2, b2))      # the tuple keeps the ids above valid
    if entry is False:
        return None
    layer = entry[0]
    y = layer.forward(_to_bf16(xs.reshape(T, d)))
    rows = [tuple(r) for r in xa[:-1]]
    rows[scatter_dim] = rows[scatter_dim] + (m,)
    return ShardedArray(x.shape, torch.bfloat16, NamedSharding(mesh, _spec(rows + [()])), {rt.rank: y.view(out_local)})


# ------------------------------------------------------------------------------------ FFN
def ffn(x: ShardedArray, w1: ShardedArray, b1: Optional[ShardedArray], w2: ShardedArray, b2: Optional[ShardedArray], *,
        act: str = "gelu_tanh", scatter_dim: Optional[int] = None) -> ShardedArray:
    """y = act(x.W1 + b1).W2 + b2 (case6_attention.py:36-40).  With W1 ``P(None,'model')`` and W2
    ``P('model',None)`` this is the column-parallel -> activation -> row-parallel block: one
    all-reduce (or reduce-scatter with ``scatter_dim``) over 'model' at the end.  With W1/W2
    ``P('data',None)`` the weights are all-gathered first (FSDP).

    One process per GPU, W1 / W2 sharded as above, x replicated over 'model' and ``scatter_dim`` given: the call is served by
    ``layers.FusedTensorParallelFFN`` (reduce-scatter fused into the second GEMM over NVLink peer memory); the returned
    array then aliases the layer's output buffer, which the next call with the same parameters overwrites."""
    fast = _fast_ffn(x, w1, b1, w2, b2, act, scatter_dim)
    if fast is not None:
        return fast
    h = dense(x, w1, b1, epilogue=act)
    return dense(h, w2, b2, scatter_dim=scatter_dim)


# ------------------------------------------------------------------- modules / parameters
def _init_normal(shape, fan_in, seed) -> torch.Tensor:
    g = torch.Generator().manual_seed(seed)
    return torch.randn(shape, generator=g, dtype=torch.float32) / math.sqrt(fan_in)


class FlaxAttention:
    """The reference's attention module (case6_attention.py:42-143): q/k/v Dense without bias,
    f32 softmax attention, to_out_0 Dense with bias, dropout(0) = identity.

    ``layout="literal"`` keeps the reference's annotations (activations sequence-sharded through the
    logical name 'embed', heads replicated, case6_attention.py:105-107,114-116); ``layout="heads"``
    is the modified annotation set BASELINE.json's case6 config asks for: rule
    ('heads','model') in place of ('embed','model'), q/k/v constrained to ('batch', None, 'heads'),
    out-projection row-parallel with a reduce-scatter back to the sequence-sharded output."""

    def __init__(self, query_dim: int, heads: int = 8, dim_head: int = 64, causal: bool = False,
                 layout: str = "literal", ring: Optional[bool] = None):
        self.ring = ring                                   # literal layout: None / True = ring attention, False = all-gather K, V
        self.query_dim, self.heads, self.dim_head = query_dim, heads, dim_head
        self.scale = dim_head ** -0.5                      # case6_attention.py:51
        self.causal = causal
        self.layout = layout

    # logical names of the parameters (case6_attention.py:56-59,85-88)
    param_axes = {"to_q": ("embed", "heads"), "to_k": ("embed", "heads"), "to_v": ("embed", "heads"),
                  "to_out_0": ("heads", "embed")}
    biased = ("to_out_0",)          # nn.Dense(query_dim) keeps its default use_bias=True (case6_attention.py:83-90)

    def param_shapes(self):
        inner = self.heads * self.dim_head
        return {"to_q": (self.query_dim, inner), "to_k": (self.query_dim, inner), "to_v": (self.query_dim, inner),
                "to_out_0": (inner, self.query_dim)}

    def init(self, key, out_shardings) -> Dict[str, object]:
        """``model.init`` under ``jit(..., out_shardings=logical_state_sharding)`` (case6_attention.py:171-196): every
        Dense kernel is lecun_normal and comes into being ALREADY SHARDED -- each device generates exactly its slice in its
        own memory (counter-based threefry, ``random.lecun_normal``), nothing is drawn on the host or moved afterwards;
        the bias is zeros (flax's default bias_init).  ``out_shardings`` = ``logical_to_mesh_sharding(
        get_partition_spec(model), mesh, rules)``; ``key`` = ``random.PRNGKey(seed)``."""
        from . import random as _random
        rt = _rt.get_runtime()
        params = {}
        for name, shape in self.param_shapes().items():
            sh = out_shardings[name]
            if name in self.biased:
                k_sh, b_sh = sh
                kernel = _random.lecun_normal(_random.path_key(key, name + "/kernel"), shape, k_sh)
                idx = b_sh.devices_indices_map((shape[1],))
                bias = ShardedArray((shape[1],), torch.float32, b_sh,
                                    {d: torch.zeros(idx[d][0].stop - idx[d][0].start, dtype=torch.float32,
                                                    device=rt.buffer_device(d)) for d in rt.local_device_ids if d in idx})
                params[name] = (kernel, bias)
            else:
                params[name] = _random.lecun_normal(_random.path_key(key, name + "/kernel"), shape, sh)
        return params

    def init_host(self, seed: int = 1) -> Dict[str, object]:
        """lecun-normal-like f32 parameters on the host (stand-in for model.init, case6_attention.py:171-178)."""
        inner = self.heads * self.dim_head
        return {
            "to_q": _init_normal((self.query_dim, inner), self.query_dim, seed),
            "to_k": _init_normal((self.query_dim, inner), self.query_dim, seed + 1),
            "to_v": _init_normal((self.query_dim, inner), self.query_dim, seed + 2),
            "to_out_0": (_init_normal((inner, self.query_dim), inner, seed + 3), torch.zeros(self.query_dim)),
        }

    def shard_params(self, host_params, mesh: Mesh, rules) -> Dict[str, object]:
        """Place parameters with the sharding their logical axes translate to
        (nn.logical_to_mesh_sharding, case6_attention.py:189-191)."""
        out = {}
        for name, value in host_params.items():
            sh = logical_to_mesh_sharding(self.param_axes[name], mesh, rules)
            if name == "to_out_0":
                kernel, bias = value
                out[name] = (device_put(kernel, sh), device_put(bias, NamedSharding(mesh, PartitionSpec())))
            else:
                out[name] = device_put(value, sh)
        return out

    def _fast_heads(self, params, x: ShardedArray):
        """layout="heads" on real GPUs: fused q/k/v projection, attention on this rank's heads, row-parallel out-projection
        with the fused reduce-scatter along S (layers.ShardedMHA); None when the shardings are not the head-sharded set."""
        rt = _rt.get_runtime()
        if getattr(rt, "kind", "") != "distributed" or not rt.on_cuda or _rt._compute_override is not None or x.ndim != 3:
            return None
        mesh = x.sharding.mesh
        w_o, b_o = params["to_out_0"]
        kq, nq = _axes(params["to_q"])
        ko, no = _axes(w_o)
        if kq or no or len(nq) != 1 or tuple(ko) != tuple(nq) or any(tuple(_axes(params[n])) != (kq, nq) for n in ("to_k", "to_v")):
            return None
        m = nq[0]
        W = mesh.axis_size(m)
        xa = _axes(x)
        if W == 1 or self.heads % W or xa[2] or any(m in r for r in xa):
            return None
        xs = x.shard(rt.rank)
        Bl, S, M = xs.shape
        if S % W or (S // W) % 128 or (Bl * S) % (W * 256):
            return None
        key = ("mha", id(params["to_q"]), id(params["to_k"]), id(params["to_v"]), id(w_o), id(b_o), (Bl, S), self.causal)
        entry = _fast_layers.get(key)
        if entry is None:
            from . import layers
            ranks, pg = _model_group(rt, mesh, m)
            if not b_o.sharding.is_fully_replicated(b_o.shape):
                b_o = reshard(b_o, NamedSharding(mesh, PartitionSpec()))
            bf = lambda t: _to_bf16(t.shard(rt.rank)).contiguous()
            try:
                layer = layers.ShardedMHA(bf(params["to_q"]), bf(params["to_k"]), bf(params["to_v"]), bf(w_o), bf(b_o),
                                          self.heads // W, self.dim_head, Bl, S, self.causal, pg, scatter="seq")
            except Exception:   # noqa: BLE001
                _fast_layers[key] = False
                return None
            entry = _fast_layers[key] = (layer, params)
        if entry is False:
            return None
        y = entry[0].forward(_to_bf16(xs))
        spec = _spec([tuple(xa[0]), tuple(xa[1]) + (m,), ()])
        return ShardedArray(x.shape[:-1] + (self.query_dim,), torch.bfloat16, NamedSharding(mesh, spec),
                            {rt.rank: y.view(Bl, S // W, self.query_dim)})

    def __call__(self, params, hidden_states: ShardedArray) -> ShardedArray:
        x = hidden_states
        if self.layout == "heads":
            x = with_logical_constraint(x, ("batch", None, None))          # gather the sequence on entry
            fast = self._fast_heads(params, x)
            if fast is not None:
                return fast
        q = dense(x, params["to_q"])                                        # :96
        k = dense(x, params["to_k"])                                        # :98
        v = dense(x, params["to_v"])                                        # :99
        act_names = ("batch", "embed", None) if self.layout == "literal" else ("batch", None, "heads")
        q, k, v = (with_logical_constraint(t, act_names) for t in (q, k, v))   # :105-107
        q, k, v = (split_heads(t, self.heads) for t in (q, k, v))               # :109-112
        names4 = ("batch", "embed", None, None) if self.layout == "literal" else ("batch", None, "heads", None)
        q, k, v = (with_logical_constraint(t, names4) for t in (q, k, v))       # :114-116
        o = dot_product_attention(q, k, v, scale=self.scale, causal=self.causal, ring=self.ring)   # :121-133
        o = merge_heads(o)                                                  # :135
        o = with_logical_constraint(o, ("batch", "kv", "heads"))            # :137
        w_o, b_o = params["to_out_0"]
        if self.layout == "heads":
            y = dense(o, w_o, b_o, scatter_dim=1)      # row-parallel + reduce-scatter to S-sharded
            return y
        y = dense(o, w_o, b_o)                                              # :139
        return with_logical_constraint(y, ("batch", "embed"))               # :141 (dropout(0) = identity, :143)


# ------------------------------------------------------------------------------------ jit
def jit(fn, in_shardings=None, out_shardings=None):
    """``jax.jit(fn, in_shardings=.., out_shardings=..)`` as the scripts use it
    (case6_attention.py:229-232): arguments that are ShardedArrays are moved to ``in_shardings``
    if they are not there already, results to ``out_shardings``.  There is no tracing compiler --
    the layers are already kernels; this only pins the boundary shardings."""

    def _apply(tree, shardings):
        if shardings is None:
            return tree
        if isinstance(tree, ShardedArray):
            return reshard(tree, shardings)
        if isinstance(tree, dict):
            return {k: _apply(v, shardings[k] if isinstance(shardings, dict) else shardings) for k, v in tree.items()}
        if isinstance(tree, (tuple, list)):
            sh = shardings if isinstance(shardings, (tuple, list)) else [shardings] * len(tree)
            return type(tree)(_apply(v, s) for v, s in zip(tree, sh))
        return tree

    def wrapped(*args):
        if in_shardings is not None:
            args = tuple(_apply(a, s) for a, s in zip(args, in_shardings))
        return _apply(fn(*args), out_shardings)

    return wrapped
