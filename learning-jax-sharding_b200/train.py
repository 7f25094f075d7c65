"""The training half of the reference's Dense module (first slice): what
``train_step`` does in case5_attention_dense.py:130-139 -- ``loss = attention.apply(params, x).sum()``, ``jax.grad``,
``state.apply_gradients`` with ``optax.adam(learning_rate=0.001)`` (:107) -- for a module that is one bf16
``nn.Dense(use_bias=False)`` (:59-67), on the sharded state the script derives from its logical rules (:109-115).

Host mirror of the pieces the script names: ``adam`` (optax.adam), ``TrainState.create / apply_gradients``
(flax.training.train_state), and ``dense_sum_loss_grad`` (the kernel gradient jax.grad produces for that loss).  The
arithmetic runs in the library: the weight-gradient product x^T . dy is b200_gemm_bf16_ex with a_layout = B200_A_KM
(tcgen05, MN-major A tiles through TMA, no transposed copy), the update is b200_adam_step.  For the attention module
(case6_attention.py:206-220) ``attention_train_step`` chains capi.attention_fwd_lse / capi.attention_bwd and the Dense
backward GEMMs over the sharded state in the head-sharded layout.
"""
from __future__ import annotations

from dataclasses import dataclass
from typing import Callable, Dict, Optional

import torch

from . import runtime as _rt
from .array import ShardedArray, reshard
from .nn import _axes, _spec, _to_bf16
from .sharding import NamedSharding


@dataclass(frozen=True)
class Adam:
    """optax.adam(learning_rate, b1=0.9, b2=0.999, eps=1e-8): scale_by_adam, then scale(-learning_rate)."""
    learning_rate: float = 1e-3
    b1: float = 0.9
    b2: float = 0.999
    eps: float = 1e-8


def adam(learning_rate: float = 1e-3, b1: float = 0.9, b2: float = 0.999, eps: float = 1e-8) -> Adam:
    return Adam(learning_rate, b1, b2, eps)


def _zeros_like(p: ShardedArray) -> ShardedArray:
    return ShardedArray(p.shape, torch.float32, p.sharding, {d: torch.zeros_like(t, dtype=torch.float32)
                                                             for d, t in p.local_shards.items()})


class TrainState:
    """flax.training.train_state.TrainState for f32 parameters held as ShardedArrays: ``step``, ``params``
    (name -> ShardedArray), ``tx`` and the optimizer state (first / second moments in the parameters' sharding, which is
    what ``nn.get_partition_spec`` + ``logical_to_mesh_sharding`` give the reference's state, case5_attention_dense.py
    :113-115)."""

    def __init__(self, step: int, apply_fn: Optional[Callable], params: Dict[str, ShardedArray], tx: Adam,
                 mu: Dict[str, ShardedArray], nu: Dict[str, ShardedArray]):
        self.step, self.apply_fn, self.params, self.tx, self.mu, self.nu = step, apply_fn, params, tx, mu, nu

    @classmethod
    def create(cls, *, apply_fn: Optional[Callable], params: Dict[str, ShardedArray], tx: Adam) -> "TrainState":
        for name, p in params.items():
            if p.dtype != torch.float32:
                raise TypeError(f"TrainState keeps f32 parameters (flax's param_dtype); {name} is {p.dtype}")
        return cls(0, apply_fn, dict(params), tx, {k: _zeros_like(p) for k, p in params.items()},
                   {k: _zeros_like(p) for k, p in params.items()})

    def apply_gradients(self, *, grads: Dict[str, ShardedArray]) -> "TrainState":
        """One optimizer update, in place on every local shard (b200_adam_step); returns the state with step + 1."""
        comp = _rt.local_compute()
        rt = _rt.get_runtime()
        step = self.step + 1
        for name, p in self.params.items():
            g = grads[name]
            if not g.sharding.same_layout(p.sharding, p.shape):
                g = reshard(g, p.sharding)
            for d in rt.local_device_ids:
                if d in p.local_shards:
                    comp.adam_step(p.shard(d), g.shard(d).contiguous(), self.mu[name].shard(d), self.nu[name].shard(d),
                                   step=step, lr=self.tx.learning_rate, b1=self.tx.b1, b2=self.tx.b2, eps=self.tx.eps)
        from . import nn as _nn
        _nn.invalidate_fast_layers()      # the fused layers hold bf16 copies of the f32 parameters just updated in place
        _nn.invalidate_param_casts(*self.params.values())     # ... and so do the parameter arrays themselves
        return TrainState(step, self.apply_fn, self.params, self.tx, self.mu, self.nu)


def dense_sum_loss_grad(x: ShardedArray, kernel: ShardedArray) -> ShardedArray:
    """d/d kernel of ``nn.dense(x, kernel).sum()`` in the kernel's sharding, f32 holding bf16-valued numbers (the oracle's
    ``dense_sum_loss_grads``): on every device the partial product x_local^T . ones[rows_local, N_local] with f32
    accumulation and f32 output, an all-reduce over the mesh axes that shard the rows of x (the data-parallel gradient
    reduction), ONE rounding to bf16 as the unsharded dot would apply, then local slicing to the kernel's sharding."""
    rt = _rt.get_runtime()
    comp = _rt.local_compute()
    mesh = x.sharding.mesh
    xa = _axes(x)
    rows, kx = xa[:-1], xa[-1]
    row_axes = tuple(a for r in rows for a in r)
    kw, nw = _axes(kernel)
    # the forward (nn.dense) cannot shard y's features over an axis the rows of x use: the cotangent follows y
    nw_eff = tuple(a for a in nw if a not in row_axes)
    kdim, ndim = kernel.shape
    part = {}
    for d in rt.local_device_ids:
        xs = x.shard(d)
        x2 = _to_bf16(xs.reshape(-1, xs.shape[-1]))
        n_local = ndim
        for a in nw_eff:
            n_local //= mesh.shape[a]
        dy = torch.ones((x2.shape[0], n_local), dtype=torch.bfloat16, device=x2.device)
        part[d] = comp.gemm_bf16_tn(x2, dy, out_dtype=torch.float32)
    if row_axes:
        part = rt.all_reduce(part, mesh.axis_groups(row_axes))
    # one rounding to bf16, kept as f32 (promote_dtype's cast of the cotangent back to the f32 param dtype)
    part = {d: comp.cast_bf16_to_f32(comp.cast_f32_to_bf16(t.contiguous())) for d, t in part.items()}
    grad = ShardedArray((kdim, ndim), torch.float32, NamedSharding(mesh, _spec([kx, nw_eff])), part)
    return reshard(grad, kernel.sharding)


def dense_train_step(state: TrainState, x: ShardedArray, name: str = "kernel") -> TrainState:
    """train_step of case5_attention_dense.py:130-139 for the single-Dense module: grads of sum(Dense(x)), then Adam."""
    return state.apply_gradients(grads={name: dense_sum_loss_grad(x, state.params[name])})


# ------------------------------------------------------------------ the attention module (case6_attention.py:206-220)
def attention_sum_loss_grads(module, params, x: ShardedArray) -> Dict[str, ShardedArray]:
    """Gradients of ``module(params, x).sum()`` for ``nn.FlaxAttention(layout="heads")`` in the parameters' shardings
    (to_q / to_k / to_v column-parallel over the model axis, to_out_0 row-parallel, batch over the data axis): what
    ``jax.grad(loss_unrolled)`` returns in the reference's train_step (case6_attention.py:206-214) under the head-sharded
    annotation set of BASELINE.json's case6 config.  Per device: the forward with its residuals (q, k, v, o, lse2), then
        dO = dy . Wo^T                         (b200_gemm_bf16, b_layout = NK; dy = ones, the cotangent of .sum())
        dWo = o^T . dy, dWq|k|v = x^T . dq|k|v (b200_gemm_bf16_ex, a_layout = KM, f32 partials)
        dq, dk, dv                             (b200_attention_bwd)
    and ONE collective kind: the all-reduce of the weight-gradient partials over the data axis.  Nothing crosses the model
    axis in the backward (x and dy are replicated over it, every head's chain is local).  Keys of the result: "to_q", "to_k",
    "to_v", "to_out_0.kernel", "to_out_0.bias"."""
    from . import nn as _nn
    if module.layout == "literal":
        return _attention_sum_loss_grads_literal(module, params, x)
    if module.layout != "heads":
        raise NotImplementedError(f"attention_sum_loss_grads: unknown layout {module.layout!r}")
    rt = _rt.get_runtime()
    comp = _rt.local_compute()
    mesh = x.sharding.mesh
    w_o, b_o = params["to_out_0"]
    for name in ("to_q", "to_k", "to_v"):
        kw, _ = _axes(params[name])
        if kw:
            raise NotImplementedError(f"{name} must be column-parallel (dim 0 unsharded) in the 'heads' layout")
    if _axes(w_o)[1]:
        raise NotImplementedError("to_out_0 must be row-parallel (dim 1 unsharded) in the 'heads' layout")
    H, D = module.heads, module.dim_head
    xg = _nn.with_logical_constraint(x, ("batch", None, None))                # the sequence gathered, as in the forward
    data_axes = tuple(_axes(xg)[0])
    q, k, v = (_nn.dense(xg, params[n]) for n in ("to_q", "to_k", "to_v"))   # features sharded like the kernels' columns
    q4, k4, v4 = (_nn.split_heads(t, H) for t in (q, k, v))
    B, S, M = x.shape
    parts = {n: {} for n in ("to_q", "to_k", "to_v", "to_out_0.kernel")}
    for d in rt.local_device_ids:
        ql, kl, vl = q4.shard(d), k4.shard(d), v4.shard(d)                    # [b_loc, S, H_loc, D]
        o, lse = comp.attention_fwd_lse(ql, kl, vl, scale=module.scale, causal=module.causal)
        rows = ql.shape[0] * S
        dy = torch.ones((rows, M), dtype=torch.bfloat16, device=ql.device)
        wo_l = _to_bf16(w_o.shard(d))                                         # [H_loc * D, M]
        do = comp.gemm_bf16(dy, wo_l, b_layout=1)                             # dy . Wo^T -> [rows, H_loc * D]
        o2 = o.reshape(rows, -1)
        parts["to_out_0.kernel"][d] = comp.gemm_bf16_tn(o2, dy, out_dtype=torch.float32)
        dq, dk, dv = comp.attention_bwd(ql, kl, vl, o, do.reshape(ql.shape), lse, scale=module.scale, causal=module.causal)
        x2 = _to_bf16(xg.shard(d).reshape(rows, M))
        for n, g in (("to_q", dq), ("to_k", dk), ("to_v", dv)):
            parts[n][d] = comp.gemm_bf16_tn(x2, g.reshape(rows, -1), out_dtype=torch.float32)
    grads = {}
    for n, part in parts.items():
        if data_axes:
            part = rt.all_reduce(part, mesh.axis_groups(data_axes))           # the data-parallel gradient reduction
        part = {d: comp.cast_bf16_to_f32(comp.cast_f32_to_bf16(t.contiguous())) for d, t in part.items()}
        target = w_o if n == "to_out_0.kernel" else params[n]
        grads[n] = ShardedArray(target.shape, torch.float32, target.sharding, part)
    # d sum(y) / d bias = number of (batch, position) rows, in the bias's dtype chain (bf16) like the oracle
    nrows = float(torch.tensor(float(B * S)).to(torch.bfloat16))
    grads["to_out_0.bias"] = ShardedArray(b_o.shape, torch.float32, b_o.sharding,
                                          {d: torch.full_like(t, nrows, dtype=torch.float32) for d, t in b_o.local_shards.items()})
    return grads


def _attention_sum_loss_grads_literal(module, params, x: ShardedArray) -> Dict[str, ShardedArray]:
    """The same gradients under the reference's LITERAL annotation set (case6_attention.py:56-59,85-88,105-116,137,141 with
    the rules of :183-187): activations split along S over 'model', to_q / to_k / to_v split on their rows (embed), to_out_0
    on its columns.  The backward is the transpose of the forward's resharding points, one by one:
        :141  all-to-all M -> S            | dy = ones arrives S-sharded; its M-sharded form is ones again
        :139  column-parallel to_out_0     | dWo = o^T . dy (all-reduce over 'data'); dO = dy . Wo^T is a PARTIAL sum over the
                                           | kernel's column shards
        :137  all-gather of o along S      | ... which a reduce-scatter along S over 'model' both completes and re-splits
        :121-135 attention, K / V whole    | b200_attention_bwd on the local query block against the gathered K, V (for a
                                           | causal mask: the keys up to the block's last query); dK, dV partials cover every
                                           | key -> reduce-scatter along S; dQ is local
        :96-99 Dense on gathered kernels   | dW = x^T . dq|dk|dv summed over every token shard (all-reduce over 'data' and
                                           | 'model'), then cut to the kernel's row shard
    Every weight gradient is rounded to bf16 once, after its full sum, like the cotangent of an unsharded bf16 dot."""
    from . import nn as _nn
    rt = _rt.get_runtime()
    comp = _rt.local_compute()
    mesh = x.sharding.mesh
    H, D = module.heads, module.dim_head
    B, S, M = x.shape
    w_o, b_o = params["to_out_0"]
    q, k, v = (_nn.dense(x, params[n]) for n in ("to_q", "to_k", "to_v"))                     # :96-99, S-sharded like x
    q, k, v = (_nn.with_logical_constraint(t, ("batch", "embed", None)) for t in (q, k, v))   # :105-107
    q4, k4, v4 = (_nn.split_heads(t, H) for t in (q, k, v))                                    # :109-112
    qa = _axes(q4)
    data_axes, seq_axes = tuple(qa[0]), tuple(qa[1])
    if qa[2] or qa[3]:
        raise NotImplementedError("the literal layout keeps heads and head dim unsharded")
    ko, no = _axes(w_o)
    if ko or (seq_axes and tuple(no) != seq_axes):
        raise NotImplementedError("to_out_0 must be column-parallel over the axis that splits S (the literal rules)")
    kv_spec = NamedSharding(mesh, _spec([data_axes, (), (), ()]))
    kf, vf = reshard(k4, kv_spec), reshard(v4, kv_spec)                                        # K, V whole along S
    seq_groups = mesh.axis_groups(seq_axes) if seq_axes else None
    o_sh, res = {}, {}
    for d in rt.local_device_ids:
        ql = q4.shard(d)
        f1 = q4.index_of(d)[1].stop
        kl, vl = kf.shard(d), vf.shard(d)
        if module.causal:
            kl, vl = kl[:, :f1].contiguous(), vl[:, :f1].contiguous()                          # keys up to the block's last query
        o_l, lse_l = comp.attention_fwd_lse(ql, kl, vl, scale=module.scale, causal=module.causal)
        o_sh[d], res[d] = o_l, (ql, kl, vl, lse_l)
    o4 = ShardedArray(q4.shape, torch.bfloat16, q4.sharding, o_sh)
    o_g = _nn.with_logical_constraint(_nn.merge_heads(o4), ("batch", "kv", "heads"))          # :135-137: gathered along S
    parts = {n: {} for n in ("to_q", "to_k", "to_v", "to_out_0.kernel")}
    do_part = {}
    for d in rt.local_device_ids:
        og = o_g.shard(d)
        rows = og.shape[0] * og.shape[1]
        wo_l = _to_bf16(w_o.shard(d))                                                          # [inner, M / m]
        dy = torch.ones((rows, wo_l.shape[1]), dtype=torch.bfloat16, device=og.device)
        parts["to_out_0.kernel"][d] = comp.gemm_bf16_tn(og.reshape(rows, -1), dy, out_dtype=torch.float32)
        do_part[d] = comp.gemm_bf16(dy, wo_l, b_layout=1, out_dtype=torch.float32).reshape(og.shape)   # partial over columns
    if seq_axes:
        do_part = rt.reduce_scatter(do_part, seq_groups, dim=1)                                # sum over 'model', split S again
    dk_part, dv_part = {}, {}
    for d in rt.local_device_ids:
        ql, kl, vl, lse_l = res[d]
        do_l = comp.cast_f32_to_bf16(do_part[d].contiguous()).reshape(ql.shape)
        dq, dk, dv = comp.attention_bwd(ql, kl, vl, o_sh[d], do_l, lse_l, scale=module.scale, causal=module.causal)
        xs = x.shard(d)
        x2 = _to_bf16(xs.reshape(-1, M))
        parts["to_q"][d] = comp.gemm_bf16_tn(x2, dq.reshape(x2.shape[0], -1), out_dtype=torch.float32)
        full = (ql.shape[0], S) + tuple(ql.shape[2:])
        for dst, g in ((dk_part, dk), (dv_part, dv)):
            buf = torch.zeros(full, dtype=torch.float32, device=g.device)                      # keys this block never sees: zero
            buf[:, :g.shape[1]] = comp.cast_bf16_to_f32(g)
            dst[d] = buf
    if seq_axes:
        dk_part, dv_part = rt.reduce_scatter(dk_part, seq_groups, dim=1), rt.reduce_scatter(dv_part, seq_groups, dim=1)
    for d in rt.local_device_ids:
        x2 = _to_bf16(x.shard(d).reshape(-1, M))
        for n, g in (("to_k", dk_part[d]), ("to_v", dv_part[d])):
            gb = comp.cast_f32_to_bf16(g.contiguous()).reshape(x2.shape[0], -1)
            parts[n][d] = comp.gemm_bf16_tn(x2, gb, out_dtype=torch.float32)
    grads = {}
    for n, part in parts.items():
        axes = data_axes if n == "to_out_0.kernel" else data_axes + seq_axes                   # the token shards summed over
        if axes:
            part = rt.all_reduce(part, mesh.axis_groups(axes))
        part = {d: comp.cast_bf16_to_f32(comp.cast_f32_to_bf16(t.contiguous())) for d, t in part.items()}
        if n == "to_out_0.kernel":
            grads[n] = ShardedArray(w_o.shape, torch.float32, w_o.sharding, part)
        else:
            full = ShardedArray(params[n].shape, torch.float32, NamedSharding(mesh, _spec([(), ()])), part)
            grads[n] = reshard(full, params[n].sharding)                                       # local cut to the row shard
    nrows = float(torch.tensor(float(B * S)).to(torch.bfloat16))
    grads["to_out_0.bias"] = ShardedArray(b_o.shape, torch.float32, b_o.sharding,
                                          {d: torch.full_like(t, nrows, dtype=torch.float32) for d, t in b_o.local_shards.items()})
    return grads


def attention_train_state(module, params, tx: Adam) -> TrainState:
    """TrainState over the module's parameters with the flat names attention_sum_loss_grads uses."""
    w_o, b_o = params["to_out_0"]
    flat = {"to_q": params["to_q"], "to_k": params["to_k"], "to_v": params["to_v"], "to_out_0.kernel": w_o,
            "to_out_0.bias": b_o}
    return TrainState.create(apply_fn=module, params=flat, tx=tx)


def attention_params_of(state: TrainState):
    """The nested parameter dict nn.FlaxAttention.__call__ takes, from a TrainState built by attention_train_state."""
    p = state.params
    return {"to_q": p["to_q"], "to_k": p["to_k"], "to_v": p["to_v"], "to_out_0": (p["to_out_0.kernel"], p["to_out_0.bias"])}


def attention_train_step(module, state: TrainState, x: ShardedArray) -> TrainState:
    """train_step of case6_attention.py:206-220: grads of ``module(params, x).sum()``, then ``state.apply_gradients``."""
    return state.apply_gradients(grads=attention_sum_loss_grads(module, attention_params_of(state), x))
