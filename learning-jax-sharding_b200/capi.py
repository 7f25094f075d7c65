"""ctypes binding of libb200sl.so (include/b200_sharded.h).

The library is the product: there is NO CPU or eager-PyTorch fallback.  If the shared
object has not been built (``learning-jax-sharding_b200/build.sh`` or
``__graft_entry__.build()``), or a tensor is not a CUDA tensor, the wrappers raise.

Tensors are only containers here: each wrapper passes ``tensor.data_ptr()`` and the
current torch CUDA stream to the ``extern "C"`` entry point, which enqueues its kernels
on that stream and returns immediately.
"""
from __future__ import annotations

import ctypes
import os

import torch

_HERE = os.path.dirname(os.path.abspath(__file__))
# B200_LIB_PATH selects another build of the same library (A/B runs of compile-time variants, tools/build_variants.sh)
LIB_PATH = os.environ.get("B200_LIB_PATH") or os.path.join(_HERE, "libb200sl.so")

# enums of include/b200_sharded.h
BF16, F32 = 0, 1
EPI_NONE, EPI_BIAS, EPI_BIAS_RELU, EPI_BIAS_GELU_TANH, EPI_RELU, EPI_GELU_TANH = range(6)
B_KN, B_NK = 0, 1
A_MK, A_KM = 0, 1
RNG_BITS, RNG_UNIFORM, RNG_NORMAL, RNG_TRUNCATED_NORMAL = range(4)

EXPORTED_SYMBOLS = (
    "b200_version",
    "b200_last_error",
    "b200_launch_count",
    "b200_gemm_bf16",
    "b200_gemm_bf16_rowblocked",
    "b200_gemm_bf16_ex",
    "b200_gemm_f32",
    "b200_attention_fwd",
    "b200_attention_fwd_ex",
    "b200_ffn_fwd",
    "b200_ffn_fwd_static_scratch",
    "b200_cast_f32_to_bf16",
    "b200_cast_bf16_to_f32",
    "b200_sum_partials_to_bf16",
    "b200_adam_step",
    "b200_attention_fwd_lse",
    "b200_attention_bwd",
    "b200_attention_combine",
    "b200_comm_unique_id",
    "b200_mesh_create",
    "b200_mesh_axis_comm",
    "b200_comm_size",
    "b200_comm_rank",
    "b200_comm_release",
    "b200_mesh_destroy",
    "b200_all_reduce",
    "b200_reduce_scatter",
    "b200_all_gather",
    "b200_all_to_all",
    "b200_ffn_fwd_tp",
    "b200_peer_alloc",
    "b200_peer_open",
    "b200_peer_close",
    "b200_peer_free",
    "b200_peer_copy2d",
    "b200_peer_gather",
    "b200_peer_push",
    "b200_peer_progressive_sum",
    "b200_peer_barrier",
    "b200_sum_bf16_partials",
    "b200_random_fill",
)


class GemmTuning(ctypes.Structure):
    """b200_gemm_tuning of include/b200_sharded.h: per-call overrides of the launch heuristics (all zero = library choice)."""
    _fields_ = [("cta_group", ctypes.c_int), ("block_n", ctypes.c_int), ("group_m", ctypes.c_int),
                ("max_stages", ctypes.c_int), ("max_ctas", ctypes.c_int), ("k_serpentine", ctypes.c_int)]


class AttentionArgs(ctypes.Structure):
    """b200_attention_args of include/b200_sharded.h."""
    _fields_ = [("Q", ctypes.c_void_p), ("K", ctypes.c_void_p), ("V", ctypes.c_void_p), ("O", ctypes.c_void_p),
                ("lse2", ctypes.c_void_p), ("B", ctypes.c_int), ("Sq", ctypes.c_int), ("Sk", ctypes.c_int),
                ("H", ctypes.c_int), ("D", ctypes.c_int), ("strides", ctypes.c_int64 * 12), ("scale", ctypes.c_float),
                ("causal", ctypes.c_int), ("variant", ctypes.c_int)]


class RandomArgs(ctypes.Structure):
    """b200_random_args of include/b200_sharded.h."""
    _fields_ = [("key", ctypes.c_uint32 * 2), ("ndim", ctypes.c_int), ("global_shape", ctypes.c_int64 * 4),
                ("shard_start", ctypes.c_int64 * 4), ("shard_shape", ctypes.c_int64 * 4), ("distribution", ctypes.c_int),
                ("lo", ctypes.c_float), ("hi", ctypes.c_float), ("scale", ctypes.c_float)]


class GemmArgs(ctypes.Structure):
    """b200_gemm_args of include/b200_sharded.h."""
    _fields_ = [("A", ctypes.c_void_p), ("lda", ctypes.c_int64), ("a_rows_per_block", ctypes.c_int),
                ("a_block_stride", ctypes.c_int), ("B", ctypes.c_void_p), ("ldb", ctypes.c_int64),
                ("b_layout", ctypes.c_int), ("bias", ctypes.c_void_p), ("C", ctypes.c_void_p), ("ldc", ctypes.c_int64),
                ("out_dtype", ctypes.c_int), ("c_block_ptrs", ctypes.c_void_p), ("m_rows_per_block", ctypes.c_int),
                ("m_block_rot", ctypes.c_int), ("M", ctypes.c_int), ("N", ctypes.c_int), ("K", ctypes.c_int),
                ("epilogue", ctypes.c_int), ("a_layout", ctypes.c_int), ("tune", GemmTuning),
                ("a_blocks_inner", ctypes.c_int), ("a_outer_stride", ctypes.c_int), ("m_block_interleave", ctypes.c_int),
                ("done_counters", ctypes.c_void_p), ("done_flag_ptrs", ctypes.c_void_p), ("done_sender", ctypes.c_int),
                ("done_groups", ctypes.c_int), ("done_epoch", ctypes.c_uint32)]


class AttentionBwdArgs(ctypes.Structure):
    """b200_attention_bwd_args of include/b200_sharded.h."""
    _fields_ = [("Q", ctypes.c_void_p), ("K", ctypes.c_void_p), ("V", ctypes.c_void_p), ("O", ctypes.c_void_p),
                ("dO", ctypes.c_void_p), ("lse2", ctypes.c_void_p), ("dQ", ctypes.c_void_p), ("dK", ctypes.c_void_p),
                ("dV", ctypes.c_void_p), ("workspace", ctypes.c_void_p), ("B", ctypes.c_int), ("Sq", ctypes.c_int),
                ("Sk", ctypes.c_int), ("H", ctypes.c_int), ("D", ctypes.c_int), ("strides", ctypes.c_int64 * 24),
                ("scale", ctypes.c_float), ("causal", ctypes.c_int)]


class B200Error(RuntimeError):
    pass


_lib = None


def load():
    """Load libb200sl.so once; raise loudly if it is missing (no fallback path exists)."""
    global _lib
    if _lib is not None:
        return _lib
    if not os.path.exists(LIB_PATH):
        raise B200Error(
            f"{LIB_PATH} is missing: build it with learning-jax-sharding_b200/build.sh "
            "(or __graft_entry__.build()). There is no CPU/PyTorch fallback for the sharded layers."
        )
    lib = ctypes.CDLL(LIB_PATH)
    i32, i64, vp, f32 = ctypes.c_int, ctypes.c_int64, ctypes.c_void_p, ctypes.c_float
    lib.b200_version.restype = i32
    lib.b200_last_error.restype = ctypes.c_char_p
    lib.b200_launch_count.restype = i64
    lib.b200_gemm_bf16.argtypes = [vp, i64, vp, i64, i32, vp, vp, i64, i32, i32, i32, i32, i32, vp]
    lib.b200_gemm_bf16_rowblocked.argtypes = [vp, i64, i32, i32, vp, i64, i32, vp, vp, i64, i32, i32, i32, i32, i32, vp]
    lib.b200_gemm_bf16_ex.argtypes = [ctypes.POINTER(GemmArgs), vp]
    lib.b200_peer_alloc.argtypes = [ctypes.c_uint64, ctypes.POINTER(vp), vp]
    lib.b200_peer_open.argtypes = [vp, ctypes.POINTER(vp)]
    lib.b200_peer_close.argtypes = [vp]
    lib.b200_peer_free.argtypes = [vp]
    lib.b200_peer_copy2d.argtypes = [vp, i64, vp, i64, i64, i64, vp]
    lib.b200_peer_gather.argtypes = [vp, i32, i32, i64, i64, vp, i64, i32, i64, i32, vp]
    lib.b200_peer_progressive_sum.argtypes = [vp, ctypes.c_uint32, i32, i32, i64, vp, i64, vp, i64, vp, i32, vp]
    lib.b200_peer_push.argtypes = [vp, i32, i32, i64, vp, i64, i32, vp]
    lib.b200_peer_barrier.argtypes = [vp, i32, i32, ctypes.c_uint32, vp]
    lib.b200_sum_bf16_partials.argtypes = [vp, i64, i32, vp, i64, vp, i64, vp]
    lib.b200_gemm_f32.argtypes = [vp, i64, i64, vp, i64, i64, vp, i64, i64, i32, i32, i32, i32, i32, vp]
    lib.b200_attention_fwd.argtypes = [vp, vp, vp, vp, i32, i32, i32, i32, i32] + [i64] * 12 + [f32, i32, vp]
    lib.b200_random_fill.argtypes = [ctypes.POINTER(RandomArgs), vp, vp]
    lib.b200_attention_fwd_ex.argtypes = [ctypes.POINTER(AttentionArgs), vp]
    lib.b200_ffn_fwd_static_scratch.argtypes = [vp, vp, vp, vp, vp, vp, i32, i32, i32, i32, i32, vp]
    lib.b200_attention_fwd_lse.argtypes = [vp, vp, vp, vp, vp, i32, i32, i32, i32, i32] + [i64] * 12 + [f32, i32, vp]
    lib.b200_attention_combine.argtypes = [vp, vp, vp, vp, i32, i32, i32, i32, i32, vp]
    lib.b200_attention_bwd.argtypes = [ctypes.POINTER(AttentionBwdArgs), vp]
    lib.b200_ffn_fwd.argtypes = [vp, vp, vp, vp, vp, vp, vp, i32, i32, i32, i32, i32, vp]
    lib.b200_cast_f32_to_bf16.argtypes = [vp, vp, i64, vp]
    lib.b200_cast_bf16_to_f32.argtypes = [vp, vp, i64, vp]
    lib.b200_sum_partials_to_bf16.argtypes = [vp, i64, i32, vp, i64, vp, i64, vp]
    lib.b200_adam_step.argtypes = [vp, vp, i32, vp, vp, i64, f32, f32, f32, f32, i32, vp]
    lib.b200_comm_unique_id.argtypes = [vp]
    lib.b200_mesh_create.argtypes = [vp, i32, i32, ctypes.POINTER(i32), i32, ctypes.POINTER(vp)]
    lib.b200_mesh_axis_comm.argtypes = [vp, i32, ctypes.POINTER(vp)]
    lib.b200_comm_size.argtypes = [vp]
    lib.b200_comm_rank.argtypes = [vp]
    lib.b200_comm_release.argtypes = [vp]
    lib.b200_mesh_destroy.argtypes = [vp]
    lib.b200_ffn_fwd_tp.argtypes = [vp] * 9 + [i32] * 6 + [vp]
    for _n in ("b200_all_reduce", "b200_reduce_scatter", "b200_all_gather", "b200_all_to_all"):
        getattr(lib, _n).argtypes = [vp, vp, vp, i64, i32, vp]
    for name in EXPORTED_SYMBOLS[3:]:
        getattr(lib, name).restype = i32
    _lib = lib
    return lib


def _check(rc: int, what: str):
    if rc != 0:
        msg = load().b200_last_error().decode("utf-8", "replace")
        raise B200Error(f"{what} failed (status {rc}): {msg}")


def _ptr(t):
    return None if t is None else ctypes.c_void_p(t.data_ptr())


def _stream(t: torch.Tensor):
    return ctypes.c_void_p(torch.cuda.current_stream(t.device).cuda_stream)


def _require_cuda(*tensors):
    for t in tensors:
        if t is not None and not t.is_cuda:
            raise B200Error(
                "b200 sharded layers run on CUDA tensors only (got a "
                f"{t.device} tensor); there is no CPU fallback"
            )


def launch_count() -> int:
    return int(load().b200_launch_count())


def _tuning(tune) -> GemmTuning:
    """``tune``: None or a dict with any of cta_group, block_n, group_m, max_stages, max_ctas, k_serpentine (see
    b200_gemm_tuning).  The overrides travel with the call; the library keeps no tuning state."""
    t = GemmTuning()
    for k, v in (tune or {}).items():
        if not hasattr(t, k):
            raise B200Error(f"unknown GEMM tuning field {k!r}")
        setattr(t, k, int(v))
    return t


def gemm_bf16(a, b, *, b_layout=B_KN, bias=None, epilogue=EPI_NONE, out=None, out_dtype=torch.bfloat16,
              a_row_blocks=None, tune=None):
    """out[M,N] = epilogue(a[M,K] @ B + bias). ``b`` is [K,N] (B_KN, a Dense kernel) or [N,K] (B_NK).

    a, b, out may be row-strided views (last dim contiguous).  ``a_row_blocks=(M, rows_per_block,
    block_stride)`` reads M logical rows out of a larger ``a`` through the row-blocked view of
    b200_gemm_bf16_rowblocked."""
    _require_cuda(a, b, bias, out)
    if a.dtype != torch.bfloat16 or b.dtype != torch.bfloat16:
        raise B200Error("gemm_bf16 takes bf16 operands")
    if a.dim() != 2 or b.dim() != 2 or a.stride(1) != 1 or b.stride(1) != 1:
        raise B200Error("gemm_bf16 takes 2-D operands with a contiguous last dim")
    M, K = a.shape
    rpb = bstride = 0
    if a_row_blocks is not None:
        M, rpb, bstride = a_row_blocks
    if b_layout == B_KN:
        Kb, N = b.shape
    else:
        N, Kb = b.shape
    if Kb != K:
        raise B200Error(f"gemm_bf16: contraction mismatch {tuple(a.shape)} x {tuple(b.shape)} (layout {b_layout})")
    if out is None:
        out = torch.empty((M, N), dtype=out_dtype, device=a.device)
    if out.shape != (M, N) or out.stride(1) != 1 or out.dtype not in (torch.bfloat16, torch.float32):
        raise B200Error("gemm_bf16: bad output tensor")
    if bias is not None and (bias.dtype != torch.bfloat16 or bias.numel() != N or not bias.is_contiguous()):
        raise B200Error("gemm_bf16: bias must be a contiguous bf16 vector of length N")
    args = GemmArgs()
    args.A, args.lda, args.a_rows_per_block, args.a_block_stride = a.data_ptr(), a.stride(0), rpb, bstride
    args.B, args.ldb, args.b_layout = b.data_ptr(), b.stride(0), b_layout
    args.bias = None if bias is None else bias.data_ptr()
    args.C, args.ldc = out.data_ptr(), out.stride(0)
    args.out_dtype = F32 if out.dtype == torch.float32 else BF16
    args.M, args.N, args.K, args.epilogue = M, N, K, epilogue
    args.tune = _tuning(tune)
    with torch.cuda.device(a.device):
        rc = load().b200_gemm_bf16_ex(ctypes.byref(args), _stream(a))
    _check(rc, "b200_gemm_bf16")
    return out


def gemm_bf16_tn(a_km, b, *, out=None, out_dtype=torch.bfloat16, tune=None):
    """out[M,N] = a_km^T @ b with a_km stored [K,M] and b [K,N], both row-major bf16: the weight-gradient product
    dW = x^T @ dy of a Dense layer (b200_gemm_bf16_ex with a_layout = B200_A_KM; no transposed copy is made, the TMA
    loads MN-major tiles)."""
    _require_cuda(a_km, b, out)
    if a_km.dtype != torch.bfloat16 or b.dtype != torch.bfloat16:
        raise B200Error("gemm_bf16_tn takes bf16 operands")
    if a_km.dim() != 2 or b.dim() != 2 or a_km.stride(1) != 1 or b.stride(1) != 1 or a_km.shape[0] != b.shape[0]:
        raise B200Error(f"gemm_bf16_tn: need [K,M] and [K,N] operands, got {tuple(a_km.shape)} and {tuple(b.shape)}")
    K, M = a_km.shape
    N = b.shape[1]
    if out is None:
        out = torch.empty((M, N), dtype=out_dtype, device=b.device)
    if out.shape != (M, N) or out.stride(1) != 1 or out.dtype not in (torch.bfloat16, torch.float32):
        raise B200Error("gemm_bf16_tn: bad output tensor")
    args = GemmArgs()
    args.A, args.lda, args.a_layout = a_km.data_ptr(), a_km.stride(0), A_KM
    args.B, args.ldb, args.b_layout = b.data_ptr(), b.stride(0), B_KN
    args.C, args.ldc = out.data_ptr(), out.stride(0)
    args.out_dtype = F32 if out.dtype == torch.float32 else BF16
    args.M, args.N, args.K, args.epilogue = M, N, K, EPI_NONE
    args.tune = _tuning(tune)
    with torch.cuda.device(b.device):
        rc = load().b200_gemm_bf16_ex(ctypes.byref(args), _stream(b))
    _check(rc, "b200_gemm_bf16_ex")
    return out


def adam_step(param, grad, mu, nu, *, step: int, lr=1e-3, b1=0.9, b2=0.999, eps=1e-8):
    """In-place optax.adam update of one f32 parameter tensor (b200_adam_step): mu, nu are the first / second moment
    buffers, `step` the 1-based count AFTER this update; grad is f32 or bf16."""
    _require_cuda(param, grad, mu, nu)
    if param.dtype != torch.float32 or mu.dtype != torch.float32 or nu.dtype != torch.float32:
        raise B200Error("adam_step: param, mu, nu must be fp32")
    if grad.dtype not in (torch.float32, torch.bfloat16):
        raise B200Error("adam_step: grad must be fp32 or bf16")
    n = param.numel()
    if not (param.is_contiguous() and grad.is_contiguous() and mu.is_contiguous() and nu.is_contiguous()) or \
            grad.numel() != n or mu.numel() != n or nu.numel() != n:
        raise B200Error("adam_step: tensors must be contiguous and of equal size")
    with torch.cuda.device(param.device):
        rc = load().b200_adam_step(_ptr(param), _ptr(grad), F32 if grad.dtype == torch.float32 else BF16, _ptr(mu), _ptr(nu),
                                   ctypes.c_int64(n), ctypes.c_float(lr), ctypes.c_float(b1), ctypes.c_float(b2),
                                   ctypes.c_float(eps), ctypes.c_int(step), _stream(param))
    _check(rc, "b200_adam_step")
    return param


def random_fill(out, key, global_shape, shard_start, distribution, *, lo=0.0, hi=0.0, scale=1.0):
    """Fill ``out`` (this device's contiguous shard: f32, or int32 for RNG_BITS) with the values jax.random gives the
    slice [shard_start, shard_start + out.shape) of an array of ``global_shape`` under ``key`` (two uint32 words)."""
    _require_cuda(out)
    nd = len(global_shape)
    if not (1 <= nd <= 4) or out.dim() != nd or len(shard_start) != nd or not out.is_contiguous():
        raise B200Error("random_fill: out must be a contiguous tensor of the same rank (1..4) as global_shape")
    want = torch.int32 if distribution == RNG_BITS else torch.float32
    if out.dtype != want:
        raise B200Error(f"random_fill: distribution {distribution} writes {want}")
    a = RandomArgs()
    a.key = (ctypes.c_uint32 * 2)(int(key[0]) & 0xFFFFFFFF, int(key[1]) & 0xFFFFFFFF)
    a.ndim = nd
    a.global_shape = (ctypes.c_int64 * 4)(*(list(global_shape) + [1] * (4 - nd)))
    a.shard_start = (ctypes.c_int64 * 4)(*(list(shard_start) + [0] * (4 - nd)))
    a.shard_shape = (ctypes.c_int64 * 4)(*(list(out.shape) + [1] * (4 - nd)))
    a.distribution, a.lo, a.hi, a.scale = distribution, float(lo), float(hi), float(scale)
    with torch.cuda.device(out.device):
        rc = load().b200_random_fill(ctypes.byref(a), _ptr(out), _stream(out))
    _check(rc, "b200_random_fill")
    return out


def gemm_f32(a, b, *, out=None, accumulate=False):
    """True-fp32 (batched) matmul: a [.., M, K] @ b [.., K, N]; leading batch dim optional."""
    _require_cuda(a, b, out)
    if a.dtype != torch.float32 or b.dtype != torch.float32:
        raise B200Error("gemm_f32 takes fp32 operands")
    batched = a.dim() == 3
    a3 = a if batched else a.unsqueeze(0)
    b3 = b if b.dim() == 3 else b.unsqueeze(0)
    if a3.dim() != 3 or b3.dim() != 3 or a3.shape[0] != b3.shape[0] or a3.shape[2] != b3.shape[1]:
        raise B200Error(f"gemm_f32: shape mismatch {tuple(a.shape)} x {tuple(b.shape)}")
    if a3.stride(2) != 1 or b3.stride(2) != 1:
        raise B200Error("gemm_f32 takes operands with a contiguous last dim")
    Bt, M, K = a3.shape
    N = b3.shape[2]
    if out is None:
        if accumulate:
            raise B200Error("gemm_f32: accumulate needs an output tensor")
        out = torch.empty((Bt, M, N) if batched else (M, N), dtype=torch.float32, device=a.device)
    o3 = out if out.dim() == 3 else out.unsqueeze(0)
    if o3.shape != (Bt, M, N) or o3.stride(2) != 1 or out.dtype != torch.float32:
        raise B200Error("gemm_f32: bad output tensor")
    with torch.cuda.device(a.device):
        rc = load().b200_gemm_f32(
            _ptr(a3), a3.stride(1), a3.stride(0), _ptr(b3), b3.stride(1), b3.stride(0), _ptr(o3), o3.stride(1),
            o3.stride(0), Bt, M, N, K, int(bool(accumulate)), _stream(a))
    _check(rc, "b200_gemm_f32")
    return out


def attention_fwd(q, k, v, *, scale=None, causal=False, out=None, lse=None, variant=0):
    """softmax(scale * q k^T [+causal]) v with q [B,Sq,H,D], k/v [B,Sk,H,D] bf16 (b s n h layout).  ``lse`` (optional
    f32 [B,H,Sq]) receives log2 of the softmax denominator; ``variant`` forces a kernel for this call (0 = library
    choice, 1 / 8 / 9 -- see b200_attention_args)."""
    _require_cuda(q, k, v, out, lse)
    for t in (q, k, v):
        if t.dtype != torch.bfloat16 or t.dim() != 4 or t.stride(3) != 1:
            raise B200Error("attention_fwd takes 4-D bf16 tensors [B,S,H,D] with contiguous D")
    B, Sq, H, D = q.shape
    Sk = k.shape[1]
    if k.shape != (B, Sk, H, D) or v.shape != (B, Sk, H, D):
        raise B200Error(f"attention_fwd: shape mismatch q{tuple(q.shape)} k{tuple(k.shape)} v{tuple(v.shape)}")
    if out is None:
        out = torch.empty((B, Sq, H, D), dtype=torch.bfloat16, device=q.device)
    if out.shape != (B, Sq, H, D) or out.dtype != torch.bfloat16 or out.stride(3) != 1:
        raise B200Error("attention_fwd: bad output tensor")
    if lse is not None and (lse.shape != (B, H, Sq) or lse.dtype != torch.float32 or not lse.is_contiguous()):
        raise B200Error("attention_fwd: lse must be a contiguous f32 [B,H,Sq] tensor")
    a = AttentionArgs()
    a.Q, a.K, a.V, a.O = q.data_ptr(), k.data_ptr(), v.data_ptr(), out.data_ptr()
    a.lse2 = None if lse is None else lse.data_ptr()
    a.B, a.Sq, a.Sk, a.H, a.D = B, Sq, Sk, H, D
    st = []
    for t in (q, k, v, out):
        st += [t.stride(0), t.stride(1), t.stride(2)]
    a.strides = (ctypes.c_int64 * 12)(*st)
    a.scale = float(D ** -0.5 if scale is None else scale)
    a.causal, a.variant = int(bool(causal)), int(variant)
    with torch.cuda.device(q.device):
        rc = load().b200_attention_fwd_ex(ctypes.byref(a), _stream(q))
    _check(rc, "b200_attention_fwd")
    return out


def attention_fwd_lse(q, k, v, *, scale=None, causal=False, variant=0):
    """attention_fwd that also returns lse2 [B,H,Sq] f32 (log2 of the softmax denominator, scale included): (out, lse2)."""
    _require_cuda(q, k, v)
    B, Sq, H, D = q.shape
    lse = torch.empty((B, H, Sq), dtype=torch.float32, device=q.device)
    out = attention_fwd(q, k, v, scale=scale, causal=causal, lse=lse, variant=variant)
    return out, lse


def attention_combine(acc, lse_acc, o_blk, lse_blk, *, first: bool):
    """Fold one key/value block's attention result (o_blk bf16 [B,S,H,D], lse_blk f32 [B,H,S]) into the running pair
    (acc f32 [B,S,H,D], lse_acc f32 [B,H,S]) in place -- b200_attention_combine, the merge step of ring attention."""
    _require_cuda(acc, lse_acc, o_blk, lse_blk)
    B, S, H, D = o_blk.shape
    if acc.shape != o_blk.shape or acc.dtype != torch.float32 or o_blk.dtype != torch.bfloat16 or \
            lse_acc.shape != (B, H, S) or lse_blk.shape != (B, H, S) or lse_acc.dtype != torch.float32 or \
            lse_blk.dtype != torch.float32:
        raise B200Error("attention_combine: acc f32 / o_blk bf16 [B,S,H,D], lse_acc / lse_blk f32 [B,H,S]")
    if not (acc.is_contiguous() and o_blk.is_contiguous() and lse_acc.is_contiguous() and lse_blk.is_contiguous()):
        raise B200Error("attention_combine takes contiguous tensors")
    with torch.cuda.device(acc.device):
        rc = load().b200_attention_combine(_ptr(acc), _ptr(lse_acc), _ptr(o_blk), _ptr(lse_blk), B, S, H, D, int(bool(first)),
                                           _stream(acc))
    _check(rc, "b200_attention_combine")
    return acc


def attention_bwd(q, k, v, o, do, lse2, *, scale=None, causal=False):
    """(dq, dk, dv) of attention_fwd for the cotangent ``do`` (b200_attention_bwd); o and lse2 from attention_fwd_lse."""
    _require_cuda(q, k, v, o, do, lse2)
    for t in (q, k, v, o, do):
        if t.dtype != torch.bfloat16 or t.dim() != 4 or t.stride(3) != 1:
            raise B200Error("attention_bwd takes 4-D bf16 tensors [B,S,H,D] with contiguous D")
    B, Sq, H, D = q.shape
    Sk = k.shape[1]
    if scale is None:
        scale = D ** -0.5
    dq = torch.empty((B, Sq, H, D), dtype=torch.bfloat16, device=q.device)
    dk = torch.empty((B, Sk, H, D), dtype=torch.bfloat16, device=q.device)
    dv = torch.empty((B, Sk, H, D), dtype=torch.bfloat16, device=q.device)
    ws = torch.empty((B * Sq * H * D + B * H * Sq,), dtype=torch.float32, device=q.device)
    a = AttentionBwdArgs()
    a.Q, a.K, a.V, a.O, a.dO = q.data_ptr(), k.data_ptr(), v.data_ptr(), o.data_ptr(), do.data_ptr()
    a.lse2, a.dQ, a.dK, a.dV, a.workspace = lse2.data_ptr(), dq.data_ptr(), dk.data_ptr(), dv.data_ptr(), ws.data_ptr()
    a.B, a.Sq, a.Sk, a.H, a.D = B, Sq, Sk, H, D
    st = []
    for t in (q, k, v, o, do, dq, dk, dv):
        st += [t.stride(0), t.stride(1), t.stride(2)]
    a.strides = (ctypes.c_int64 * 24)(*st)
    a.scale, a.causal = float(scale), int(bool(causal))
    with torch.cuda.device(q.device):
        rc = load().b200_attention_bwd(ctypes.byref(a), _stream(q))
    _check(rc, "b200_attention_bwd")
    return dq, dk, dv


def ffn_fwd(x, w1, b1, w2, b2, *, act=EPI_BIAS_GELU_TANH, h=None, out=None, out_dtype=torch.bfloat16):
    """y = act(x @ w1 + b1) @ w2 (+ b2) for one device's shard (w1 [d_model,F], w2 [F,d_model])."""
    _require_cuda(x, w1, b1, w2, b2, h, out)
    T, d_model = x.shape
    F = w1.shape[1]
    for t in (x, w1, w2):
        if t.dtype != torch.bfloat16 or not t.is_contiguous():
            raise B200Error("ffn_fwd takes contiguous bf16 tensors")
    if w1.shape != (d_model, F) or w2.shape != (F, d_model):
        raise B200Error("ffn_fwd: weight shapes do not match x")
    if h is None:
        h = torch.empty((T, F), dtype=torch.bfloat16, device=x.device)
    if out is None:
        out = torch.empty((T, d_model), dtype=out_dtype, device=x.device)
    with torch.cuda.device(x.device):
        rc = load().b200_ffn_fwd(
            _ptr(x), _ptr(w1), _ptr(b1), _ptr(w2), _ptr(b2), _ptr(h), _ptr(out),
            F32 if out.dtype == torch.float32 else BF16, T, d_model, F, act, _stream(x))
    _check(rc, "b200_ffn_fwd")
    return out


def cast_f32_to_bf16(src, out=None):
    _require_cuda(src, out)
    src = src.contiguous()
    if out is None:
        out = torch.empty(src.shape, dtype=torch.bfloat16, device=src.device)
    with torch.cuda.device(src.device):
        rc = load().b200_cast_f32_to_bf16(_ptr(src), _ptr(out), src.numel(), _stream(src))
    _check(rc, "b200_cast_f32_to_bf16")
    return out


def cast_bf16_to_f32(src, out=None):
    _require_cuda(src, out)
    src = src.contiguous()
    if out is None:
        out = torch.empty(src.shape, dtype=torch.float32, device=src.device)
    with torch.cuda.device(src.device):
        rc = load().b200_cast_bf16_to_f32(_ptr(src), _ptr(out), src.numel(), _stream(src))
    _check(rc, "b200_cast_bf16_to_f32")
    return out


def sum_partials_to_bf16(partials, bias=None, out=None):
    """partials: f32 [R, rows, cols] (contiguous) -> bf16 [rows, cols] = bf16(sum_r) (+ bias)."""
    _require_cuda(partials, bias, out)
    R, rows, cols = partials.shape
    if out is None:
        out = torch.empty((rows, cols), dtype=torch.bfloat16, device=partials.device)
    with torch.cuda.device(partials.device):
        rc = load().b200_sum_partials_to_bf16(
            _ptr(partials), rows * cols, R, _ptr(bias), cols, _ptr(out), rows * cols, _stream(partials))
    _check(rc, "b200_sum_partials_to_bf16")
    return out


# ---------------------------------------------------------------- fused GEMM -> reduce-scatter plumbing
def gemm_bf16_scatter(a, b, block_ptrs, rows_per_block: int, rot: int, ldc: int, *, b_layout=B_KN, a_block_stride: int = 0,
                      M: int = 0, tune=None, a_view=None, interleave: bool = False, progress=None):
    """a[M,K] @ b -> bf16 rows scattered to ``block_ptrs`` (a CUDA int64 tensor of M/rows_per_block base
    pointers, e.g. peer GPUs' receive slots): b200_gemm_bf16_ex with the row-block scatter.  With ``a_block_stride`` > 0
    block j of the product is read from rows [j * a_block_stride, j * a_block_stride + rows_per_block) of ``a`` (``M`` =
    logical rows of the product): one token chunk of every owner's row block.  ``a_view`` = (rows, inner, inner_stride,
    outer_stride) is the two-level view of b200_gemm_args (reduce-scatter along the sequence dim of a [batch, seq] input)."""
    _require_cuda(a, b, block_ptrs)
    K = a.shape[1]
    M = M or a.shape[0]
    N = b.shape[1] if b_layout == B_KN else b.shape[0]
    args = GemmArgs()
    args.A, args.lda = a.data_ptr(), a.stride(0)
    args.B, args.ldb, args.b_layout = b.data_ptr(), b.stride(0), b_layout
    args.C, args.ldc, args.out_dtype = None, ldc, BF16
    args.c_block_ptrs = block_ptrs.data_ptr()
    args.m_rows_per_block, args.m_block_rot = rows_per_block, rot
    args.m_block_interleave = int(bool(interleave))
    if progress is not None:       # (counters tensor, owners' flag-array pointers tensor, sender index, groups, epoch)
        counters, flag_ptrs, sender, groups, epoch = progress
        args.done_counters, args.done_flag_ptrs = counters.data_ptr(), flag_ptrs.data_ptr()
        args.done_sender, args.done_groups, args.done_epoch = int(sender), int(groups), int(epoch) & 0xFFFFFFFF
    if a_view is not None:      # two-level view: (rows_per_block, blocks_inner, inner_stride, outer_stride)
        args.a_rows_per_block, args.a_blocks_inner, args.a_block_stride, args.a_outer_stride = (int(v) for v in a_view)
    elif a_block_stride:
        args.a_rows_per_block, args.a_block_stride = rows_per_block, a_block_stride
    args.M, args.N, args.K, args.epilogue = M, N, K, EPI_NONE
    args.tune = _tuning(tune)
    with torch.cuda.device(a.device):
        rc = load().b200_gemm_bf16_ex(ctypes.byref(args), _stream(a))
    _check(rc, "b200_gemm_bf16_ex")


def comm_unique_id() -> bytes:
    """128-byte id for mesh_create (rank 0 makes it, every rank passes the same bytes)."""
    buf = ctypes.create_string_buffer(128)
    _check(load().b200_comm_unique_id(buf), "b200_comm_unique_id")
    return buf.raw


class MeshComm:
    """b200_mesh_* / b200_all_reduce ... of include/b200_sharded.h on torch tensors: a device mesh with one NCCL
    communicator per mesh axis, for one process per GPU.  Collective at construction and at the first use of an axis."""

    def __init__(self, unique_id: bytes, world: int, rank: int, shape):
        self.shape = tuple(int(v) for v in shape)
        arr = (ctypes.c_int * len(self.shape))(*self.shape)
        self._mesh = ctypes.c_void_p()
        _check(load().b200_mesh_create(ctypes.create_string_buffer(unique_id, 128), world, rank, arr, len(self.shape),
                                       ctypes.byref(self._mesh)), "b200_mesh_create")
        self._comms = {}

    def _comm(self, axis: int):
        if axis not in self._comms:
            c = ctypes.c_void_p()
            _check(load().b200_mesh_axis_comm(self._mesh, axis, ctypes.byref(c)), "b200_mesh_axis_comm")
            self._comms[axis] = c
        return self._comms[axis]

    def axis_size(self, axis: int) -> int:
        return load().b200_comm_size(self._comm(axis))

    def axis_rank(self, axis: int) -> int:
        return load().b200_comm_rank(self._comm(axis))

    @staticmethod
    def _dt(t):
        if t.dtype == torch.bfloat16:
            return BF16
        if t.dtype == torch.float32:
            return F32
        raise B200Error("mesh collectives take bf16 or fp32 tensors")

    def all_reduce(self, t, axis: int):
        """In-place sum of ``t`` over the mesh axis."""
        _require_cuda(t)
        _check(load().b200_all_reduce(self._comm(axis), _ptr(t), _ptr(t), t.numel(), self._dt(t), _stream(t)), "b200_all_reduce")
        return t

    def reduce_scatter(self, t, axis: int):
        """``t`` = axis_size consecutive blocks along dim 0; returns the sum of everybody's block [axis_rank]."""
        _require_cuda(t)
        n = self.axis_size(axis)
        out = torch.empty((t.shape[0] // n,) + tuple(t.shape[1:]), dtype=t.dtype, device=t.device)
        _check(load().b200_reduce_scatter(self._comm(axis), _ptr(t), _ptr(out), out.numel(), self._dt(t), _stream(t)),
               "b200_reduce_scatter")
        return out

    def all_gather(self, t, axis: int):
        """Concatenation of the ranks' ``t`` along dim 0, in axis order."""
        _require_cuda(t)
        n = self.axis_size(axis)
        out = torch.empty((t.shape[0] * n,) + tuple(t.shape[1:]), dtype=t.dtype, device=t.device)
        _check(load().b200_all_gather(self._comm(axis), _ptr(t), _ptr(out), t.numel(), self._dt(t), _stream(t)), "b200_all_gather")
        return out

    def all_to_all(self, t, axis: int):
        """Block p of ``t`` (dim 0 split in axis_size blocks) goes to rank p; returns the blocks received, in rank order."""
        _require_cuda(t)
        n = self.axis_size(axis)
        out = torch.empty_like(t)
        _check(load().b200_all_to_all(self._comm(axis), _ptr(t), _ptr(out), t.numel() // n, self._dt(t), _stream(t)),
               "b200_all_to_all")
        return out

    def ffn_fwd_tp(self, axis: int, x, w1, b1, w2, b2, *, act=EPI_BIAS_GELU_TANH, reduce_scatter=False, chunks=1):
        """b200_ffn_fwd_tp: this rank's W1 columns / W2 rows, the sum over the mesh axis inside the call.  Returns y
        [T, d_model] (all-reduce) or [T / axis_size, d_model] (reduce-scatter: this rank's sequence block)."""
        _require_cuda(x, w1, b1, w2, b2)
        T, d_model = x.shape
        F = w1.shape[1]
        n = self.axis_size(axis)
        h = torch.empty((T, F), dtype=torch.bfloat16, device=x.device)
        partial = torch.empty((T, d_model), dtype=torch.float32, device=x.device)
        y = torch.empty((T // n if reduce_scatter else T, d_model), dtype=torch.bfloat16, device=x.device)
        with torch.cuda.device(x.device):
            rc = load().b200_ffn_fwd_tp(self._comm(axis), _ptr(x), _ptr(w1), _ptr(b1), _ptr(w2), _ptr(b2), _ptr(h),
                                        _ptr(partial), _ptr(y), 1 if reduce_scatter else 0, T, d_model, F, act, chunks,
                                        _stream(x))
        _check(rc, "b200_ffn_fwd_tp")
        return y

    def destroy(self):
        for c in self._comms.values():
            load().b200_comm_release(c)
        self._comms = {}
        if self._mesh:
            _check(load().b200_mesh_destroy(self._mesh), "b200_mesh_destroy")
            self._mesh = ctypes.c_void_p()


def peer_alloc(nbytes: int):
    """(device pointer, 64-byte IPC handle) of a zero-filled buffer other ranks can map. Set-up only."""
    ptr = ctypes.c_void_p()
    handle = ctypes.create_string_buffer(64)
    _check(load().b200_peer_alloc(nbytes, ctypes.byref(ptr), handle), "b200_peer_alloc")
    return int(ptr.value), handle.raw


def peer_open(handle: bytes) -> int:
    ptr = ctypes.c_void_p()
    buf = ctypes.create_string_buffer(handle, 64)
    _check(load().b200_peer_open(buf, ctypes.byref(ptr)), "b200_peer_open")
    return int(ptr.value)


def peer_barrier(flag_ptrs, rank: int, world: int, epoch: int, stream=None):
    st = _stream(flag_ptrs) if stream is None else ctypes.c_void_p(stream.cuda_stream)
    with torch.cuda.device(flag_ptrs.device):
        rc = load().b200_peer_barrier(_ptr(flag_ptrs), rank, world, epoch & 0xFFFFFFFF, st)
    _check(rc, "b200_peer_barrier")


def peer_progressive_sum(flags_ptr: int, epoch: int, senders: int, groups: int, rows_per_group: int, src_ptr: int,
                         src_stride: int, bias, cols: int, out, ctas: int, stream):
    """b200_peer_progressive_sum: out rows of group g = bf16(sum over senders) (+ bias) as soon as every sender's progress
    flag for g shows `epoch`; enqueued on ``stream`` to run beside the pushing GEMM."""
    _require_cuda(bias, out)
    with torch.cuda.device(out.device):
        rc = load().b200_peer_progressive_sum(ctypes.c_void_p(flags_ptr), epoch & 0xFFFFFFFF, senders, groups, rows_per_group,
                                              ctypes.c_void_p(src_ptr), src_stride, _ptr(bias), cols, _ptr(out), ctas,
                                              ctypes.c_void_p(stream.cuda_stream))
    _check(rc, "b200_peer_progressive_sum")


def peer_push(dst_bases, world: int, rank: int, dst_off: int, src: torch.Tensor, ctas: int, stream):
    """This rank's contiguous shard ``src`` into offset ``dst_off`` (bytes) of every rank's gathered buffer (dst_bases: CUDA
    int64 tensor of base pointers), enqueued on ``stream`` -- b200_peer_push."""
    _require_cuda(dst_bases, src)
    if not src.is_contiguous():
        raise B200Error("peer_push takes a contiguous source")
    with torch.cuda.device(src.device):
        rc = load().b200_peer_push(_ptr(dst_bases), world, rank, dst_off, _ptr(src), src.numel() * src.element_size(), ctas,
                                   ctypes.c_void_p(stream.cuda_stream))
    _check(rc, "b200_peer_push")


def sum_bf16_partials(src_ptr: int, src_stride: int, nsrc: int, bias, cols: int, out):
    """out[rows, cols] = bf16(sum_r partial_r) (+bias); partials are raw device memory at src_ptr."""
    _require_cuda(bias, out)
    with torch.cuda.device(out.device):
        rc = load().b200_sum_bf16_partials(ctypes.c_void_p(src_ptr), src_stride, nsrc, _ptr(bias), cols, _ptr(out),
                                           out.numel(), _stream(out))
    _check(rc, "b200_sum_bf16_partials")
    return out


def peer_copy2d(dst_ptr: int, dst_pitch: int, src_ptr: int, src_pitch: int, width_bytes: int, rows: int, stream):
    """Strided D2D / peer copy on the copy engines, enqueued on `stream` (a torch.cuda.Stream)."""
    rc = load().b200_peer_copy2d(ctypes.c_void_p(dst_ptr), dst_pitch, ctypes.c_void_p(src_ptr), src_pitch, width_bytes,
                                 rows, ctypes.c_void_p(stream.cuda_stream))
    _check(rc, "b200_peer_copy2d")


def peer_gather(src_bases, world: int, rank: int, src_off: int, src_pitch: int, dst_ptr: int, dst_pitch: int, rows: int,
                width_bytes: int, ctas: int, stream):
    """One-kernel pull of a column chunk of every rank's shard (src_bases: CUDA int64 tensor of base pointers)."""
    rc = load().b200_peer_gather(_ptr(src_bases), world, rank, src_off, src_pitch, ctypes.c_void_p(dst_ptr), dst_pitch,
                                 rows, width_bytes, ctas, ctypes.c_void_p(stream.cuda_stream))
    _check(rc, "b200_peer_gather")
