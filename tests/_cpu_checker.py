"""TEST INFRASTRUCTURE: a stand-in for ``capi`` (same function names and semantics) built on the
CPU oracle, so that the collective PLANS of the host layer can be exercised on a GPU-less box
(virtual runtime on cpu, gloo world_size 2).  The product never imports this."""
import torch

from oracle import reference_math as rm

BF16, F32 = torch.bfloat16, torch.float32
EPI_NONE, EPI_BIAS, EPI_BIAS_RELU, EPI_BIAS_GELU_TANH, EPI_RELU, EPI_GELU_TANH = range(6)
B_KN, B_NK = 0, 1


def _r(x):
    return x.to(BF16).to(F32)


def gemm_bf16(a, b, *, b_layout=B_KN, bias=None, epilogue=EPI_NONE, out=None, out_dtype=BF16):
    bb = b.to(F32) if b_layout == B_KN else b.to(F32).t()
    acc = a.to(F32) @ bb
    if out_dtype == F32 and epilogue == EPI_NONE:
        res = acc
    else:
        v = _r(acc)
        if epilogue in (EPI_BIAS, EPI_BIAS_RELU, EPI_BIAS_GELU_TANH):
            v = _r(v + bias.to(F32))
        if epilogue in (EPI_BIAS_RELU, EPI_RELU):
            v = torch.relu(v)
        if epilogue in (EPI_BIAS_GELU_TANH, EPI_GELU_TANH):
            v = _r(rm.gelu_tanh(v))
        res = v.to(out_dtype)
    if out is not None:
        out.copy_(res)
        return out
    return res


def gemm_f32(a, b, *, out=None, accumulate=False):
    res = (a.double() @ b.double()).float()
    if out is not None:
        out.copy_(out + res if accumulate else res)
        return out
    return res


def attention_fwd(q, k, v, *, scale=None, causal=False, out=None):
    return rm.attention_core(q, k, v, scale, causal)


def cast_f32_to_bf16(src, out=None):
    return src.to(BF16)


def cast_bf16_to_f32(src, out=None):
    return src.to(F32)


def sum_partials_to_bf16(partials, bias=None, out=None):
    v = _r(partials.sum(dim=0))
    if bias is not None:
        v = _r(v + bias.to(F32))
    return v.to(BF16)


def gemm_bf16_tn(a_km, b, *, out=None, out_dtype=BF16):
    acc = a_km.to(F32).t() @ b.to(F32)
    res = acc if out_dtype == F32 else _r(acc).to(out_dtype)
    if out is not None:
        out.copy_(res)
        return out
    return res


def adam_step(param, grad, mu, nu, *, step, lr=1e-3, b1=0.9, b2=0.999, eps=1e-8):
    p2, mu2, nu2 = rm.adam_update(param, grad, mu, nu, step, lr=lr, b1=b1, b2=b2, eps=eps)
    param.copy_(p2)
    mu.copy_(mu2)
    nu.copy_(nu2)
    return param


def attention_fwd_lse(q, k, v, *, scale=None, causal=False):
    o = rm.attention_core(q, k, v, scale, causal)
    _, _, _, lse = rm.attention_core_grads(q, k, v, torch.zeros_like(q), scale, causal)
    return o, lse


def attention_bwd(q, k, v, o, do, lse2, *, scale=None, causal=False):
    dq, dk, dv, _ = rm.attention_core_grads(q, k, v, do, scale, causal)
    return dq, dk, dv


def attention_combine(acc, lse_acc, o_blk, lse_blk, *, first):
    """b200_attention_combine on the host (f64 inside)."""
    lb = lse_blk.double()
    if first:
        acc.copy_(o_blk.to(F32))
        lse_acc.copy_(lse_blk)
        return acc
    la = lse_acc.double()
    ln = torch.log2(torch.exp2(la) + torch.exp2(lb))
    w_old = torch.exp2(la - ln).permute(0, 2, 1).unsqueeze(-1)
    w_new = torch.exp2(lb - ln).permute(0, 2, 1).unsqueeze(-1)
    acc.copy_((acc.double() * w_old + o_blk.double() * w_new).to(F32))
    lse_acc.copy_(ln.to(F32))
    return acc
