"""Generates tests/golden/*.pt: outputs of the CPU oracle (oracle/reference_math.py) on seeded inputs.

The reference cannot be run in this image (no jax / flax), so these fixtures do NOT come from the
reference itself -- parity is unpinned (see the oracle's header).  They pin the ORACLE: any change
to its arithmetic shows up as a fixture mismatch, and the GPU tests compare the CUDA path with
the same values without recomputing them.  Inputs are regenerated from their seeds; only outputs
(small bf16/f32 tensors) are stored.

    python tests/golden/make_golden.py
"""
import os
import sys

import torch

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, os.path.dirname(os.path.dirname(HERE)))
from oracle import reference_math as rm  # noqa: E402


def case6_literal_inputs():
    """B8,S256,M640, 8 heads x 64 (case6_attention.py:149-151,43-45); bias non-zero so that path is exercised."""
    x = rm.synthetic((8, 256, 640), 100)
    params = {
        "to_q": rm.synthetic((640, 512), 101, "weight"),
        "to_k": rm.synthetic((640, 512), 102, "weight"),
        "to_v": rm.synthetic((640, 512), 103, "weight"),
        "to_out_0": (rm.synthetic((512, 640), 104, "weight"), rm.synthetic((640,), 105, "bias")),
    }
    return x, params


def ffn_small_inputs():
    x = rm.synthetic((4, 64, 256), 200)
    return (x, rm.synthetic((256, 1024), 201, "weight"), rm.synthetic((1024,), 202, "bias"),
            rm.synthetic((1024, 256), 203, "weight"), rm.synthetic((256,), 204, "bias"))


def dot_inputs():
    g = torch.Generator().manual_seed(0)
    return torch.randn(4, 16, generator=g), torch.randn(16, 4, generator=g)


def attention_inputs(d):
    shape = (2, 192, 4, d)
    return tuple(rm.synthetic(shape, 300 + i + d).to(torch.bfloat16) for i in range(3))


def main():
    out = {}
    a, b = dot_inputs()
    out["dot_4x16x4"] = rm.dot_f32(a, b)
    g = torch.Generator().manual_seed(0)
    out["einsum_8x4x16x4"] = rm.einsum_abc_acd_abd(torch.randn(8, 4, 16, generator=g), torch.randn(8, 16, 4, generator=g))
    x, params = case6_literal_inputs()
    # the full outputs are 2.6 MB / 2 MB: keep a strided sample (every 2nd batch row, every 4th token)
    out["case6_literal_y"] = rm.mha_forward(x, params, 8, 64)[::2, ::4].clone()
    out["case6_literal_q"] = rm.dense_bf16(x, params["to_q"])[::2, ::4].clone()    # = case5's to_q projection
    fx = ffn_small_inputs()
    out["ffn_gelu"] = rm.ffn_forward(*fx, act="gelu_tanh")
    out["ffn_relu"] = rm.ffn_forward(*fx, act="relu")
    for d in (64, 128):
        q, k, v = attention_inputs(d)
        out[f"attn_d{d}"] = rm.attention_core(q, k, v)
        out[f"attn_d{d}_causal"] = rm.attention_core(q, k, v, causal=True)
    torch.save(out, os.path.join(HERE, "oracle_outputs.pt"))
    for k, v in out.items():
        print(k, tuple(v.shape), v.dtype)


if __name__ == "__main__":
    main()
