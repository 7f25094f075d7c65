"""Attention forward / backward timings at the bench shapes (bring-up diagnostic, one GPU).  usage: attn_time.py [fwd|bwd|all]"""
import os, sys
import torch
ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)
import learning_jax_sharding_b200 as ljs
capi = ljs.capi
what = sys.argv[1] if len(sys.argv) > 1 else "all"


def timed(fn, iters=10, warm=3):
    for _ in range(warm):
        fn()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(iters):
        fn()
    e1.record(); torch.cuda.synchronize()
    return e0.elapsed_time(e1) / iters


for (B, S, H, D, causal) in [(8, 4096, 32, 128, False), (8, 4096, 32, 128, True), (8, 8192, 8, 128, True), (8, 4096, 32, 64, False),
                             (8, 4096, 32, 64, True)]:
    q, k, v, do = (torch.randn(B, S, H, D, device="cuda").to(torch.bfloat16) for _ in range(4))
    fl = 2.0 * B * H * S * (S + 1) * D if causal else 4.0 * B * H * S * S * D
    line = f"B{B} S{S} H{H} D{D} causal={int(causal)}:"
    if what in ("fwd", "all"):
        o = torch.empty_like(q)
        for rep in range(2):
            ms = timed(lambda: capi.attention_fwd(q, k, v, causal=causal, out=o))
            line += f" fwd {ms:.3f} ms {fl / ms / 1e9:.0f} TFLOP/s |"
    if what in ("bwd", "all"):
        o, lse = capi.attention_fwd_lse(q, k, v, causal=causal)
        ms = timed(lambda: capi.attention_bwd(q, k, v, o, do, lse, causal=causal), iters=5, warm=2)
        line += f" bwd {ms:.3f} ms {2.5 * fl / ms / 1e9:.0f} TFLOP/s"
    print(line, flush=True)
    del q, k, v, do
