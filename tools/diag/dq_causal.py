"""Where does the causal dQ error sit?  (bring-up diagnostic, run on the GPU box)"""
import os, sys
import torch
ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)
import learning_jax_sharding_b200 as ljs
from oracle import reference_math as rm
from tests import _tol

def run(B, S, H, D, causal, seed=90):
    q = rm.synthetic((B, S, H, D), seed).to(torch.bfloat16); k = rm.synthetic((B, S, H, D), seed + 1).to(torch.bfloat16)
    v = rm.synthetic((B, S, H, D), seed + 2).to(torch.bfloat16); do = rm.synthetic((B, S, H, D), seed + 3).to(torch.bfloat16)
    dq_w, dk_w, dv_w, _ = rm.attention_core_grads(q, k, v, do, causal=causal)
    outs = []
    for rep in range(3):
        o, lse = ljs.capi.attention_fwd_lse(q.cuda(), k.cuda(), v.cuda(), causal=causal)
        dq, dk, dv = ljs.capi.attention_bwd(q.cuda(), k.cuda(), v.cuda(), o, do.cuda(), lse, causal=causal)
        outs.append((dq.cpu().float(), dk.cpu().float(), dv.cpu().float()))
    print(f"== B{B} S{S} H{H} D{D} causal={causal}: deterministic dq={all(torch.equal(outs[0][0], x[0]) for x in outs)} "
          f"dk={all(torch.equal(outs[0][1], x[1]) for x in outs)}")
    for name, got, want in (("dq", outs[0][0], dq_w.float()), ("dk", outs[0][1], dk_w.float()), ("dv", outs[0][2], dv_w.float())):
        rms = _tol.row_rms(want).float()
        exc = ((got - want).abs() - 2e-2 * want.abs()).clamp_min(0) / (2 ** -8 * rms)
        per_row = exc.amax(dim=-1)            # [B, S, H]
        top = torch.topk(per_row.flatten(), 8)
        idx = [tuple(int(x) for x in torch.unravel_index(i, per_row.shape)) for i in top.indices]
        print(f"  {name}: c_needed={exc.max():.2f}; worst rows (b, s, h): " + ", ".join(f"{i}:{v:.1f}" for i, v in zip(idx, top.values)))
        b, s, h = idx[0]
        print(f"     row {idx[0]} rms_row={rms[b, s, h, 0]:.4f} got[:6]={got[b, s, h, :6].tolist()} want[:6]={want[b, s, h, :6].tolist()}")
        hist = torch.histc(per_row.flatten().clamp(max=24), bins=8, min=0, max=24)
        print(f"     per-row c histogram (0..24 in 8 bins): {hist.int().tolist()}")

for args in [(2, 256, 2, 128, True), (2, 256, 2, 128, False), (1, 1024, 2, 128, True), (1, 1024, 1, 64, True), (1, 4096, 1, 128, True)]:
    run(*args)
