"""Interleaved A/B timing of two builds of the library on the attention forward (same process, same tensors, alternating
launches, so that clock / power drift hits both alike).  usage: attn_ab.py libA.so libB.so [more.so ...]"""
import ctypes, os, statistics, sys
import torch
ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)
import learning_jax_sharding_b200 as ljs
capi = ljs.capi

libs = []
for path in sys.argv[1:]:
    lib = ctypes.CDLL(os.path.abspath(path))
    lib.b200_attention_fwd_ex.argtypes = [ctypes.POINTER(capi.AttentionArgs), ctypes.c_void_p]
    lib.b200_attention_fwd_ex.restype = ctypes.c_int
    libs.append((os.path.basename(path), lib))


def call(lib, q, k, v, o, causal):
    a = capi.AttentionArgs()
    a.Q, a.K, a.V, a.O = q.data_ptr(), k.data_ptr(), v.data_ptr(), o.data_ptr()
    B, S, H, D = q.shape
    a.B, a.Sq, a.Sk, a.H, a.D = B, S, S, H, D
    st = []
    for t in (q, k, v, o):
        st += [t.stride(0), t.stride(1), t.stride(2)]
    a.strides = (ctypes.c_int64 * 12)(*st)
    a.scale, a.causal, a.variant = D ** -0.5, int(causal), 0
    rc = lib.b200_attention_fwd_ex(ctypes.byref(a), ctypes.c_void_p(torch.cuda.current_stream().cuda_stream))
    assert rc == 0


shapes = [(8, 4096, 32, 128, False), (8, 8192, 8, 128, True), (8, 4096, 32, 64, False), (8, 4096, 32, 64, True)]
for (B, S, H, D, causal) in shapes:
    q, k, v = (torch.randn(B, S, H, D, device="cuda").to(torch.bfloat16) for _ in range(3))
    o = torch.empty_like(q)
    fl = 2.0 * B * H * S * (S + 1) * D if causal else 4.0 * B * H * S * S * D
    times = {n: [] for n, _ in libs}
    for n, lib in libs:
        call(lib, q, k, v, o, causal)
    torch.cuda.synchronize()
    for rnd in range(8):
        for n, lib in libs:
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record()
            for _ in range(4):
                call(lib, q, k, v, o, causal)
            e1.record(); torch.cuda.synchronize()
            times[n].append(e0.elapsed_time(e1) / 4)
    print(f"B{B} S{S} H{H} D{D} causal={int(causal)}: " + " | ".join(
        f"{n}: {statistics.median(t):.3f} ms {fl / statistics.median(t) / 1e9:.0f} TFLOP/s (min {fl / min(t) / 1e9:.0f})" for n, t in times.items()),
        flush=True)
