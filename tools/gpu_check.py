"""Bring-up checks run on the GPU box (not part of the test suite): each invocation tests one
kernel configuration against a torch fp32 computation ON THE GPU and prints error statistics
and timings.  Run each configuration in its own process (a device trap poisons the context).

  python tools/gpu_check.py gemm --g 1 --layout kn
  python tools/gpu_check.py gemm_time
  python tools/gpu_check.py f32
"""
import argparse
import importlib.util
import os
import sys
import time

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
spec = importlib.util.spec_from_file_location("capi", os.path.join(ROOT, "learning-jax-sharding_b200", "capi.py"))
capi = importlib.util.module_from_spec(spec)
spec.loader.exec_module(capi)

# The library keeps no tuning state (overrides travel with each call: b200_gemm_args.tune, b200_attention_args.variant).
# This tool's sweeps were written as "set, then call": the setters below only edit TOOL-LOCAL dictionaries that the wrapped
# call functions pass along.
_TUNE, _ATTN = {}, {"variant": 0}
_gemm_bf16, _attention_fwd, _attention_fwd_lse = capi.gemm_bf16, capi.attention_fwd, capi.attention_fwd_lse
capi.gemm_bf16 = lambda *a, **k: _gemm_bf16(*a, **{"tune": dict(_TUNE), **k})
capi.attention_fwd = lambda *a, **k: _attention_fwd(*a, **{"variant": _ATTN["variant"], **k})
capi.attention_fwd_lse = lambda *a, **k: _attention_fwd_lse(*a, **{"variant": _ATTN["variant"], **k})
capi.set_gemm_tuning = lambda g=0, bn=0: _TUNE.update(cta_group=g, block_n=bn)
capi.set_gemm_group_m = lambda gm=0: _TUNE.update(group_m=gm)
capi.set_gemm_k_serpentine = lambda mode=-1: _TUNE.update(k_serpentine=1 if mode == 1 else 0)
capi.set_gemm_max_stages = lambda st=0: _TUNE.update(max_stages=st)
capi.set_attention_variant = lambda v=0: _ATTN.update(variant=v)     # 0, 1, 8, 9 (2-7 are archived in tools/variants/)


def ref_epilogue(acc, bias, epi):
    v = acc.to(torch.bfloat16).float()
    if epi in (capi.EPI_BIAS, capi.EPI_BIAS_RELU, capi.EPI_BIAS_GELU_TANH):
        v = (v + bias.float()).to(torch.bfloat16).float()
    if epi in (capi.EPI_BIAS_RELU, capi.EPI_RELU):
        v = torch.relu(v)
    if epi in (capi.EPI_BIAS_GELU_TANH, capi.EPI_GELU_TANH):
        v = torch.nn.functional.gelu(v, approximate="tanh")
    return v


def check_gemm(M, N, K, layout, epi=capi.EPI_NONE, out_dtype=torch.bfloat16, seed=0):
    g = torch.Generator(device="cuda").manual_seed(seed)
    a = torch.randn(M, K, device="cuda", generator=g).to(torch.bfloat16)
    if layout == "kn":
        b = (torch.randn(K, N, device="cuda", generator=g) / K ** 0.5).to(torch.bfloat16)
        ref = a.float() @ b.float()
    else:
        b = (torch.randn(N, K, device="cuda", generator=g) / K ** 0.5).to(torch.bfloat16)
        ref = a.float() @ b.float().t()
    bias = (torch.randn(N, device="cuda", generator=g) * 0.5).to(torch.bfloat16)
    out = capi.gemm_bf16(a, b, b_layout=capi.B_KN if layout == "kn" else capi.B_NK, bias=bias, epilogue=epi,
                         out_dtype=out_dtype)
    torch.cuda.synchronize()
    if out_dtype == torch.float32 and epi == capi.EPI_NONE:
        want = ref
    else:
        want = ref_epilogue(ref, bias, epi)
    got = out.float()
    err = (got - want).abs()
    denom = want.abs().clamp_min(1e-2)
    rel = (err / denom).max().item()
    bad = (err > 2e-2 * want.abs() + 2e-2).sum().item()
    print(f"  gemm M={M} N={N} K={K} layout={layout} epi={epi} out={str(out_dtype)[6:]}: "
          f"max_abs={err.max().item():.3e} max_rel={rel:.3e} bad={bad}/{got.numel()} "
          f"{'OK' if bad == 0 else 'FAIL'}", flush=True)
    if bad:
        idx = (err > 2e-2 * want.abs() + 2e-2).nonzero()[:8]
        for i, j in idx.tolist():
            print(f"    [{i},{j}] got={got[i, j].item():.5f} want={want[i, j].item():.5f}")
        # error structure: which rows / cols are wrong
        rows = (err > 2e-2 * want.abs() + 2e-2).any(dim=1).nonzero().flatten()
        cols = (err > 2e-2 * want.abs() + 2e-2).any(dim=0).nonzero().flatten()
        print(f"    bad rows: n={rows.numel()} first={rows[:6].tolist()} last={rows[-3:].tolist()}; "
              f"bad cols: n={cols.numel()} first={cols[:6].tolist()} last={cols[-3:].tolist()}")
    return bad == 0


def time_fn(fn, iters=20, warmup=3):
    for _ in range(warmup):
        fn()
    torch.cuda.synchronize()
    s, e = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    s.record()
    for _ in range(iters):
        fn()
    e.record()
    torch.cuda.synchronize()
    return s.elapsed_time(e) / iters


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("what")
    ap.add_argument("--g", type=int, default=0)
    ap.add_argument("--bn", type=int, default=0)
    ap.add_argument("--layout", default="kn")
    args = ap.parse_args()
    print(f"== {args.what} g={args.g} bn={args.bn} layout={args.layout} on {torch.cuda.get_device_name(0)}",
          flush=True)
    capi.load()
    ok = True
    if args.what == "gemm":
        capi.set_gemm_tuning(args.g, args.bn)
        shapes = [(128, 256, 64), (256, 256, 128), (256, 512, 256), (512, 640, 512), (2048, 512, 640),
                  (1000, 264, 136), (4096, 4096, 1024)]
        for (M, N, K) in shapes:
            ok &= check_gemm(M, N, K, args.layout)
        ok &= check_gemm(512, 768, 256, args.layout, epi=capi.EPI_BIAS)
        ok &= check_gemm(512, 768, 256, args.layout, epi=capi.EPI_BIAS_RELU)
        ok &= check_gemm(512, 768, 256, args.layout, epi=capi.EPI_BIAS_GELU_TANH)
        ok &= check_gemm(512, 768, 256, args.layout, epi=capi.EPI_NONE, out_dtype=torch.float32)
    elif args.what == "gemm_time":
        for (g, bn) in [(1, 256), (2, 256), (1, 128), (2, 128)]:
            capi.set_gemm_tuning(g, bn)
            for layout in ("kn", "nk"):
                for (M, N, K) in [(8192, 8192, 8192), (16384, 16384, 4096), (16384, 4096, 16384)]:
                    a = torch.randn(M, K, device="cuda").to(torch.bfloat16)
                    b = torch.randn((K, N) if layout == "kn" else (N, K), device="cuda").to(torch.bfloat16)
                    out = torch.empty(M, N, device="cuda", dtype=torch.bfloat16)
                    lay = capi.B_KN if layout == "kn" else capi.B_NK
                    ms = time_fn(lambda: capi.gemm_bf16(a, b, b_layout=lay, out=out))
                    print(f"  g={g} bn={bn} {layout} M={M} N={N} K={K}: {ms:.3f} ms "
                          f"{2 * M * N * K / ms / 1e9:.1f} TFLOP/s", flush=True)
        # cuBLAS comparator
        for (M, N, K) in [(8192, 8192, 8192), (16384, 16384, 4096), (16384, 4096, 16384)]:
            a = torch.randn(M, K, device="cuda").to(torch.bfloat16)
            b = torch.randn(K, N, device="cuda").to(torch.bfloat16)
            ms = time_fn(lambda: torch.matmul(a, b))
            print(f"  cublas M={M} N={N} K={K}: {ms:.3f} ms {2 * M * N * K / ms / 1e9:.1f} TFLOP/s", flush=True)
    elif args.what == "f32":
        for (Bt, M, N, K) in [(1, 4, 4, 16), (8, 4, 4, 16), (1, 2, 1, 16), (1, 257, 129, 77), (1, 1024, 1024, 1024)]:
            a = torch.randn(Bt, M, K, device="cuda")
            b = torch.randn(Bt, K, N, device="cuda")
            out = capi.gemm_f32(a, b)
            want = (a.double() @ b.double())
            err = ((out.double() - want).abs() / want.abs().clamp_min(1e-3)).max().item()
            good = torch.allclose(out.double(), want, rtol=1e-5, atol=1e-5 * K ** 0.5)
            print(f"  f32 B={Bt} M={M} N={N} K={K}: max_rel={err:.3e} {'OK' if good else 'FAIL'}", flush=True)
            ok &= good
    elif args.what == "attn":
        if args.g:
            capi.set_attention_variant(args.g)
        sys.path.insert(0, ROOT)
        from oracle import reference_math as rm
        for (B, Sq, Sk, H, D, causal) in [(1, 128, 128, 1, 64, False), (1, 128, 128, 1, 128, False),
                                          (1, 128, 256, 1, 128, False), (2, 256, 512, 2, 128, False),
                                          (1, 384, 384, 2, 64, True), (1, 200, 333, 2, 128, False),
                                          (1, 1024, 1024, 2, 128, True)]:
            if args.g == 9 and D != 128:
                continue
            q = rm.synthetic((B, Sq, H, D), 1).to(torch.bfloat16)
            k = rm.synthetic((B, Sk, H, D), 2).to(torch.bfloat16)
            v = rm.synthetic((B, Sk, H, D), 3).to(torch.bfloat16)
            o = capi.attention_fwd(q.cuda(), k.cuda(), v.cuda(), causal=causal)
            torch.cuda.synchronize()
            want = rm.attention_core(q, k, v, causal=causal).float()
            err = (o.cpu().float() - want).abs()
            bad = (err > 2e-2 * want.abs() + 2e-2)
            print(f"  attn B={B} Sq={Sq} Sk={Sk} H={H} D={D} causal={causal}: max_abs={err.max().item():.3e} "
                  f"bad={bad.sum().item()}/{bad.numel()} {'OK' if not bad.any() else 'FAIL'}", flush=True)
            if bad.any():
                rows = bad.any(dim=3).any(dim=2).any(dim=0).nonzero().flatten()
                cols = bad.any(dim=2).any(dim=1).any(dim=0).nonzero().flatten()
                print(f"    bad query rows n={rows.numel()} first={rows[:8].tolist()}; bad d n={cols.numel()} "
                      f"first={cols[:8].tolist()}; got[0,0,0,:4]={o[0,0,0,:4].tolist()} want={want[0,0,0,:4].tolist()}")
                ok = False
    elif args.what == "gemm_group":
        for (M, N, K) in [(16384, 16384, 4096), (16384, 4096, 16384)]:
            a = torch.randn(M, K, device="cuda").to(torch.bfloat16)
            b = torch.randn(K, N, device="cuda").to(torch.bfloat16)
            out = torch.empty(M, N, device="cuda", dtype=torch.bfloat16)
            for g in (1, 2):
                for gm in (2, 4, 8, 16, 32):
                    capi.set_gemm_tuning(g, 256)
                    capi.set_gemm_group_m(gm)
                    ms = time_fn(lambda: capi.gemm_bf16(a, b, out=out), iters=30)
                    print(f"  g={g} group_m={gm} M={M} N={N} K={K}: {ms:.3f} ms {2 * M * N * K / ms / 1e9:.1f} TFLOP/s",
                          flush=True)
            ms = time_fn(lambda: torch.matmul(a, b, out=out), iters=30)
            print(f"  cublas M={M} N={N} K={K}: {ms:.3f} ms {2 * M * N * K / ms / 1e9:.1f} TFLOP/s", flush=True)
    elif args.what == "gemm_stages":
        import ctypes
        lib = capi.load()
        import statistics
        for (M, N, K) in [(16384, 16384, 4096), (16384, 4096, 16384)]:
            a = torch.randn(M, K, device="cuda").to(torch.bfloat16)
            b = torch.randn(K, N, device="cuda").to(torch.bfloat16)
            out = torch.empty(M, N, device="cuda", dtype=torch.bfloat16)
            variants = [(2, 256, st) for st in (2, 3, 4, 5, 6)] + [(1, 256, st) for st in (2, 3, 4)] + [(0, 0, -1)]
            res = {v: [] for v in variants}
            for rnd in range(4):                       # interleave variants so clocks / thermals are shared
                for v in variants:
                    g, bn, st = v
                    if st < 0:
                        ms = time_fn(lambda: torch.matmul(a, b, out=out), iters=10, warmup=2)
                    else:
                        capi.set_gemm_tuning(g, bn)
                        capi.set_gemm_max_stages(st)
                        ms = time_fn(lambda: capi.gemm_bf16(a, b, out=out), iters=10, warmup=2)
                    res[v].append(ms)
            for v in variants:
                ms = statistics.median(res[v])
                name = "cublas" if v[2] < 0 else f"g={v[0]} bn={v[1]} stages={v[2]}"
                print(f"  {name:24s} M={M} N={N} K={K}: {ms:.3f} ms {2 * M * N * K / ms / 1e9:.1f} TFLOP/s", flush=True)
            capi.set_gemm_max_stages(0)
    elif args.what == "gemm_sweep2":
        import ctypes
        import statistics
        lib = capi.load()
        for (M, N, K) in [(16384, 16384, 4096), (16384, 4096, 16384), (32768, 12288, 4096), (8192, 8192, 8192)]:
            a = torch.randn(M, K, device="cuda").to(torch.bfloat16)
            b = torch.randn(K, N, device="cuda").to(torch.bfloat16)
            out = torch.empty(M, N, device="cuda", dtype=torch.bfloat16)
            variants = [(2, 256, 4, gm) for gm in (4, 8, 16, 32, 64)] + [(2, 256, 3, 16), (1, 256, 4, 16), (0, 0, -1, 0)]
            res = {v: [] for v in variants}
            for rnd in range(4):
                for v in variants:
                    g, bn, st, gm = v
                    if st < 0:
                        ms = time_fn(lambda: torch.matmul(a, b, out=out), iters=10, warmup=2)
                    else:
                        capi.set_gemm_tuning(g, bn)
                        capi.set_gemm_group_m(gm)
                        capi.set_gemm_max_stages(st)
                        ms = time_fn(lambda: capi.gemm_bf16(a, b, out=out), iters=10, warmup=2)
                    res[v].append(ms)
            for v in variants:
                ms = statistics.median(res[v])
                name = "cublas" if v[2] < 0 else f"g={v[0]} bn={v[1]} stages={v[2]} group_m={v[3]}"
                print(f"  {name:36s} M={M} N={N} K={K}: {ms:.3f} ms {2 * M * N * K / ms / 1e9:.1f} TFLOP/s", flush=True)
            capi.set_gemm_max_stages(0)
            capi.set_gemm_group_m(0)
    elif args.what == "attn_phases":
        # needs a build with B200_EXTRA_NVCC_FLAGS=-DB200_ATTN_PROFILE=1 (the hooks are compiled out by default)
        import ctypes
        lib = capi.load()
        lib.b200_attention_set_profile_buffer.argtypes = [ctypes.c_void_p]
        capi.set_attention_variant(args.g if args.g else 2)
        buf = torch.zeros(16, dtype=torch.int64, device="cuda")
        B, S, H, D = 8, 4096, 32, 128
        q, k, v = (torch.randn(B, S, H, D, device="cuda").to(torch.bfloat16) for _ in range(3))
        o = torch.empty_like(q)
        for _ in range(2):
            capi.attention_fwd(q, k, v, out=o)
        lib.b200_attention_set_profile_buffer(ctypes.c_void_p(buf.data_ptr()))
        capi.attention_fwd(q, k, v, out=o)
        torch.cuda.synchronize()
        lib.b200_attention_set_profile_buffer(None)
        t = buf.cpu().tolist()
        if args.g in (5, 7, 9):
            nt = max(t[7], 1)
            names5 = ["wait s_full", "tmem ld S + s_free", "mask + half max", "xchg barrier", "decide / rescale / P-buffer wait",
                      "exp", "tmem st P + arrive"]
            if args.g in (7, 9):
                names5[2:4] = ["mask + row max", "wait reference max of tile j-1"]
            print(f"  v{args.g} softmax warp 4 of block (0,0,0): {nt} KV tiles of this warp, cycles per tile:")
            for n, c in zip(names5, t[:7]):
                print(f"    {n:34s} {c / nt:8.0f}")
            print(f"    {'total':34s} {sum(t[:7]) / nt:8.0f}")
            if args.g == 9:
                print(f"    {'epilogues (per own tile)':34s} {t[11] / nt:8.0f}")
            print(f"  MMA warp: {t[8] / nt:.0f} cycles per tile in the loop, {t[9] / nt:.0f} issuing, {t[10] / nt:.1f} polls per tile",
                  flush=True)
            print("RESULT PASS")
            return
        nt = max(t[6], 1)
        names = (["wait s_full", "tmem ld S", "mask+max+decide", "rescale O (if any)", "exp (P in regs)", "tmem st P + arrive"]
                 if args.g == 4 else
                 ["wait s_full", "tmem ld S", "mask+max+decide", "wait pv_done", "exp + P store", "fence + arrive"])
        print(f"  softmax warp 4 of block (0,0,0): {nt} KV tiles, cycles per tile:")
        for n, c in zip(names, t[:6]):
            print(f"    {n:18s} {c / nt:8.0f}")
        print(f"    {'total':18s} {sum(t[:6]) / nt:8.0f}", flush=True)
    elif args.what == "gemm_traffic":
        # run under: ncu --metrics dram__bytes_read.sum,dram__bytes_write.sum,gpu__time_duration.sum -k regex:gemm_bf16
        # launches, in order: for each of the two FFN GEMM shapes, (group_m, k_serpentine) combinations
        for (M, N, K, combos) in [(16384, 16384, 4096, ((16, 0), (16, 1), (8, 1), (32, 1))),
                                  (16384, 4096, 16384, ((8, 0), (8, 1), (4, 1), (16, 1), (64, 1)))]:
            a = torch.randn(M, K, device="cuda").to(torch.bfloat16)
            b = torch.randn(K, N, device="cuda").to(torch.bfloat16)
            out = torch.empty(M, N, device="cuda", dtype=torch.bfloat16)
            for (gm, ks) in combos:
                capi.set_gemm_group_m(gm)
                capi.set_gemm_k_serpentine(ks)
                capi.gemm_bf16(a, b, out=out)
                torch.cuda.synchronize()
                print(f"  launched M={M} N={N} K={K} group_m={gm} k_serpentine={ks}", flush=True)
        capi.set_gemm_group_m(0)
        capi.set_gemm_k_serpentine(-1)
    elif args.what == "gemm_serp":
        # timing of the two FFN GEMM shapes with / without the serpentine K order, interleaved
        import statistics
        for (M, N, K, combos) in [(16384, 16384, 4096, ((16, 0), (16, 1), (8, 1))),
                                  (16384, 4096, 16384, ((8, 0), (8, 1), (16, 1), (64, 1)))]:
            a = torch.randn(M, K, device="cuda").to(torch.bfloat16)
            b = torch.randn(K, N, device="cuda").to(torch.bfloat16)
            out = torch.empty(M, N, device="cuda", dtype=torch.bfloat16)
            res = {c: [] for c in combos}
            for rnd in range(4):
                for c in combos:
                    capi.set_gemm_group_m(c[0])
                    capi.set_gemm_k_serpentine(c[1])
                    res[c].append(time_fn(lambda: capi.gemm_bf16(a, b, out=out), iters=10, warmup=2))
            for c in combos:
                ms = statistics.median(res[c])
                print(f"  M={M} N={N} K={K} group_m={c[0]} k_serpentine={c[1]}: {ms:.3f} ms {2.0 * M * N * K / ms / 1e9:.1f} TFLOP/s", flush=True)
            # the serpentine order only permutes the f32 summation: same result within bf16 rounding
            capi.set_gemm_k_serpentine(0)
            ref = capi.gemm_bf16(a, b).float()
            capi.set_gemm_k_serpentine(1)
            got = capi.gemm_bf16(a, b).float()
            d = (got - ref).abs().max().item()
            print(f"  max |serpentine - plain| = {d:.3e} (|C| max {ref.abs().max().item():.1f})", flush=True)
        capi.set_gemm_group_m(0)
        capi.set_gemm_k_serpentine(-1)
    elif args.what == "attn_prof":
        q, k, v = (torch.randn(2, 4096, 16, 128, device="cuda").to(torch.bfloat16) for _ in range(3))
        for _ in range(3):
            capi.attention_fwd(q, k, v)
        torch.cuda.synchronize()
    elif args.what == "attn_clock":
        # sustained run of each variant (and cuDNN SDPA) with NVML clock / power sampling: converts the measured time
        # into SM cycles per (CTA, KV tile), the unit the clock64 phase profiles are in
        sys.path.insert(0, ROOT)
        from bench import ClockSampler
        B, S, H, D = 8, 4096, 32, 128
        q, k, v = (torch.randn(B, S, H, D, device="cuda").to(torch.bfloat16) for _ in range(3))
        o = torch.empty_like(q)
        qt, kt, vt = (t.transpose(1, 2) for t in (q, k, v))
        fl = 4.0 * B * H * S * S * D
        ctas = B * H * (S // 128)
        waves = -(-ctas // torch.cuda.get_device_properties(0).multi_processor_count)
        for name in ["v8", "v9", "sdpa", "v9", "v8", "sdpa"]:
            if name == "sdpa":
                fn = lambda: torch.nn.functional.scaled_dot_product_attention(qt, kt, vt)
            else:
                capi.set_attention_variant(int(name[1:]))
                fn = lambda: capi.attention_fwd(q, k, v, out=o)
            for _ in range(20):
                fn()
            torch.cuda.synchronize()
            cs = ClockSampler(0)
            cs.start()
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            iters = 400
            e0.record()
            for _ in range(iters):
                fn()
            e1.record()
            torch.cuda.synchronize()
            c = cs.stop()
            ms = e0.elapsed_time(e1) / iters
            mhz = c.get("sm_mhz") or 0
            cyc = ms * 1e-3 * mhz * 1e6 / (waves * (S // 128)) if name != "sdpa" else float("nan")
            print(f"  {name:5s} {ms:.3f} ms {fl / ms / 1e9:7.1f} TFLOP/s  sm {mhz} MHz (min {c.get('sm_min_mhz')}) "
                  f"power max {c.get('power_w_max')} W {c.get('reasons')}  -> {cyc:.0f} cycles per CTA-tile", flush=True)
        capi.set_attention_variant(0)
    elif args.what == "train_time":
        # the two kernels of the Dense train_step at the case5 config size: weight-gradient GEMM (x^T . dy) and Adam
        T, M, N = 32768, 4096, 4096
        x = torch.randn(T, M, device="cuda").to(torch.bfloat16)
        dy = torch.randn(T, N, device="cuda").to(torch.bfloat16)
        out = torch.empty(M, N, device="cuda", dtype=torch.float32)
        ms = time_fn(lambda: capi.gemm_bf16_tn(x, dy, out=out), iters=10, warmup=3)
        print(f"  dW = x^T . dy  T={T} M={M} N={N}: {ms:.3f} ms {2.0 * T * M * N / ms / 1e9:.1f} TFLOP/s", flush=True)
        xt = x.t().contiguous()
        ms2 = time_fn(lambda: capi.gemm_bf16(xt, dy, out=out), iters=10, warmup=3)
        print(f"  same product from a pre-transposed x (K-major A): {ms2:.3f} ms {2.0 * T * M * N / ms2 / 1e9:.1f} TFLOP/s", flush=True)
        want = torch.matmul(x.float().t()[:256], dy.float())
        err = (out[:256] - want).abs().max().item()
        print(f"  max |err| on 256 rows vs fp32 matmul: {err:.3e}", flush=True)
        ok &= err < 0.5
        n = M * N
        p, mu, nu = (torch.randn(n, device="cuda") for _ in range(3))
        nu.abs_()
        g = torch.randn(n, device="cuda")
        ms3 = time_fn(lambda: capi.adam_step(p, g, mu, nu, step=3), iters=20, warmup=3)
        print(f"  adam_step n={n} (f32 grad): {ms3:.4f} ms {28.0 * n / ms3 / 1e6:.0f} GB/s (28 B per element)", flush=True)
        gb = g.to(torch.bfloat16)
        ms4 = time_fn(lambda: capi.adam_step(p, gb, mu, nu, step=3), iters=20, warmup=3)
        print(f"  adam_step n={n} (bf16 grad): {ms4:.4f} ms {26.0 * n / ms4 / 1e6:.0f} GB/s (26 B per element)", flush=True)
    elif args.what == "attn_bwd_time":
        # backward of the attention core at the case5 shape, against torch SDPA's (cuDNN / flash) backward
        for (B, S, H, D, causal) in [(8, 4096, 32, 128, False), (8, 4096, 32, 128, True), (8, 4096, 32, 64, False)]:
            q, k, v, do = (torch.randn(B, S, H, D, device="cuda").to(torch.bfloat16) for _ in range(4))
            o, lse = capi.attention_fwd_lse(q, k, v, causal=causal)
            ms = time_fn(lambda: capi.attention_bwd(q, k, v, o, do, lse, causal=causal), iters=5, warmup=2)
            fl = 2.5 * ((2.0 * B * H * S * (S + 1) * D) if causal else 4.0 * B * H * S * S * D)   # 5 matmuls vs the forward's 2
            print(f"  attn bwd B={B} S={S} H={H} D={D} causal={causal}: {ms:.3f} ms {fl / ms / 1e9:.1f} TFLOP/s", flush=True)
            qt, kt, vt = (t.transpose(1, 2).detach().requires_grad_(True) for t in (q, k, v))
            ot = torch.nn.functional.scaled_dot_product_attention(qt, kt, vt, is_causal=causal)
            dot = do.transpose(1, 2)
            def sdpa_bwd():
                torch.autograd.grad(ot, (qt, kt, vt), dot, retain_graph=True)
            ms2 = time_fn(sdpa_bwd, iters=5, warmup=2)
            print(f"    torch sdpa backward: {ms2:.3f} ms {fl / ms2 / 1e9:.1f} TFLOP/s", flush=True)
    elif args.what == "attn_bwd_prof":
        q, k, v, do = (torch.randn(2, 4096, 16, 128, device="cuda").to(torch.bfloat16) for _ in range(4))
        o, lse = capi.attention_fwd_lse(q, k, v)
        for _ in range(3):
            capi.attention_bwd(q, k, v, o, do, lse)
        torch.cuda.synchronize()
    elif args.what == "attn_bwd_phases":
        # needs a build with B200_EXTRA_NVCC_FLAGS=-DB200_ATTN_PROFILE=1
        import ctypes
        lib = capi.load()
        lib.b200_attention_set_profile_buffer.argtypes = [ctypes.c_void_p]
        buf = torch.zeros(16, dtype=torch.int64, device="cuda")
        B, S, H, D = 8, 4096, 32, 128
        q, k, v, do = (torch.randn(B, S, H, D, device="cuda").to(torch.bfloat16) for _ in range(4))
        o, lse = capi.attention_fwd_lse(q, k, v)
        for _ in range(2):
            capi.attention_bwd(q, k, v, o, do, lse)
        lib.b200_attention_set_profile_buffer(ctypes.c_void_p(buf.data_ptr()))
        capi.attention_bwd(q, k, v, o, do, lse)
        torch.cuda.synchronize()
        lib.b200_attention_set_profile_buffer(None)
        t = buf.cpu().tolist()
        n = max(t[15], 1)
        print(f"  attention backward, block (0,0,0), {n} steps, cycles per step:")
        for name, c in zip(["wait S", "wait staging read + barrier", "P pass + arrive", "wait dP", "dS pass + arrive",
                            "wait step_done (dV, dK, dQ MMAs)", "dQ read-out + staging + TMA issue"], t[:7]):
            print(f"    compute warp 4: {name:38s} {c / n:8.0f}")
        print(f"    {'compute total':54s} {sum(t[:7]) / n:8.0f}")
        for name, c in zip(["wait Q", "wait dq_free", "issue calls (S, dV, dK+dQ)", "wait P", "wait dS", "issue dP of next step (incl. wait dO)",
                            "first dP"], t[8:15]):
            print(f"    MMA warp: {name:44s} {c / n:8.0f}")
    elif args.what == "attn_bwd_var":
        # run-to-run spread of the causal backward inside ONE process (the figure differed between boxes): 8 x 5 launches
        B, S, H, D = 8, 4096, 32, 128
        q, k, v, do = (torch.randn(B, S, H, D, device="cuda").to(torch.bfloat16) for _ in range(4))
        for causal in (True, False):
            o, lse = capi.attention_fwd_lse(q, k, v, causal=causal)
            fl = 2.5 * ((2.0 * B * H * S * (S + 1) * D) if causal else 4.0 * B * H * S * S * D)
            vals = [fl / time_fn(lambda: capi.attention_bwd(q, k, v, o, do, lse, causal=causal), iters=5, warmup=1) / 1e9
                    for _ in range(8)]
            print(f"  attn bwd causal={causal}: " + " ".join(f"{x:.0f}" for x in vals) + " TFLOP/s", flush=True)
    elif args.what == "attn_quick":
        # one line per call, for A/B runs of library variants: burst (5 iterations) non-causal / causal and a sustained
        # run (300 launches) with the NVML clock, all at B8 S4096 H32 D128
        sys.path.insert(0, ROOT)
        from bench import ClockSampler
        if args.g:
            capi.set_attention_variant(args.g)
        B, S, H, D = 8, 4096, 32, 128
        q, k, v = (torch.randn(B, S, H, D, device="cuda").to(torch.bfloat16) for _ in range(3))
        o = torch.empty_like(q)
        out = []
        for causal in (False, True):
            ms = time_fn(lambda: capi.attention_fwd(q, k, v, causal=causal, out=o), iters=5, warmup=2)
            fl = (2.0 * B * H * S * (S + 1) * D) if causal else 4.0 * B * H * S * S * D
            out.append(fl / ms / 1e9)
        cs = ClockSampler(0)
        cs.start()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for _ in range(300):
            capi.attention_fwd(q, k, v, out=o)
        e1.record()
        torch.cuda.synchronize()
        c = cs.stop()
        ms = e0.elapsed_time(e1) / 300
        print(f"  QUICK g={args.g} burst {out[0]:.1f} / causal {out[1]:.1f}  sustained {4.0 * B * H * S * S * D / ms / 1e9:.1f} TFLOP/s "
              f"@ {c.get('sm_mhz')} MHz", flush=True)
    elif args.what == "attn_time":
        if args.g:
            capi.set_attention_variant(args.g)
        for (B, S, H, D, causal) in [(8, 4096, 32, 128, False), (8, 4096, 32, 128, True), (4, 8192, 32, 128, True),
                                     (8, 4096, 32, 64, False)]:
            if args.g == 9 and D != 128:
                continue
            q, k, v = (torch.randn(B, S, H, D, device="cuda").to(torch.bfloat16) for _ in range(3))
            o = torch.empty_like(q)
            ms = time_fn(lambda: capi.attention_fwd(q, k, v, causal=causal, out=o), iters=5, warmup=2)
            fl = (2.0 * B * H * S * (S + 1) * D) if causal else 4.0 * B * H * S * S * D
            print(f"  attn B={B} S={S} H={H} D={D} causal={causal}: {ms:.3f} ms {fl / ms / 1e9:.1f} TFLOP/s", flush=True)
            qt, kt, vt = (t.transpose(1, 2) for t in (q, k, v))
            ms = time_fn(lambda: torch.nn.functional.scaled_dot_product_attention(qt, kt, vt, is_causal=causal), iters=5,
                         warmup=2)
            print(f"    torch sdpa: {ms:.3f} ms {fl / ms / 1e9:.1f} TFLOP/s", flush=True)
    else:
        raise SystemExit(f"unknown check {args.what}")
    print("RESULT", "PASS" if ok else "FAIL", flush=True)
    sys.exit(0 if ok else 1)


if __name__ == "__main__":
    main()
