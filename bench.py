#!/usr/bin/env python
"""Headline benchmark: the sharded feed-forward block of BASELINE.json configs[1]
(case4_gspmd_ff: d_model 4096, d_ff 16384, batch 8 x seq 2048, bf16; W1 column-parallel, W2
row-parallel over the 'model' axis), one forward pass per step, synthetic data.

  python bench.py [--gpus N] [--steps K] [--warmup W]            # N>1: launched under torchrun
  python bench.py --impl reference ...                           # CPU restatement on the host cores

Prints ONE JSON line (rank 0).  `value` = whole-job TFLOP/s with inputs resident in HBM, timed
with CUDA events, max over ranks.  `e2e` = the same metric through the public API with the
step's activations in pinned HOST memory (H2D of x and D2H of y inside the timed region).
Other workloads (--workload case5_attention_dense / case6_attention / case3_fully_sharded)
are measurement aids; the driver's default line is case4_gspmd_ff.
"""
import argparse
import json
import os
import subprocess
import sys
import threading
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)

import torch  # noqa: E402

METRIC = "sharded FFN + attention TFLOP/s at 1/2/4/8 B200; % of bf16 tensor peak"
UNIT = "TFLOP/s"

WORKLOADS = {
    # name: (batch, seq, d_model, d_ff)
    "case4_gspmd_ff": dict(batch=8, seq=2048, d_model=4096, d_ff=16384, parallelism="tp"),
    "case3_fully_sharded": dict(batch=8, seq=2048, d_model=4096, d_ff=16384, parallelism="fsdp"),
}
ATTN_WORKLOADS = {
    "case5_attention_dense": dict(batch=8, seq=4096, heads=32, dim_head=128, d_model=4096, causal=False),
    "case6_attention": dict(batch=16, seq=8192, heads=32, dim_head=128, d_model=4096, causal=True),
}


def measured_peaks():
    path = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(path):
        with open(path) as f:
            p = json.load(f)
        return {"bf16_burst": p.get("bf16_tflops"), "bf16_sustained": p.get("bf16_tflops_sustained"),
                "hbm_gbs": p.get("hbm_gbs"), "source": "measured (MEASURED_PEAKS.json)"}
    return {"bf16_burst": 1590.0, "bf16_sustained": 1400.0, "hbm_gbs": 6650.0,
            "source": "fallback (B200_PROFILING.md)"}


class ClockSampler:
    """Samples SM clock, power and throttle reasons through NVML (~every 2 ms) while the timed region runs."""

    def __init__(self, gpu_index: int):
        self.gpu_index = gpu_index
        self.samples = []
        self._stop = threading.Event()
        self._thread = None
        self._nvml = None

    def start(self):
        try:
            import pynvml
            pynvml.nvmlInit()
            self._nvml = pynvml
            self._handle = pynvml.nvmlDeviceGetHandleByIndex(self._physical_index())
        except Exception as e:   # noqa: BLE001 - no NVML: report it in the JSON line
            self._err = repr(e)
            self._nvml = None
            return
        self._thread = threading.Thread(target=self._loop, daemon=True)
        self._thread.start()

    def _physical_index(self):
        vis = os.environ.get("CUDA_VISIBLE_DEVICES")
        if vis:
            ids = [v for v in vis.split(",") if v.strip() != ""]
            try:
                return int(ids[self.gpu_index])
            except (ValueError, IndexError):
                return self.gpu_index
        return self.gpu_index

    def _loop(self):
        n = self._nvml
        while not self._stop.is_set():
            try:
                sm = n.nvmlDeviceGetClockInfo(self._handle, n.NVML_CLOCK_SM)
                pw = n.nvmlDeviceGetPowerUsage(self._handle) / 1000.0
                rs = n.nvmlDeviceGetCurrentClocksEventReasons(self._handle)
                self.samples.append((sm, pw, rs))
            except Exception:   # noqa: BLE001
                pass
            time.sleep(0.002)

    def stop(self):
        if self._nvml is None:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvml unavailable: " + getattr(self, "_err", "")]}
        self._stop.set()
        self._thread.join(timeout=2)
        n = self._nvml
        try:
            smax = n.nvmlDeviceGetMaxClockInfo(self._handle, n.NVML_CLOCK_SM)
        except Exception:   # noqa: BLE001
            smax = None
        names = {"hw_slowdown": getattr(n, "nvmlClocksEventReasonHwSlowdown", 0x8),
                 "hw_thermal_slowdown": getattr(n, "nvmlClocksEventReasonHwThermalSlowdown", 0x40),
                 "sw_thermal_slowdown": getattr(n, "nvmlClocksEventReasonSwThermalSlowdown", 0x20),
                 "sw_power_cap": getattr(n, "nvmlClocksEventReasonSwPowerCap", 0x4)}
        sm = sorted(s[0] for s in self.samples)
        reasons = sorted(k for k, bit in names.items() if any(s[2] & bit for s in self.samples))
        return {"sm_mhz": sm[len(sm) // 2] if sm else None, "sm_min_mhz": sm[0] if sm else None, "sm_max_mhz": smax,
                "power_w_max": max((s[1] for s in self.samples), default=None), "samples": len(sm), "reasons": reasons}


def train_kernel_sample(dev, peaks):
    """CUDA-event timings of the backward-side kernels at the case5 shape (B8, S4096, 32 heads x 128, d_model 4096):
    attention backward, the weight-gradient GEMM dW = x^T . dy, and the Adam update of one 4096 x 4096 kernel."""
    import torch
    from learning_jax_sharding_b200 import capi

    def timed(fn, iters=5, warmup=2):
        for _ in range(warmup):
            fn()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for _ in range(iters):
            fn()
        e1.record()
        torch.cuda.synchronize(dev)
        return e0.elapsed_time(e1) / iters

    B, S, H, D, M = 8, 4096, 32, 128, 4096
    q, k, v, do = (torch.randn(B, S, H, D, device=dev).to(torch.bfloat16) for _ in range(4))
    o, lse = capi.attention_fwd_lse(q, k, v)
    ms_bwd = timed(lambda: capi.attention_bwd(q, k, v, o, do, lse))
    bwd_flops = 2.5 * 4.0 * B * H * S * S * D
    del q, k, v, do, o, lse
    x = torch.randn(B * S, M, device=dev).to(torch.bfloat16)
    dy = torch.randn(B * S, M, device=dev).to(torch.bfloat16)
    dw = torch.empty(M, M, device=dev, dtype=torch.float32)
    ms_dw = timed(lambda: capi.gemm_bf16_tn(x, dy, out=dw))
    n = M * M
    p, mu, nu = (torch.randn(n, device=dev) for _ in range(3))
    nu.abs_()
    g = dw.reshape(-1)
    ms_adam = timed(lambda: capi.adam_step(p, g, mu, nu, step=3), iters=20)
    return {"workload": "case5 shape: attention backward (dQ, dK, dV; 2.5 x the forward's FLOPs), dW = x^T . dy "
                        "(32768 x 4096 x 4096), Adam on 4096 x 4096 f32",
            "attention_bwd_tflops": bwd_flops / (ms_bwd * 1e-3) / 1e12, "attention_bwd_ms": ms_bwd,
            "dw_gemm_tflops": 2.0 * B * S * M * M / (ms_dw * 1e-3) / 1e12, "dw_gemm_ms": ms_dw,
            "dw_gemm_frac_of_peak": 2.0 * B * S * M * M / (ms_dw * 1e-3) / 1e12 / peaks["bf16_burst"],
            "adam_gbs": 28.0 * n / (ms_adam * 1e-3) / 1e9, "adam_ms": ms_adam,
            "adam_frac_of_hbm_peak": 28.0 * n / (ms_adam * 1e-3) / 1e9 / peaks.get("hbm_gbs", 6551.4)}


def workload_config(workload, cfg, act, world):
    """The `config` object of the JSON line: what is computed, identical in the product arm and in `--impl reference` (how each
    arm computes it is described elsewhere in its line: `implementation`, `cpu_baseline.sample`)."""
    mode = cfg.get("parallelism", "tp")
    return {"workload": workload, "batch": cfg["batch"], "seq_len": cfg["seq"], "d_model": cfg["d_model"], "d_ff": cfg["d_ff"],
            "activation": act, "bias": True, "flops_per_step": ffn_flops(cfg),
            "parallelism": "1 GPU" if world == 1 else (f"{mode}{world}" + (" (W1 columns / W2 rows over 'model')" if mode == "tp"
                                                                       else " (tokens and weight rows over 'data')")),
            "l2": "inputs larger than L2 (x 134 MB, W1 134 MB, W2 134 MB, h 537 MB per step vs 126 MB L2)"}


def ffn_flops(cfg):
    return 4.0 * cfg["batch"] * cfg["seq"] * cfg["d_model"] * cfg["d_ff"]


def attn_flops(cfg):
    b, s, h, d, m = cfg["batch"], cfg["seq"], cfg["heads"], cfg["dim_head"], cfg["d_model"]
    core = 2.0 * b * h * s * (s + 1) * d if cfg["causal"] else 4.0 * b * h * s * s * d
    proj = 8.0 * b * s * m * (h * d)
    return core, proj


# ------------------------------------------------------------------------------ CPU legs
def cpu_ffn_sample(cfg, seconds=12.0, tokens=2048, act="gelu_tanh"):
    """The oracle (CPU restatement) timed on the host cores on a bounded sample of the workload:
    `tokens` rows of the same FFN, repeated until ~`seconds` of CPU work."""
    from oracle import reference_math as rm
    d, f = cfg["d_model"], cfg["d_ff"]
    torch.set_num_threads(os.cpu_count() or 1)
    x = rm.synthetic((tokens, d), 0).to(torch.bfloat16)
    w1, b1 = rm.synthetic((d, f), 1, "weight").to(torch.bfloat16), rm.synthetic((f,), 2, "bias").to(torch.bfloat16)
    w2, b2 = rm.synthetic((f, d), 3, "weight").to(torch.bfloat16), rm.synthetic((d,), 4, "bias").to(torch.bfloat16)
    rm.ffn_forward(x[:256], w1, b1, w2, b2, act)       # warm-up
    reps, t0 = 0, time.perf_counter()
    while True:
        rm.ffn_forward(x, w1, b1, w2, b2, act)
        reps += 1
        el = time.perf_counter() - t0
        if el >= seconds or reps >= 50:
            break
    flops = 4.0 * tokens * d * f * reps
    return {"value": flops / el / 1e12, "unit": UNIT, "cores": torch.get_num_threads(), "kind": "port",
            "sample": f"{tokens} of {cfg['batch'] * cfg['seq']} tokens of the same FFN (oracle ffn_forward, torch-CPU "
                      f"bf16-rounded/f32-accumulate), {reps} passes in {el:.1f} s; JAX is not installable here",
            "seconds": el, "ms_per_pass": el / reps * 1e3}


def cpu_attention_sample(cfg, seq=1024, seconds=10.0):
    """The oracle's MHA forward (projections + attention core + out-projection) on the host cores on a bounded sample of the
    workload: ONE batch item cut to `seq` tokens, same widths and heads (attention cost is quadratic in the sequence, so the
    sample's own FLOPs are what the rate is computed from)."""
    from oracle import reference_math as rm
    torch.set_num_threads(os.cpu_count() or 1)
    M, H, D = cfg["d_model"], cfg["heads"], cfg["dim_head"]
    x = rm.synthetic((1, seq, M), 0)
    params = {"to_q": rm.synthetic((M, H * D), 1, "weight"), "to_k": rm.synthetic((M, H * D), 2, "weight"),
              "to_v": rm.synthetic((M, H * D), 3, "weight"),
              "to_out_0": (rm.synthetic((H * D, M), 4, "weight"), rm.synthetic((M,), 5, "bias"))}
    rm.mha_forward(x[:, :128], params, H, D, cfg["causal"])
    reps, t0 = 0, time.perf_counter()
    while True:
        rm.mha_forward(x, params, H, D, cfg["causal"])
        reps += 1
        el = time.perf_counter() - t0
        if el >= seconds or reps >= 50:
            break
    core = 2.0 * H * seq * (seq + 1) * D if cfg["causal"] else 4.0 * H * seq * seq * D
    flops = (core + 8.0 * seq * M * H * D) * reps
    return {"value": flops / el / 1e12, "unit": UNIT, "cores": torch.get_num_threads(), "kind": "port",
            "sample": f"oracle mha_forward on 1 batch item x {seq} of {cfg['seq']} tokens (same widths, {H} heads x {D}), {reps} "
                      f"passes in {el:.1f} s; rate from the sample's own FLOPs; JAX is not installable here"}


def run_reference(args):
    """--impl reference: the reference's own CPU path is jax/XLA:CPU, which cannot be installed in this
    image; the timed stand-in is the oracle port on all host cores, each step a bounded sample."""
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    cfg = WORKLOADS["case4_gspmd_ff"]
    tokens = 1024
    from oracle import reference_math as rm
    d, f = cfg["d_model"], cfg["d_ff"]
    torch.set_num_threads(os.cpu_count() or 1)
    x = rm.synthetic((tokens, d), 0).to(torch.bfloat16)
    w1, b1 = rm.synthetic((d, f), 1, "weight").to(torch.bfloat16), rm.synthetic((f,), 2, "bias").to(torch.bfloat16)
    w2, b2 = rm.synthetic((f, d), 3, "weight").to(torch.bfloat16), rm.synthetic((d,), 4, "bias").to(torch.bfloat16)
    for _ in range(max(1, min(args.warmup, 2))):
        rm.ffn_forward(x, w1, b1, w2, b2, args.act)
    steps = min(args.steps, 200)            # ~0.2 s per step on 16+ cores: the default 50 steps take ~10 s
    t0 = time.perf_counter()
    for _ in range(steps):
        rm.ffn_forward(x, w1, b1, w2, b2, args.act)
    el = time.perf_counter() - t0
    value = 4.0 * tokens * d * f * steps / el / 1e12
    sample = (f"each step = {tokens} of {cfg['batch'] * cfg['seq']} tokens of the case4_gspmd_ff FFN through the oracle "
              f"port (torch-CPU); the reference's jax/XLA:CPU path is not installable in this image")
    line = {"impl": "reference", "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": args.gpus, "steps": steps,
            "warmup": args.warmup, "ms_per_step": el / steps * 1e3, "higher_is_better": True, "scaling": "strong",
            "vs_baseline": None, "dtype": "bf16", "data": "synthetic",
            "config": workload_config("case4_gspmd_ff", cfg, args.act, args.gpus),
            "cpu_baseline": {"value": value, "unit": UNIT, "cores": torch.get_num_threads(), "kind": "port",
                             "sample": sample},
            "e2e": {"value": value, "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0}}
    print(json.dumps(line), flush=True)


# ------------------------------------------------------------------------------ parity checks (oracle = checker only)
def _scaled_need(got, want, rtol=2e-2):
    """The tolerance tests/_tol.py uses, restated here so that bench.py needs nothing from tests/: returns
    (c needed by |got - want| <= rtol |want| + c 2^-8 rms_row, max error in bf16 ulps of max(|want|, rms_row), max abs)."""
    got, want = got.double(), want.double()
    rms = want.pow(2).mean(dim=-1, keepdim=True).sqrt()
    rms = torch.maximum(rms, 0.05 * want.pow(2).mean().sqrt().expand_as(rms)).clamp_min(1e-30)
    err = (got - want).abs()
    need = ((err - rtol * want.abs()).clamp_min(0.0) / (2.0 ** -8 * rms)).max()
    ulps = (err / (2.0 ** -7 * torch.maximum(want.abs(), rms))).max()
    return float(need), float(ulps), float(err.max())


def _reduce_check(dist, dev, rows, need, ulps, max_abs, what):
    if dist is not None:
        t = torch.tensor([need, ulps, max_abs], device=dev, dtype=torch.float64)
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        r = torch.tensor([rows], device=dev, dtype=torch.float64)
        dist.all_reduce(r)
        need, ulps, max_abs, rows = float(t[0]), float(t[1]), float(t[2]), int(r.item())
    return {"rows": rows, "max_abs": max_abs, "max_ulp": ulps, "c_needed": need, "ok": bool(need <= 4.0),
            "bound": "|got - oracle| <= 2e-2 |oracle| + 4 * 2^-8 * rms_row(oracle)", "what": what}


def ffn_parity_check(bench, act, dist, dev, rows_per_rank=16):
    """After the timed region: this rank's output block of one more resident step against the CPU oracle on a seeded
    sample of its rows (every rank checks its own rows; max over ranks).  The oracle is the checker, never timed."""
    from oracle import reference_math as rm
    torch.set_num_threads(max(1, (os.cpu_count() or 1) // max(bench.world, 1)))
    y = bench.step_resident()
    torch.cuda.synchronize(dev)
    g = torch.Generator().manual_seed(1000 + bench.rank)
    local = torch.randint(0, y.shape[0], (rows_per_rank,), generator=g)
    local[0], local[-1] = 0, y.shape[0] - 1
    x_rows = bench.x_host[local] if bench.mode == "fsdp" else bench.x_host[bench.rank * y.shape[0] + local]
    w1, b1, w2, b2 = bench.host_weights
    want = rm.ffn_forward(x_rows.float(), w1.float(), b1.float(), w2.float(), b2.float(), act)
    need, ulps, max_abs = _scaled_need(y[local.to(dev)].cpu().float(), want.float())
    return _reduce_check(dist, dev, rows_per_rank, need, ulps, max_abs,
                         f"{rows_per_rank} seeded rows of every rank's y block, oracle ffn_forward on the host")


def attention_parity_check(r, cfg, dist, dev, rank, world):
    """Stage-wise parity of the MHA layer at the config shape, each stage's GPU output against the oracle applied to that
    stage's GPU inputs: (1) fused q/k/v projection on sampled token rows, (2) the attention core on one (batch, local head)
    pair per rank over the full sequence, (3) the row-parallel out-projection + cross-rank reduction on sampled rows."""
    from oracle import reference_math as rm
    torch.set_num_threads(max(1, (os.cpu_count() or 1) // max(world, 1)))
    layer, x = r["layer"], r["x"]
    B, S, Hl, D, M = layer.B, layer.S, layer.Hl, layer.D, layer.M
    inner = Hl * D
    y = layer.forward(x)
    torch.cuda.synchronize(dev)
    g = torch.Generator().manual_seed(2000 + rank)
    worst = [0.0, 0.0, 0.0]

    def upd(got, want):
        n, u, m = _scaled_need(got.float(), want.float())
        worst[0], worst[1], worst[2] = max(worst[0], n), max(worst[1], u), max(worst[2], m)

    # (1) projections: 8 token rows of x against x_row @ Wqkv
    rows = torch.randint(0, B * S, (8,), generator=g)
    xr = x.reshape(B * S, M)[rows.to(dev)].cpu()
    upd(layer.qkv[rows.to(dev)].cpu(), rm.dense_bf16(xr.float(), layer.wqkv.cpu().float()))
    # (2) core: one (batch, head) of the local shard
    b, h = int(torch.randint(0, B, (1,), generator=g)), int(torch.randint(0, Hl, (1,), generator=g))
    qkv = layer.qkv.view(B, S, 3, Hl, D)
    q, k, v = (qkv[b:b + 1, :, i, h:h + 1].cpu() for i in range(3))
    want = rm.attention_core_blocked(q, k, v, causal=cfg["causal"], q_block=1024)
    upd(layer.o.view(B, S, Hl, D)[b:b + 1, :, h:h + 1].cpu(), want)
    # (3) out-projection: rows of this rank's y block need the o rows of every rank of its model group
    d_ax, m_ax = r["mesh"]
    yrows = torch.randint(0, y.shape[0], (8,), generator=torch.Generator().manual_seed(3000 + rank // m_ax))
    m_rank = rank % m_ax
    blk = (B * S) // m_ax
    # every rank of the model group evaluates the partial of EVERY member's sampled rows (8 x m_ax rows): tiny
    parts = []
    for owner in range(m_ax):
        grows = layer.rs.global_rows(yrows, owner) if layer.rs is not None else yrows
        o_rows = layer.o[grows.to(dev)].cpu().float()
        parts.append(o_rows @ layer.wo.cpu().float())
    mine = torch.stack(parts)                                            # [m_ax, 8, M] f32: this rank's partial products
    if m_ax > 1:
        tot = mine.to(dev)
        dist.all_reduce(tot, group=r["group"])
        mine = tot.cpu()
    want_y = ((mine[m_rank].to(torch.bfloat16).float() + layer.bo.cpu().float())).to(torch.bfloat16)
    upd(y[yrows.to(dev)].cpu(), want_y)
    return _reduce_check(dist, dev, 24 + S, worst[0], worst[1], worst[2],
                         "stage-wise: 8 projection rows, one (batch, head) attention core over the full sequence, 8 "
                         "out-projection rows incl. the cross-rank sum, per rank")


# ------------------------------------------------------------------------------ extra legs of the default line
def host_link_leg(bench, dist, dev, reps=10):
    """Copy-only roofline of the end-to-end step: every rank concurrently uploads its share of x from pinned memory and
    downloads its share of y to pinned memory (both directions at once, no compute)."""
    main = torch.cuda.current_stream(dev)
    if bench.world > 1 and bench.mode == "tp":
        rows = bench.T // bench.world
        sl = slice(bench.rank * rows, (bench.rank + 1) * rows)
    else:
        sl = slice(0, bench.T_in)
    xs, ys = bench._x_stages[0], bench._y_devs[0]

    def one():
        with torch.cuda.stream(bench.copy_in):
            xs[sl].copy_(bench.x_host[sl] if (bench.world > 1 and bench.mode == "tp") else bench.x_host, non_blocking=True)
        with torch.cuda.stream(bench.copy_out):
            bench._y_hosts[0].copy_(ys, non_blocking=True)

    for _ in range(2):
        one()
    torch.cuda.synchronize(dev)
    if dist is not None:
        dist.barrier()
    bench.copy_in.wait_stream(main)
    bench.copy_out.wait_stream(main)
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record(main)
    bench.copy_in.wait_event(e0)
    bench.copy_out.wait_event(e0)
    for _ in range(reps):
        one()
    main.wait_stream(bench.copy_in)
    main.wait_stream(bench.copy_out)
    e1.record(main)
    torch.cuda.synchronize(dev)
    ms = e0.elapsed_time(e1) / reps
    if dist is not None:
        t = torch.tensor([ms], device=dev)
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        ms = float(t.item())
    return ms


def sustained_leg(bench, dist, dev, local_rank, total_flops, peaks, seconds=2.5):
    """The same resident step looped for >= `seconds` with NVML sampled inside the loop: the power-capped steady state."""
    ms_est = None
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(10):
        bench.step_resident()
    e1.record()
    torch.cuda.synchronize(dev)
    ms_est = max(e0.elapsed_time(e1) / 10, 1e-3)
    steps = int(min(max(seconds * 1e3 / ms_est, 50), 20000))
    if dist is not None:
        t = torch.tensor([steps], device=dev)
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        steps = int(t.item())
        dist.barrier()
    torch.cuda.synchronize(dev)
    sampler = ClockSampler(local_rank)
    sampler.start()
    e0.record()
    for i in range(steps):
        bench.step_resident()
        if i % 64 == 63:
            e1.record()
            e1.synchronize()          # keep the launch queue bounded (the host stays well ahead of the GPU)
    e1.record()
    torch.cuda.synchronize(dev)
    clocks = sampler.stop()
    ms = e0.elapsed_time(e1) / steps
    if dist is not None:
        t = torch.tensor([ms], device=dev)
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        ms = float(t.item())
    val = total_flops / (ms * 1e-3) / 1e12
    return {"value": val, "unit": UNIT, "ms_per_step": ms, "steps": steps, "seconds": ms * steps * 1e-3,
            "per_gpu_frac_of_sustained_peak": val / bench.world / peaks["bf16_sustained"],
            "per_gpu_frac_of_burst_peak": val / bench.world / peaks["bf16_burst"], "clocks": clocks}


def api_leg(ljs, host_tensors, cfg, act, dist, dev, rank, world, total_flops, layers_value, steps=20):
    """The same FFN step through the reference-shaped public API: nn.jit(apply_fn, in_shardings, out_shardings) around
    nn.ffn on NamedSharding operands of a (1, N) data x model mesh -- W1 P(None,'model'), b1 P('model'), W2 P('model',None),
    x replicated over 'model', y left sequence-sharded P('data','model') like the scripts' out_shardings
    (case6_attention.py:229-232).  On N > 1 GPUs nn.ffn dispatches to the fused tensor-parallel layer."""
    from oracle import reference_math as rm
    if world == 1:
        ljs.init_virtual(1, str(dev))
    P = ljs.PartitionSpec
    mesh = ljs.Mesh(ljs.create_device_mesh((1, world)), ("data", "model"))
    ns = lambda spec: ljs.NamedSharding(mesh, spec)
    m = "model" if world > 1 else None
    x, w1, b1, w2, b2 = host_tensors
    B, S, d = cfg["batch"], cfg["seq"], cfg["d_model"]
    params = (ljs.device_put(w1, ns(P(None, m))), ljs.device_put(b1, ns(P(m))), ljs.device_put(w2, ns(P(m, None))),
              ljs.device_put(b2, ns(P())))
    xin = ljs.device_put(x.reshape(B, S, d), ns(P("data")))
    out_sh = ns(P("data", m))
    apply_fn = ljs.nn.jit(lambda prm, xx: ljs.nn.ffn(xx, *prm, act=act, scatter_dim=1 if world > 1 else None),
                          in_shardings=(None, ns(P("data"))), out_shardings=out_sh)
    for _ in range(3):
        y = apply_fn(params, xin)
    torch.cuda.synchronize(dev)
    if dist is not None:
        dist.barrier()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(steps):
        y = apply_fn(params, xin)
    e1.record()
    torch.cuda.synchronize(dev)
    ms = e0.elapsed_time(e1) / steps
    if dist is not None:
        t = torch.tensor([ms], device=dev)
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        ms = float(t.item())
    # parity of what the public call returned: 8 rows of this rank's shard against the oracle
    sh = y.addressable_shards[0]
    yl = sh.data.reshape(-1, d)
    rows_b = torch.randint(0, sh.data.shape[0], (8,), generator=torch.Generator().manual_seed(4000 + rank))
    rows_s = torch.randint(0, sh.data.shape[1], (8,), generator=torch.Generator().manual_seed(5000 + rank))
    grow = (sh.index[0].start + rows_b) * S + sh.index[1].start + rows_s
    torch.set_num_threads(max(1, (os.cpu_count() or 1) // max(world, 1)))
    want = rm.ffn_forward(x[grow].float(), w1.float(), b1.float(), w2.float(), b2.float(), act)
    need, ulps, max_abs = _scaled_need(yl[(rows_b * sh.data.shape[1] + rows_s).to(dev)].cpu().float(), want.float())
    val = total_flops / (ms * 1e-3) / 1e12
    fused = any(k[0] == "ffn" and v is not False for k, v in ljs.nn._fast_layers.items())
    return {"api": "nn.jit(apply_fn, in_shardings, out_shardings)(params, x) with apply_fn = nn.ffn on NamedSharding operands",
            "value": val, "unit": UNIT, "ms_per_step": ms, "steps": steps, "ratio_to_layers_value": val / layers_value,
            "served_by": "layers.FusedTensorParallelFFN (fused reduce-scatter along seq)" if fused else "generic nn.dense plan",
            "y_shard_shape": list(sh.data.shape),
            "parity_check": _reduce_check(dist, dev, 8, need, ulps, max_abs, "8 seeded rows of every rank's y shard")}


def case1a_leg(ljs, dist, dev, rank, world):
    """BASELINE.json configs[0]: the sharded fp32 matmul of case1a.py on REAL devices.  At every N: the contraction
    dim split over a 1-D mesh of the N GPUs (x P(None,'m'), W P('m',None) -> partial products + all-reduce), literal
    size (4,16)@(16,4), microseconds per lax.dot and parity at rtol 1e-5; at N = 8 also the five literal 2x4
    PositionalSharding layouts of case1a/1b/2/3/4; on one GPU those five run on 8 virtual devices.  Plus one large
    true-fp32 product through the same kernel (b200_gemm_f32) in TFLOP/s."""
    from oracle import reference_math as rm
    g = torch.Generator().manual_seed(0)
    A, B = torch.randn(4, 16, generator=g), torch.randn(16, 4, generator=g)
    want = rm.dot_f32(A, B)
    out = {}

    def timed(fn, iters=50):
        for _ in range(3):
            fn()
        torch.cuda.synchronize(dev)
        if dist is not None:
            dist.barrier()
        t0 = time.perf_counter()
        for _ in range(iters):
            c = fn()
        torch.cuda.synchronize(dev)
        us = (time.perf_counter() - t0) / iters * 1e6
        if dist is not None:
            t = torch.tensor([us], device=dev)
            dist.all_reduce(t, op=dist.ReduceOp.MAX)
            us = float(t.item())
        return us, c

    def check(c):
        ok = True
        for sh in c.addressable_shards:
            ok &= bool(torch.allclose(sh.data.cpu(), want[sh.index], rtol=1e-5, atol=1e-6))
        if dist is not None:
            t = torch.tensor([1 if ok else 0], device=dev)
            dist.all_reduce(t, op=dist.ReduceOp.MIN)
            ok = bool(t.item())
        return ok

    if world > 1:
        mesh = ljs.Mesh(ljs.create_device_mesh((world,)), ("m",))
        a = ljs.device_put(A, ljs.NamedSharding(mesh, ljs.PartitionSpec(None, "m")))
        b = ljs.device_put(B, ljs.NamedSharding(mesh, ljs.PartitionSpec("m", None)))
        us, c = timed(lambda: ljs.lax.dot(a, b))
        out["mesh_1d"] = {"devices": world, "runtime": "one process per GPU (NCCL)", "us_per_dot": us, "parity_ok": check(c),
                          "layout": "x P(None,'m'), W P('m',None): local (4,16/N)@(16/N,4) + all-reduce"}
    lay_world = 8 if world in (1, 8) else 0
    if lay_world:
        if world == 1:
            ljs.init_virtual(8, str(dev))
        s = ljs.PositionalSharding(ljs.create_device_mesh((2, 4)))
        layouts = {"case1a": (s.replicate(axis=0, keepdims=True), s.reshape(4, 2).replicate(axis=1, keepdims=True)),
                   "case1b": (s.replicate(axis=0, keepdims=True), s.replicate(axis=1, keepdims=True)),
                   "case2": (s, s.replicate(axis=1, keepdims=True)), "case3": (s, s),
                   "case4": (s.replicate(axis=1, keepdims=True), s.replicate(axis=0, keepdims=True))}
        res = {}
        for name, (sa, sb) in layouts.items():
            a, b = ljs.device_put(A, sa), ljs.device_put(B, sb)
            us, c = timed(lambda: ljs.lax.dot(a, b), iters=20)
            res[name] = {"us_per_dot": us, "parity_ok": check(c)}
        out["literal_2x4"] = {"runtime": "8 virtual devices on one GPU" if world == 1 else "8 GPUs, one process each (NCCL)",
                              "layouts": res}
    # large true-fp32 product on this GPU
    n = 4096
    a32, b32 = torch.randn(n, n, device=dev), torch.randn(n, n, device=dev)
    c32 = torch.empty(n, n, device=dev)
    for _ in range(2):
        ljs.capi.gemm_f32(a32, b32, out=c32)
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(5):
        ljs.capi.gemm_f32(a32, b32, out=c32)
    e1.record()
    torch.cuda.synchronize(dev)
    ms = e0.elapsed_time(e1) / 5
    idx = torch.tensor([0, 1, 2047, 4095], device=dev)
    ref = (a32[idx].double() @ b32.double()).cpu()
    out["large_fp32"] = {"shape": [n, n, n], "ms": ms, "tflops": 2.0 * n ** 3 / (ms * 1e-3) / 1e12,
                         "kernel": "gemm_f32_kernel (FFMA, no TF32)",
                         "parity_ok": bool(torch.allclose(c32[idx].cpu().double(), ref, rtol=1e-5, atol=1e-5 * n ** 0.5))}
    out["all_parity_ok"] = all(v.get("parity_ok", True) for v in out.values() if isinstance(v, dict)) and \
        all(l["parity_ok"] for l in out.get("literal_2x4", {}).get("layouts", {}).values())
    return out


def bind_to_gpu_numa_node(local_rank: int):
    """Pin this rank's host threads to the CPUs NVML reports as local to its GPU, BEFORE any pinned host buffer is
    allocated: the staging buffers of the host-buffer step are then first-touched on the GPU's own NUMA node and its
    DMA does not cross the socket interconnect.  Returns the CPU count bound to, or None."""
    try:
        import pynvml
        pynvml.nvmlInit()
        vis = os.environ.get("CUDA_VISIBLE_DEVICES")
        idx = local_rank
        if vis:
            ids = [v for v in vis.split(",") if v.strip() != ""]
            try:
                idx = int(ids[local_rank])
            except (ValueError, IndexError):
                idx = local_rank
        h = pynvml.nvmlDeviceGetHandleByIndex(idx)
        ncpu = os.cpu_count() or 1
        words = (ncpu + 63) // 64
        mask = pynvml.nvmlDeviceGetCpuAffinity(h, words)
        cpus = {w * 64 + b for w, m in enumerate(mask) for b in range(64) if (m >> b) & 1}
        cpus &= os.sched_getaffinity(0)
        if cpus:
            os.sched_setaffinity(0, cpus)
            return len(cpus)
    except Exception:   # noqa: BLE001 - no NVML / no permission: run unbound
        return None
    return None


# ------------------------------------------------------------------------------ GPU legs
def time_resident(bench, dist, dev, steps, warmup, capi):
    def sync_all():
        torch.cuda.synchronize(dev)
        if dist is not None:
            dist.barrier()
            torch.cuda.synchronize(dev)

    for _ in range(warmup):
        bench.step_resident()
    sync_all()
    launches0 = capi.launch_count()
    ev = [torch.cuda.Event(enable_timing=True) for _ in range(2)]
    bench.reset_kernel_events()
    sync_all()
    ev[0].record()
    for _ in range(steps):
        bench.step_resident(record_kernels=True)
    ev[1].record()
    sync_all()
    launches = capi.launch_count() - launches0
    ms = ev[0].elapsed_time(ev[1]) / steps
    if dist is not None:
        t = torch.tensor([ms], device=dev)
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        ms = float(t.item())
    return ms, int(launches)


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=50)
    ap.add_argument("--warmup", type=int, default=5)
    ap.add_argument("--impl", default="b200", choices=["b200", "reference"])
    ap.add_argument("--workload", default="case4_gspmd_ff")
    ap.add_argument("--act", default="gelu_tanh", choices=["gelu_tanh", "relu"])
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--no-attention", action="store_true", help="skip the supplementary attention measurements")
    ap.add_argument("--lean", action="store_true", help="main line only: no fsdp / attention / case1a / sustained legs")
    ap.add_argument("--chunks", type=int, default=4, help="token chunks for collective / copy overlap")
    args = ap.parse_args()
    args.warmup = max(args.warmup, 3)
    if args.impl == "reference":
        return run_reference(args)

    import learning_jax_sharding_b200 as ljs
    from learning_jax_sharding_b200 import layers

    world = int(os.environ.get("WORLD_SIZE", "1"))
    rank = int(os.environ.get("RANK", "0"))
    local_rank = int(os.environ.get("LOCAL_RANK", "0"))
    if world != args.gpus:
        if world == 1 and args.gpus > 1:
            raise SystemExit("bench.py --gpus N>1 must be launched under torch.distributed.run (one rank per GPU)")
    if not torch.cuda.is_available():
        raise SystemExit("bench.py needs a CUDA device: the sharded layers have no CPU fallback")
    numa = bind_to_gpu_numa_node(local_rank) if os.environ.get("B200_BENCH_NUMA", "1") != "0" else None
    torch.cuda.set_device(local_rank)
    dist = None
    if world > 1:
        ljs.init_distributed("nccl")
        import torch.distributed as dist
    dev = torch.device("cuda", local_rank)
    peaks = measured_peaks()

    if args.workload in ATTN_WORKLOADS:
        return layers.bench_attention(args, ATTN_WORKLOADS[args.workload], rank, world, dev, METRIC, UNIT,
                                      peaks, ClockSampler(local_rank), attn_flops, cpu_sample=cpu_attention_sample)
    cfg = WORKLOADS[args.workload]
    bench = layers.FFNBench(cfg, args.act, rank, world, dev, chunks=args.chunks)
    total_flops = ffn_flops(cfg)
    capi = ljs.capi

    def sync_all():
        torch.cuda.synchronize(dev)
        if dist is not None:
            dist.barrier()
            torch.cuda.synchronize(dev)

    # ---------------- kernel-resident timing (`value`) ----------------
    sampler = ClockSampler(local_rank)
    sampler.start()
    ms, launches = time_resident(bench, dist, dev, args.steps, args.warmup, capi)
    clocks = sampler.stop()
    value = total_flops / (ms * 1e-3) / 1e12
    kernel_ms, n_gemm = bench.mean_kernel_ms()    # mean duration of one GEMM launch inside the timed region
    parts_ms = bench.kernel_parts_ms(args.steps)
    gemm_per_step = max(n_gemm // max(args.steps, 1), 1)

    # ---------------- end-to-end through the public API with host buffers (`e2e`) ----------------
    for _ in range(3):
        bench.step_e2e()
    bench.e2e_drain()
    sync_all()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e2e_steps = max(3, min(args.steps, 30))   # a streaming call: enough steps for the ramp of the first one to amortise
    e0.record()
    for _ in range(e2e_steps):
        bench.step_e2e()
    bench.e2e_drain()                              # the last downloads are part of the timed region
    e1.record()
    sync_all()
    e2e_ms = e0.elapsed_time(e1) / e2e_steps
    if dist is not None:
        t = torch.tensor([e2e_ms], device=dev)
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        e2e_ms = float(t.item())
    h2d, d2h = bench.e2e_bytes()
    if dist is not None:
        t = torch.tensor([h2d, d2h], device=dev, dtype=torch.float64)
        dist.all_reduce(t)
        h2d, d2h = int(t[0].item()), int(t[1].item())
    copy_ms = host_link_leg(bench, dist, dev)
    sync_all()

    # ---------------- parity of what was just timed (every rank, max over ranks) ----------------
    parity = ffn_parity_check(bench, args.act, dist, dev)
    par_name = bench.parallelism_name()
    bench_depth = bench.e2e_depth

    flops_per_launch = total_flops / world / gemm_per_step   # GEMM launches of one step on one GPU share its FLOPs equally
    achieved = flops_per_launch / (kernel_ms * 1e-3) / 1e12
    e2e_bytes_gbs = (h2d + d2h) / (e2e_ms * 1e-3) / 1e9
    link_gbs = (h2d + d2h) / (copy_ms * 1e-3) / 1e9
    traffic = layers.ncu_traffic_bytes() if world == 1 else None
    line = {
        "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": world, "steps": args.steps, "warmup": args.warmup,
        "ms_per_step": ms, "higher_is_better": True, "scaling": "strong", "vs_baseline": None, "dtype": "bf16",
        "data": "synthetic",
        "config": workload_config(args.workload, cfg, args.act, world),
        "implementation": {"parallelism": par_name,
                           "api": "layers.FFNBench.step_resident -> b200_gemm_bf16_ex (C ABI) per shard; the same step through "
                                  "nn.jit(apply_fn) is the `api` object"},
        "clocks": clocks,
        "e2e": {"value": total_flops / (e2e_ms * 1e-3) / 1e12, "unit": UNIT, "ms_per_step": e2e_ms,
                "h2d_bytes_per_step": h2d, "d2h_bytes_per_step": d2h,
                "what": "x copied from pinned host memory and y copied back every step, weights resident",
                "host_link_gbs": link_gbs, "host_link_ms_per_step": copy_ms,
                "host_link_what": "copy-only leg: all ranks upload their x share and download their y share concurrently, "
                                  "no compute; aggregate GB/s over both directions",
                "achieved_gbs": e2e_bytes_gbs, "frac_of_host_link": e2e_bytes_gbs / link_gbs,
                "staging_depth": bench_depth, "host_threads_bound_to_gpu_numa_cpus": numa,
                "compute_ms_per_step": ms},
        "gpu_launches": int(launches),
        "parity_check": parity,
        "roofline": {"bound": "tensor", "kernel": "gemm_bf16_kernel (tcgen05, fused bias+activation epilogue)",
                     "achieved": achieved, "peak": peaks["bf16_burst"], "unit": UNIT,
                     "frac": achieved / peaks["bf16_burst"], "peak_sustained": peaks["bf16_sustained"],
                     "frac_of_sustained": achieved / peaks["bf16_sustained"], "frac_of_nominal_2250": achieved / 2250.0,
                     "peak_source": peaks["source"], "launch_ms": kernel_ms, "flops_per_launch": flops_per_launch,
                     "gemm_launches_per_step": gemm_per_step, "parts_ms": parts_ms,
                     "traffic": traffic,
                     "traffic_source": ("profiles/ncu_traffic.json: one-GPU ncu --set full capture of this shape" if world == 1
                                        else "not captured at N > 1 (ncu is single-GPU here); algorithmic bytes per launch below"),
                     "algorithmic_bytes_per_launch": layers.ffn_algorithmic_bytes(cfg, world)},
    }
    # Everything below adds sub-objects to the line.  A watchdog bounds their total time: if a supplementary leg stalls
    # (a rank lost inside a collective), every rank leaves and rank 0 prints the line with what it has.
    extras = line
    done = threading.Event()
    print_lock = threading.Lock()

    def emit():
        with print_lock:
            if not done.is_set():
                done.set()
                if rank == 0:
                    print(json.dumps(line, default=str), flush=True)

    def watchdog():
        if not done.wait(float(os.environ.get("B200_BENCH_EXTRAS_DEADLINE_S", "420"))):
            line["extras_timeout"] = True
            emit()
            os._exit(0)

    threading.Thread(target=watchdog, daemon=True).start()
    if not args.lean:
        extras["sustained"] = sustained_leg(bench, dist, dev, local_rank, total_flops, peaks)
    host_tensors = bench.host_tensors()
    del bench
    torch.cuda.empty_cache()

    if not args.lean:
        # ---- configs[2]: the same FFN with FSDP-sharded weights (tokens over 'data', weights dim0 over 'data') ----
        try:
            if world > 1:
                fb = layers.FFNBench(WORKLOADS["case3_fully_sharded"], args.act, rank, world, dev, chunks=args.chunks,
                                     host_tensors=host_tensors)
                fms, fl = time_resident(fb, dist, dev, 20, 3, capi)
                extras["fsdp"] = {"workload": "case3_fully_sharded (same FFN; x batch-sharded, W1/W2 dim0-sharded over 'data', "
                                              "all-gather of each weight before its GEMM, every step)",
                                  "value": total_flops / (fms * 1e-3) / 1e12, "unit": UNIT, "ms_per_step": fms, "steps": 20,
                                  "parallelism": fb.parallelism_name(), "gpu_launches": fl,
                                  "per_gpu_frac_of_burst_peak": total_flops / (fms * 1e-3) / 1e12 / world / peaks["bf16_burst"],
                                  "parity_check": ffn_parity_check(fb, args.act, dist, dev)}
                del fb
            else:
                extras["fsdp"] = {"workload": "case3_fully_sharded", "note": "on one GPU there is nothing to gather: the step is "
                                  "the main line's two GEMMs", "value": value, "unit": UNIT, "ms_per_step": ms}
        except Exception as e:   # noqa: BLE001 - never lose the main line over a supplement
            extras["fsdp"] = {"error": repr(e)}
        # ---- the public, reference-shaped call for the same step ----
        try:
            extras["api"] = api_leg(ljs, host_tensors, cfg, args.act, dist, dev, rank, world, total_flops, value)
        except Exception as e:   # noqa: BLE001
            extras["api"] = {"error": repr(e)}
        ljs.nn.invalidate_fast_layers()
        del host_tensors
        torch.cuda.empty_cache()
        # ---- configs[4]: head-sharded causal MHA, B16 S8192 32x128, on the (1,1)/(1,2)/(2,2)/(2,4) mesh ----
        if not args.no_attention:
            try:
                acfg = ATTN_WORKLOADS["case6_attention"]
                r = layers.run_attention_workload(acfg, rank, world, dev, steps=3, warmup=2, keep=True)
                tot = r["core_flops"] + r["proj_flops"]
                core_tf = r["core_flops"] / (r["parts_ms"]["attn"] * 1e-3) / 1e12
                extras["attention_case6"] = {
                    "workload": "case6_attention (B16, S8192, 32 heads x 128, causal; fused q/k/v projection + core + "
                                "row-parallel out-projection with the fused reduce-scatter)",
                    "mesh": "data x model = %dx%d (batch over data, heads over model)" % r["mesh"],
                    "value": tot / (r["ms_per_step"] * 1e-3) / 1e12, "unit": UNIT, "ms_per_step": r["ms_per_step"], "steps": 3,
                    "parts_ms": r["parts_ms"], "core_tflops": core_tf, "core_tflops_per_gpu": core_tf / world,
                    "core_frac_of_burst_peak_per_gpu": core_tf / world / peaks["bf16_burst"],
                    "per_gpu_frac_of_burst_peak": tot / (r["ms_per_step"] * 1e-3) / 1e12 / world / peaks["bf16_burst"],
                    "gpu_launches": r["launches"], "parity_check": attention_parity_check(r, acfg, dist, dev, rank, world)}
                del r
            except Exception as e:   # noqa: BLE001
                extras["attention_case6"] = {"error": repr(e)}
            torch.cuda.empty_cache()
        # ---- configs[0]: sharded fp32 lax.dot on real devices ----
        try:
            extras["case1a"] = case1a_leg(ljs, dist, dev, rank, world)
        except Exception as e:   # noqa: BLE001
            extras["case1a"] = {"error": repr(e)}

    if rank != 0:
        if dist is not None:
            dist.barrier()
        done.set()
        return
    if world == 1 and not args.no_attention and not args.lean:
        # supplementary: the MHA layer of configs[3] (case5_attention_dense shape) on this GPU
        acfg = ATTN_WORKLOADS["case5_attention_dense"]
        try:
            r = layers.run_attention_workload(acfg, 0, 1, dev, steps=5, warmup=2, keep=True)
            tot = r["core_flops"] + r["proj_flops"]
            line["attention"] = {
                "workload": "case5_attention_dense (B8, S4096, 32 heads x 128, non-causal; q/k/v/out projections + core)",
                "value": tot / (r["ms_per_step"] * 1e-3) / 1e12, "unit": UNIT, "ms_per_step": r["ms_per_step"],
                "core_tflops": r["core_flops"] / (r["parts_ms"]["attn"] * 1e-3) / 1e12,
                "core_frac_of_peak": r["core_flops"] / (r["parts_ms"]["attn"] * 1e-3) / 1e12 / peaks["bf16_burst"],
                "parts_ms": r["parts_ms"], "gpu_launches": r["launches"],
                "parity_check": attention_parity_check(r, acfg, None, dev, 0, 1)}
            del r
            torch.cuda.empty_cache()
            line["attention"]["d64"] = layers.attention_core_sample(dev, 8, 4096, 32, 64, False, peaks)
        except Exception as e:   # noqa: BLE001
            line["attention"] = {"error": repr(e)}
        # supplementary: the kernels the reference's train_step adds at the same shape
        try:
            line["train_kernels"] = train_kernel_sample(dev, peaks)
        except Exception as e:   # noqa: BLE001 - never lose the main line over the supplement
            line["train_kernels"] = {"error": repr(e)}
    if not args.no_cpu_baseline and world == 1:
        line["cpu_baseline"] = cpu_ffn_sample(cfg, act=args.act)
    else:
        line["cpu_baseline"] = None
    emit()
    if dist is not None:
        dist.barrier()


if __name__ == "__main__":
    main()
