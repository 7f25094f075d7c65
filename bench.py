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
    steps = min(args.steps, 20)
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
            "config": {"workload": "case4_gspmd_ff", "batch": cfg["batch"], "seq_len": cfg["seq"],
                       "d_model": d, "d_ff": f, "activation": args.act, "bias": True},
            "cpu_baseline": {"value": value, "unit": UNIT, "cores": torch.get_num_threads(), "kind": "port",
                             "sample": sample},
            "e2e": {"value": value, "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0}}
    print(json.dumps(line), flush=True)


# ------------------------------------------------------------------------------ GPU legs
def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=50)
    ap.add_argument("--warmup", type=int, default=5)
    ap.add_argument("--impl", default="b200", choices=["b200", "reference"])
    ap.add_argument("--workload", default="case4_gspmd_ff")
    ap.add_argument("--act", default="gelu_tanh", choices=["gelu_tanh", "relu"])
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--no-attention", action="store_true", help="skip the supplementary attention measurement")
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
    torch.cuda.set_device(local_rank)
    dist = None
    if world > 1:
        ljs.init_distributed("nccl")
        import torch.distributed as dist
    dev = torch.device("cuda", local_rank)

    if args.workload in ATTN_WORKLOADS:
        return layers.bench_attention(args, ATTN_WORKLOADS[args.workload], rank, world, dev, METRIC, UNIT,
                                      measured_peaks(), ClockSampler(local_rank), attn_flops)
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
    for _ in range(args.warmup):
        bench.step_resident()
    sync_all()
    sampler = ClockSampler(local_rank)
    sampler.start()
    launches0 = capi.launch_count()
    ev = [torch.cuda.Event(enable_timing=True) for _ in range(2)]
    bench.reset_kernel_events()
    sync_all()
    ev[0].record()
    for _ in range(args.steps):
        bench.step_resident(record_kernels=True)
    ev[1].record()
    sync_all()
    launches = capi.launch_count() - launches0
    clocks = sampler.stop()
    ms = ev[0].elapsed_time(ev[1]) / args.steps
    if dist is not None:
        t = torch.tensor([ms], device=dev)
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        ms = float(t.item())
    value = total_flops / (ms * 1e-3) / 1e12
    kernel_ms, n_gemm = bench.mean_kernel_ms()    # mean duration of one GEMM launch inside the timed region
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

    if rank != 0:
        if dist is not None:
            dist.barrier()
        return
    peaks = measured_peaks()
    flops_per_launch = total_flops / world / gemm_per_step   # GEMM launches of one step on one GPU share its FLOPs equally
    achieved = flops_per_launch / (kernel_ms * 1e-3) / 1e12
    line = {
        "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": world, "steps": args.steps, "warmup": args.warmup,
        "ms_per_step": ms, "higher_is_better": True, "scaling": "strong", "vs_baseline": None, "dtype": "bf16",
        "data": "synthetic",
        "config": {"workload": args.workload, "batch": cfg["batch"], "seq_len": cfg["seq"], "d_model": cfg["d_model"],
                   "d_ff": cfg["d_ff"], "activation": args.act, "bias": True, "parallelism": bench.parallelism_name(),
                   "l2": "inputs larger than L2 (x 134 MB, W1 134 MB, W2 134 MB, h 537 MB per step vs 126 MB L2)",
                   "flops_per_step": total_flops},
        "clocks": clocks,
        "e2e": {"value": total_flops / (e2e_ms * 1e-3) / 1e12, "unit": UNIT, "ms_per_step": e2e_ms,
                "h2d_bytes_per_step": h2d, "d2h_bytes_per_step": d2h,
                "what": "x copied from pinned host memory and y copied back every step, weights resident"},
        "gpu_launches": int(launches),
        "roofline": {"bound": "tensor", "kernel": "gemm_bf16_kernel (tcgen05, fused bias+activation epilogue)",
                     "achieved": achieved, "peak": peaks["bf16_burst"], "unit": UNIT,
                     "frac": achieved / peaks["bf16_burst"], "peak_sustained": peaks["bf16_sustained"],
                     "frac_of_sustained": achieved / peaks["bf16_sustained"], "frac_of_nominal_2250": achieved / 2250.0,
                     "peak_source": peaks["source"], "launch_ms": kernel_ms, "flops_per_launch": flops_per_launch, "gemm_launches_per_step": gemm_per_step,
                     "traffic": bench.ncu_traffic_bytes()},
    }
    if world == 1 and not args.no_attention:
        # supplementary: the MHA layer of configs[3] (case5_attention_dense shape) on this GPU
        del bench
        torch.cuda.empty_cache()
        acfg = ATTN_WORKLOADS["case5_attention_dense"]
        r = layers.run_attention_workload(acfg, 0, 1, dev, steps=5, warmup=2)
        tot = r["core_flops"] + r["proj_flops"]
        line["attention"] = {
            "workload": "case5_attention_dense (B8, S4096, 32 heads x 128, non-causal; q/k/v/out projections + core)",
            "value": tot / (r["ms_per_step"] * 1e-3) / 1e12, "unit": UNIT, "ms_per_step": r["ms_per_step"],
            "core_tflops": r["core_flops"] / (r["parts_ms"]["attn"] * 1e-3) / 1e12,
            "core_frac_of_peak": r["core_flops"] / (r["parts_ms"]["attn"] * 1e-3) / 1e12 / peaks["bf16_burst"],
            "parts_ms": r["parts_ms"], "gpu_launches": r["launches"]}
        # supplementary: the kernels the reference's train_step adds at the same shape
        try:
            line["train_kernels"] = train_kernel_sample(dev, peaks)
        except Exception as e:   # noqa: BLE001 - never lose the main line over the supplement
            line["train_kernels"] = {"error": repr(e)}
    if not args.no_cpu_baseline and world == 1:
        line["cpu_baseline"] = cpu_ffn_sample(cfg, act=args.act)
    else:
        line["cpu_baseline"] = None
    print(json.dumps(line), flush=True)
    if dist is not None:
        dist.barrier()


if __name__ == "__main__":
    main()
