"""What does pushing the row-parallel partial product over NVLink cost the GEMM?  (bring-up diagnostic; torchrun, >= 2 GPUs)
The second FFN GEMM at its 8-GPU per-rank shape (16384 x 4096 x 2048), timed alone with its output (a) in local memory,
plain store, (b) scattered to LOCAL row-block slots, (c) scattered to the peers' slots (the fused reduce-scatter push)."""
import os, sys
import torch
ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)
import learning_jax_sharding_b200 as ljs
from learning_jax_sharding_b200 import capi, layers

rt = ljs.init_distributed("nccl")
import torch.distributed as dist
rank, world, dev = rt.rank, rt.world, rt.torch_device
W = int(os.environ.get("PUSH_W", "8"))             # logical owner blocks (peers are reused round-robin when world < W)


def timed(fn, iters=30):
    for _ in range(5):
        fn()
    torch.cuda.synchronize(); dist.barrier(); torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(iters):
        fn()
    e1.record(); torch.cuda.synchronize()
    t = torch.tensor([e0.elapsed_time(e1) / iters], device=dev)
    dist.all_reduce(t, op=dist.ReduceOp.MAX)
    return float(t.item())


for (M, N, K) in [(16384, 4096, 2048), (16384, 4096, 4096), (65536, 4096, 1024)]:
    a = torch.randn(M, K, device=dev).to(torch.bfloat16)
    w = (torch.randn(K, N, device=dev) / K ** 0.5).to(torch.bfloat16)
    out = torch.empty(M, N, device=dev, dtype=torch.bfloat16)
    R = M // W
    base, handle = capi.peer_alloc(M * N * 2)
    mine = torch.tensor(list(handle), dtype=torch.uint8, device=dev)
    handles = [torch.empty_like(mine) for _ in range(world)]
    dist.all_gather(handles, mine)
    peer = [base if p == rank else capi.peer_open(bytes(handles[p].cpu().tolist())) for p in range(world)]
    local_ptrs = torch.tensor([base + j * R * N * 2 for j in range(W)], dtype=torch.int64, device=dev)
    # owner j = rank (j mod world): every rank writes slot `rank-th R rows`... distinct regions per sender are not needed for timing
    remote_ptrs = torch.tensor([peer[j % world] + j * R * N * 2 for j in range(W)], dtype=torch.int64, device=dev)
    fl = 2.0 * M * N * K
    t_plain = timed(lambda: capi.gemm_bf16(a, w, out=out))
    t_local = timed(lambda: capi.gemm_bf16_scatter(a, w, local_ptrs, R, (rank + 1) % W, N))
    torch_slots = torch.empty(M, N, device=dev, dtype=torch.bfloat16)
    tptrs = torch.tensor([torch_slots.data_ptr() + j * R * N * 2 for j in range(W)], dtype=torch.int64, device=dev)
    t_torch = timed(lambda: capi.gemm_bf16_scatter(a, w, tptrs, R, (rank + 1) % W, N))
    t_rot0 = timed(lambda: capi.gemm_bf16_scatter(a, w, tptrs, R, 0, N))
    t_plain2 = timed(lambda: capi.gemm_bf16(a, w, out=out))
    t_remote = timed(lambda: capi.gemm_bf16_scatter(a, w, remote_ptrs, R, (rank + 1) % W, N))
    t_remote_il = timed(lambda: capi.gemm_bf16_scatter(a, w, remote_ptrs, R, (rank + 1) % W, N, interleave=True))
    if rank == 0:
        print(f"M={M} N={N} K={K}: plain {t_plain:.4f} ms ({fl / t_plain / 1e9:.0f} TFLOP/s) | scatter local {t_local:.4f} "
              f"({fl / t_local / 1e9:.0f}) | scatter to peers {t_remote:.4f} ({fl / t_remote / 1e9:.0f}); "
              f"{(world - 1) / world * M * N * 2 / t_remote / 1e6:.0f} GB/s leave each GPU | scatter into torch memory {t_torch:.4f}, "
              f"rot 0 {t_rot0:.4f}, plain again {t_plain2:.4f}", flush=True)
    dist.barrier()
dist.destroy_process_group()
