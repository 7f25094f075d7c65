"""Progressive reduce-scatter bring-up on real GPUs (torchrun)."""
import os, sys
import torch
ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)
os.environ["B200_TP_PROGRESSIVE"] = "1"
import learning_jax_sharding_b200 as ljs
from learning_jax_sharding_b200 import layers, capi
from oracle import reference_math as rm
rt = ljs.init_distributed("nccl")
import torch.distributed as dist
rank, world, dev = rt.rank, rt.world, rt.torch_device
os.makedirs(os.path.join(ROOT, "gpurun_out"), exist_ok=True)
_log = open(os.path.join(ROOT, "gpurun_out", f"prog_rank{rank}.log"), "w")
def print(*a, **k):      # noqa: A001 - per-rank log file: survives a device trap
    _log.write(" ".join(str(x) for x in a) + "\n"); _log.flush(); os.fsync(_log.fileno())
T, d, f = 8192, 1024, 4096
x = rm.synthetic((T, d), 0)
w1, b1 = rm.synthetic((d, f), 1, "weight"), rm.synthetic((f,), 2, "bias")
w2, b2 = rm.synthetic((f, d), 3, "weight"), rm.synthetic((d,), 4, "bias")
want = rm.ffn_forward(x, w1, b1, w2, b2, "gelu_tanh").float()
bf = lambda t: t.to(torch.bfloat16).to(dev)
fl = f // world
for batch in (1, 4):
    fp = layers.FusedTensorParallelFFN(bf(w1[:, rank * fl:(rank + 1) * fl].contiguous()), bf(b1[rank * fl:(rank + 1) * fl]),
                                       bf(w2[rank * fl:(rank + 1) * fl].contiguous()), bf(b2), "gelu_tanh", T, dist.group.WORLD, batch=batch)
    print(f"[rank {rank}] batch={batch} progressive={fp.rs.progressive} groups={getattr(fp.rs, 'groups', None)}", flush=True)
    xd = bf(x)
    for it in range(4):
        print(f"[rank {rank}] step {it} enqueue")
        y = fp.forward(xd)
        print(f"[rank {rank}] step {it} enqueued; epoch={fp.rs.pepoch} buf={fp.rs.buf}")
        torch.cuda.synchronize()
        fl_local = torch.as_tensor(layers._RawCudaBuffer(fp.rs.base + 1024, 64), device=dev).view(torch.int32)
        print(f"[rank {rank}] flags={fl_local[:8].tolist()}")
        print(f"[rank {rank}] batch={batch} step {it} done; counters={fp.rs.counters.tolist()}", flush=True)
    ref = want[fp.rs.global_rows(torch.arange(T // world), rank)]
    err = (y.cpu().float() - ref).abs()
    print(f"[rank {rank}] batch={batch} max_abs={err.max().item():.3e} ok={bool((err <= 2e-2 * ref.abs() + 2e-2).all())}", flush=True)
    del fp
dist.barrier()
dist.destroy_process_group()
