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
        y = fp.forward(xd)
        torch.cuda.synchronize()
        print(f"[rank {rank}] batch={batch} step {it} done; counters={fp.rs.counters.tolist()}", flush=True)
    ref = want[fp.rs.global_rows(torch.arange(T // world), rank)]
    err = (y.cpu().float() - ref).abs()
    print(f"[rank {rank}] batch={batch} max_abs={err.max().item():.3e} ok={bool((err <= 2e-2 * ref.abs() + 2e-2).all())}", flush=True)
    del fp
dist.barrier()
dist.destroy_process_group()
