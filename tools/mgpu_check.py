"""Multi-GPU bring-up (run under torchrun on the GPU box): the resident-weight sharded FFN layers
against the CPU oracle on a reduced shape, rank by rank.

  python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29533 tools/mgpu_check.py
"""
import os
import sys

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import learning_jax_sharding_b200 as ljs  # noqa: E402
from learning_jax_sharding_b200 import layers  # noqa: E402
from oracle import reference_math as rm  # noqa: E402


def main():
    rt = ljs.init_distributed("nccl")
    import torch.distributed as dist
    rank, world = rt.rank, rt.world
    dev = rt.torch_device
    T, d, f = 4096, 1024, 4096
    x = rm.synthetic((T, d), 0)
    w1, b1 = rm.synthetic((d, f), 1, "weight"), rm.synthetic((f,), 2, "bias")
    w2, b2 = rm.synthetic((f, d), 3, "weight"), rm.synthetic((d,), 4, "bias")
    want = rm.ffn_forward(x, w1, b1, w2, b2, "gelu_tanh").float()
    bf = lambda t: t.to(torch.bfloat16).to(dev)
    ok = True
    # tensor parallel: W1 columns / W2 rows over the ranks, reduce-scatter to sequence blocks
    fl = f // world
    for chunks in (1, 4):
        tp = layers.TensorParallelFFN(bf(w1[:, rank * fl:(rank + 1) * fl].contiguous()), bf(b1[rank * fl:(rank + 1) * fl]),
                                      bf(w2[rank * fl:(rank + 1) * fl].contiguous()), bf(b2), "gelu_tanh", T,
                                      dist.group.WORLD, chunks=chunks)
        for it in range(2):
            y = tp.forward(bf(x))
        torch.cuda.synchronize()
        blk = T // world
        err = (y.cpu().float() - want[rank * blk:(rank + 1) * blk]).abs()
        good = bool((err <= 2e-2 * want[rank * blk:(rank + 1) * blk].abs() + 2e-2).all())
        print(f"[rank {rank}] TP chunks={tp.chunks}: max_abs={err.max().item():.3e} {'OK' if good else 'FAIL'}", flush=True)
        ok &= good
    # fused: GEMM2 epilogue pushes partial tiles into the owners' receive slots over NVLink peer memory
    fused = layers.FusedTensorParallelFFN(bf(w1[:, rank * fl:(rank + 1) * fl].contiguous()), bf(b1[rank * fl:(rank + 1) * fl]),
                                          bf(w2[rank * fl:(rank + 1) * fl].contiguous()), bf(b2), "gelu_tanh", T,
                                          dist.group.WORLD, chunks=2)
    for it in range(5):     # several steps: both receive buffers and the barrier epochs are exercised
        y = fused.forward(bf(x))
    torch.cuda.synchronize()
    blk = T // world
    err = (y.cpu().float() - want[rank * blk:(rank + 1) * blk]).abs()
    good = bool((err <= 2e-2 * want[rank * blk:(rank + 1) * blk].abs() + 2e-2).all())
    print(f"[rank {rank}] fused TP (peer-memory push, {fused.rs.chunks} chunk(s)): max_abs={err.max().item():.3e} "
          f"{'OK' if good else 'FAIL'}", flush=True)
    ok &= good
    # the single-launch form gives the same bits (same tiles, same reduction order), and the chunked one is not slower
    fused1 = layers.FusedTensorParallelFFN(bf(w1[:, rank * fl:(rank + 1) * fl].contiguous()), bf(b1[rank * fl:(rank + 1) * fl]),
                                           bf(w2[rank * fl:(rank + 1) * fl].contiguous()), bf(b2), "gelu_tanh", T,
                                           dist.group.WORLD, chunks=1)
    for it in range(3):
        y1 = fused1.forward(bf(x))
    torch.cuda.synchronize()
    good = bool(torch.equal(y1, y))
    print(f"[rank {rank}] fused TP chunks=1 vs chunks={fused.rs.chunks}: {'identical' if good else 'DIFFERENT'}", flush=True)
    ok &= good
    del fused1
    # progressive form: per-raster-group progress flags from the pushing GEMM, reduction beside it on a side stream
    os.environ["B200_TP_PROGRESSIVE"] = "1"
    Tp = 2 * T                                         # two raster groups of 4096 rows
    xp = torch.cat([x, rm.synthetic((T, d), 50)])
    wantp = rm.ffn_forward(xp, w1, b1, w2, b2, "gelu_tanh").float()
    for batch in (1, 4):
        fp = layers.FusedTensorParallelFFN(bf(w1[:, rank * fl:(rank + 1) * fl].contiguous()), bf(b1[rank * fl:(rank + 1) * fl]),
                                           bf(w2[rank * fl:(rank + 1) * fl].contiguous()), bf(b2), "gelu_tanh", Tp,
                                           dist.group.WORLD, batch=batch)
        for it in range(5):
            yp = fp.forward(bf(xp))
        torch.cuda.synchronize()
        ref = wantp[fp.rs.global_rows(torch.arange(Tp // world), rank)]
        err = (yp.cpu().float() - ref).abs()
        good = bool((err <= 2e-2 * ref.abs() + 2e-2).all()) and fp.rs.progressive
        print(f"[rank {rank}] fused TP, progressive reduction ({fp.rs.groups} groups, batch={batch}): max_abs={err.max().item():.3e} "
              f"{'OK' if good else 'FAIL'}", flush=True)
        ok &= good
        del fp
    os.environ["B200_TP_PROGRESSIVE"] = "0"
    # scatter along seq of a [batch, seq] activation (two-level A view): [4, T/4/W, d] per rank, chunks over batch items
    fused_s = layers.FusedTensorParallelFFN(bf(w1[:, rank * fl:(rank + 1) * fl].contiguous()), bf(b1[rank * fl:(rank + 1) * fl]),
                                            bf(w2[rank * fl:(rank + 1) * fl].contiguous()), bf(b2), "gelu_tanh", T,
                                            dist.group.WORLD, batch=4, chunks=2)
    for it in range(3):
        ys = fused_s.forward(bf(x))
    torch.cuda.synchronize()
    ref = want[fused_s.rs.global_rows(torch.arange(T // world), rank)]
    err = (ys.cpu().float() - ref).abs()
    good = bool((err <= 2e-2 * ref.abs() + 2e-2).all())
    print(f"[rank {rank}] fused TP, scatter along seq of [4, {T // 4}] ({fused_s.rs.chunks} chunk(s)): max_abs={err.max().item():.3e} "
          f"{'OK' if good else 'FAIL'}", flush=True)
    ok &= good
    del fused_s
    # head-sharded MHA layer: fused qkv GEMM, local-head attention, row-parallel out-projection with the push reduce-scatter
    Bm, Sm, Mm, Hm, Dm = 2, 1024, 512, 8, 64
    xm = rm.synthetic((Bm, Sm, Mm), 20)
    pm = {"to_q": rm.synthetic((Mm, Hm * Dm), 21, "weight"), "to_k": rm.synthetic((Mm, Hm * Dm), 22, "weight"),
          "to_v": rm.synthetic((Mm, Hm * Dm), 23, "weight"),
          "to_out_0": (rm.synthetic((Hm * Dm, Mm), 24, "weight"), rm.synthetic((Mm,), 25, "bias"))}
    want_m = rm.mha_forward(xm, pm, Hm, Dm, causal=True).float().reshape(Bm * Sm, Mm)
    hl = Hm // world
    cs = slice(rank * hl * Dm, (rank + 1) * hl * Dm)
    mha = layers.ShardedMHA(bf(pm["to_q"][:, cs].contiguous()), bf(pm["to_k"][:, cs].contiguous()),
                            bf(pm["to_v"][:, cs].contiguous()), bf(pm["to_out_0"][0][cs].contiguous()),
                            bf(pm["to_out_0"][1]), hl, Dm, Bm, Sm, True, dist.group.WORLD)
    for it in range(3):
        ym = mha.forward(bf(xm))
    torch.cuda.synchronize()
    blk = Bm * Sm // world
    ref = want_m[mha.rs.global_rows(torch.arange(blk), rank)]          # this rank's block along S of every batch item
    err = (ym.cpu().float() - ref).abs()
    good = bool((err <= 2e-2 * ref.abs() + 2e-2).all())
    print(f"[rank {rank}] head-sharded MHA (causal): max_abs={err.max().item():.3e} {'OK' if good else 'FAIL'}", flush=True)
    ok &= good
    # FSDP: tokens over ranks, weights dim0-sharded, all-gather in column chunks
    tl, k1, k2 = T // world, d // world, f // world
    fs = layers.FsdpFFN(bf(w1[rank * k1:(rank + 1) * k1].contiguous()), bf(b1), bf(w2[rank * k2:(rank + 1) * k2].contiguous()),
                        bf(b2), "gelu_tanh", tl, dist.group.WORLD, chunks=4)
    for it in range(3):                           # both buffer parities of the forward prefetch
        y = fs.forward(bf(x[rank * tl:(rank + 1) * tl].contiguous()))
    torch.cuda.synchronize()
    err = (y.cpu().float() - want[rank * tl:(rank + 1) * tl]).abs()
    good = bool((err <= 2e-2 * want[rank * tl:(rank + 1) * tl].abs() + 2e-2).all())
    print(f"[rank {rank}] FSDP: max_abs={err.max().item():.3e} {'OK' if good else 'FAIL'}", flush=True)
    ok &= good
    pg = layers.PeerGatherFFN(bf(w1[rank * k1:(rank + 1) * k1].contiguous()), bf(b1), bf(w2[rank * k2:(rank + 1) * k2].contiguous()),
                              bf(b2), "gelu_tanh", tl, dist.group.WORLD, chunks=4)
    for it in range(3):
        y = pg.forward(bf(x[rank * tl:(rank + 1) * tl].contiguous()))
    torch.cuda.synchronize()
    err = (y.cpu().float() - want[rank * tl:(rank + 1) * tl]).abs()
    good = bool((err <= 2e-2 * want[rank * tl:(rank + 1) * tl].abs() + 2e-2).all())
    print(f"[rank {rank}] FSDP peer-pull gather: max_abs={err.max().item():.3e} {'OK' if good else 'FAIL'}", flush=True)
    ok &= good
    ph = layers.PushGatherFFN(bf(w1[rank * k1:(rank + 1) * k1].contiguous()), bf(b1), bf(w2[rank * k2:(rank + 1) * k2].contiguous()),
                              bf(b2), "gelu_tanh", tl, dist.group.WORLD)
    for it in range(4):                           # both buffer parities, barrier epochs, the prefetch of the next call's W1
        y = ph.forward(bf(x[rank * tl:(rank + 1) * tl].contiguous()))
    torch.cuda.synchronize()
    err = (y.cpu().float() - want[rank * tl:(rank + 1) * tl]).abs()
    good = bool((err <= 2e-2 * want[rank * tl:(rank + 1) * tl].abs() + 2e-2).all())
    good &= bool(torch.equal(ph.w1g[0].cpu(), w1.to(torch.bfloat16))) and bool(torch.equal(ph.w2g[1].cpu(), w2.to(torch.bfloat16)))
    print(f"[rank {rank}] FSDP push gather: max_abs={err.max().item():.3e} {'OK' if good else 'FAIL'}", flush=True)
    ok &= good
    # the generic host layer on a real 1 x world mesh: nn.ffn with NCCL collectives
    P = ljs.PartitionSpec
    mesh = ljs.Mesh(ljs.create_device_mesh((1, world)), ("data", "model"))
    ns = lambda s: ljs.NamedSharding(mesh, s)
    y = ljs.nn.ffn(ljs.device_put(x.reshape(4, T // 4, d), ns(P("data"))), ljs.device_put(w1, ns(P(None, "model"))),
                   ljs.device_put(b1, ns(P("model"))), ljs.device_put(w2, ns(P("model", None))), ljs.device_put(b2, ns(P())),
                   scatter_dim=1)
    torch.cuda.synchronize()
    blk = T // 4 // world
    got = y.device_buffers[0].cpu().float()
    ref = want.reshape(4, T // 4, d)[:, rank * blk:(rank + 1) * blk]
    good = bool(((got - ref).abs() <= 2e-2 * ref.abs() + 2e-2).all())
    fast_used = any(k[0] == "ffn" and v is not False for k, v in ljs.nn._fast_layers.items())
    print(f"[rank {rank}] nn.ffn (NamedSharding, scatter_dim=1; fused fast path used: {fast_used}): {'OK' if good else 'FAIL'}",
          flush=True)
    ok &= good and fast_used
    # ... and the same call through jit(apply_fn, in_shardings, out_shardings), twice (cached layer, reused buffers)
    params_ffn = (ljs.device_put(w1, ns(P(None, "model"))), ljs.device_put(b1, ns(P("model"))),
                  ljs.device_put(w2, ns(P("model", None))), ljs.device_put(b2, ns(P())))
    apply_fn = ljs.nn.jit(lambda prm, xx: ljs.nn.ffn(xx, *prm, scatter_dim=1), in_shardings=(None, ns(P("data"))),
                          out_shardings=ns(P("data", "model")))
    xin = ljs.device_put(x.reshape(4, T // 4, d), ns(P("data")))
    for it in range(2):
        y = apply_fn(params_ffn, xin)
    torch.cuda.synchronize()
    good = bool(((y.device_buffers[0].cpu().float() - ref).abs() <= 2e-2 * ref.abs() + 2e-2).all())
    print(f"[rank {rank}] nn.jit(apply_fn) FFN: {'OK' if good else 'FAIL'}", flush=True)
    ok &= good
    # the attention module through the reference-shaped API on real devices (case6_attention.py:229-237): the head-sharded
    # annotation set (fused fast path: layers.ShardedMHA, reduce-scatter along S) and the LITERAL annotation set (generic
    # plans over NCCL: weight all-gather, K / V all-gather along S, all-to-all M -> S)
    d_ax = 2 if world >= 4 else 1
    mesh6 = ljs.Mesh(ljs.create_device_mesh((d_ax, world // d_ax)), ("data", "model"))
    x6_host = rm.synthetic((4, 1024, 512), 30)
    p6 = {"to_q": rm.synthetic((512, 512), 31, "weight"), "to_k": rm.synthetic((512, 512), 32, "weight"),
          "to_v": rm.synthetic((512, 512), 33, "weight"),
          "to_out_0": (rm.synthetic((512, 512), 34, "weight"), rm.synthetic((512,), 35, "bias"))}
    want6 = rm.mha_forward(x6_host, p6, 8, 64).float()
    x_sh = ljs.NamedSharding(mesh6, P("data", "model"))
    x6 = ljs.device_put(x6_host, x_sh)
    for layout, rules in (("heads", (("batch", "data"), ("heads", "model"), ("hidden", "model"))),
                          ("literal", (("batch", "data"), ("embed", "model"), ("hidden", "model")))):
        model = ljs.nn.FlaxAttention(512, heads=8, dim_head=64, layout=layout)
        prm = model.shard_params(p6, mesh6, rules)
        fn = ljs.nn.jit(lambda pp, xx: model(pp, xx), in_shardings=(None, x_sh), out_shardings=x_sh)
        n_before = len(ljs.nn._fast_layers)
        with mesh6, ljs.nn.axis_rules(rules):
            for it in range(2):
                y6 = fn(prm, x6)
        torch.cuda.synchronize()
        sh = y6.addressable_shards[0]
        got6, ref6 = sh.data.cpu().float(), want6[sh.index]
        good = bool(((got6 - ref6).abs() <= 2e-2 * ref6.abs() + 2e-2).all())
        fast = len(ljs.nn._fast_layers) > n_before
        print(f"[rank {rank}] FlaxAttention layout={layout} on mesh {d_ax}x{world // d_ax} via nn.jit (fused layer: {fast}): "
              f"max_abs={(got6 - ref6).abs().max().item():.3e} {'OK' if good else 'FAIL'}", flush=True)
        ok &= good and (fast or layout == "literal")
        if layout == "literal":
            # K / V blocks went round the ring of the S-sharding group (NCCL send / receive), never gathered
            ringed = "ring_shift" in [n for n, _, _ in rt.collective_log] or world // d_ax == 1
            # ... and the train_step gradients under the same annotations (reduce-scatter along S, all-reduce of dW)
            with mesh6, ljs.nn.axis_rules(rules):
                g6 = ljs.train.attention_sum_loss_grads(model, prm, x6)
            w6 = rm.mha_sum_loss_grads(x6_host, p6, 8, 64)
            gg = all(bool(torch.allclose(g6[n].global_array().cpu(), w6[n], rtol=3e-2, atol=3e-2 * w6[n].abs().max().item()))
                     for n in ("to_q", "to_k", "to_v", "to_out_0.kernel", "to_out_0.bias"))
            print(f"[rank {rank}] literal layout: ring attention used: {ringed}; train_step gradients "
                  f"{'OK' if gg else 'FAIL'}", flush=True)
            ok &= ringed and gg
    # the sharded fp32 lax.dot of case1a/1b/2/3/4 on real devices: the literal 2x4 PositionalSharding layouts need 8 GPUs;
    # at any size the 1-D mesh form of BASELINE.json configs[0]
    gA = torch.Generator().manual_seed(0)
    A4, B4 = torch.randn(4, 16, generator=gA), torch.randn(16, 4, generator=gA)
    want4 = rm.dot_f32(A4, B4)
    lay = {}
    if world == 8:
        s8 = ljs.PositionalSharding(ljs.create_device_mesh((2, 4)))
        lay = {"case1a": (s8.replicate(axis=0, keepdims=True), s8.reshape(4, 2).replicate(axis=1, keepdims=True)),
               "case1b": (s8.replicate(axis=0, keepdims=True), s8.replicate(axis=1, keepdims=True)),
               "case2": (s8, s8.replicate(axis=1, keepdims=True)), "case3": (s8, s8),
               "case4": (s8.replicate(axis=1, keepdims=True), s8.replicate(axis=0, keepdims=True))}
    m1 = ljs.Mesh(ljs.create_device_mesh((world,)), ("m",))
    lay["mesh_1d"] = (ljs.NamedSharding(m1, P(None, "m")), ljs.NamedSharding(m1, P("m", None)))
    for name, (sa, sb) in lay.items():
        C4 = ljs.lax.dot(ljs.device_put(A4, sa), ljs.device_put(B4, sb))
        good = all(bool(torch.allclose(s.data.cpu(), want4[s.index], rtol=1e-5, atol=1e-6)) for s in C4.addressable_shards)
        print(f"[rank {rank}] lax.dot {name} on {world} GPUs: shard {tuple(C4.addressable_shards[0].data.shape)} "
              f"{'OK' if good else 'FAIL'}", flush=True)
        ok &= good
    # the C-ABI mesh communicators (b200_mesh_* / b200_all_reduce ...): every collective on every axis of a (2, world/2)
    # mesh (or a 1-D mesh) against values computed locally (every rank can reconstruct the others' inputs)
    uid = [ljs.capi.comm_unique_id() if rank == 0 else None]
    dist.broadcast_object_list(uid, src=0)
    shape = (2, world // 2) if world % 2 == 0 and world > 2 else (world,)
    mc = ljs.capi.MeshComm(uid[0], world, rank, shape)
    def coords_of(q):                              # row-major: last axis fastest
        c = []
        for sz in reversed(shape):
            c.insert(0, q % sz)
            q //= sz
        return c
    coords = coords_of(rank)
    good = True
    for axis, n in enumerate(shape):
        members = [q for q in range(world)
                   if all(cq == cr for i, (cq, cr) in enumerate(zip(coords_of(q), coords)) if i != axis)]
        good &= mc.axis_size(axis) == n and mc.axis_rank(axis) == coords[axis] and len(members) == n
        for dt in (torch.float32, torch.bfloat16):
            base = torch.arange(n * 8 * 16, device=dev, dtype=torch.float32).reshape(n * 8, 16) / 64.0
            mine = (base + rank).to(dt)
            theirs = [(base + q).to(dt) for q in members]
            want_sum = torch.stack([t.float() for t in theirs]).sum(0)
            got = mc.all_reduce(mine.clone(), axis)
            good &= bool(torch.allclose(got.float(), want_sum, rtol=2e-2, atol=2e-2))
            got = mc.reduce_scatter(mine.clone(), axis)
            blk = slice(coords[axis] * 8, (coords[axis] + 1) * 8)
            good &= bool(torch.allclose(got.float(), want_sum[blk], rtol=2e-2, atol=2e-2))
            got = mc.all_gather(mine[:8].contiguous(), axis)
            good &= bool(torch.equal(got, torch.cat([t[:8] for t in theirs])))
            got = mc.all_to_all(mine.clone(), axis)
            good &= bool(torch.equal(got, torch.cat([t[blk] for t in theirs])))
    # b200_ffn_fwd_tp: the tensor-parallel FFN with its reduction inside the C call, over the LAST mesh axis; ranks that
    # differ along the other axis hold the same shard and must agree
    ax = len(shape) - 1
    n_ax, r_ax = shape[ax], coords[ax]
    fl2 = f // n_ax
    w1s, b1s = bf(w1[:, r_ax * fl2:(r_ax + 1) * fl2].contiguous()), bf(b1[r_ax * fl2:(r_ax + 1) * fl2])
    w2s = bf(w2[r_ax * fl2:(r_ax + 1) * fl2].contiguous())
    for rs in (False, True):
        for ch in (1, 2):
            y = mc.ffn_fwd_tp(ax, bf(x), w1s, b1s, w2s, bf(b2), reduce_scatter=rs, chunks=ch)
            torch.cuda.synchronize()
            blk2 = T // n_ax
            ref = want[r_ax * blk2:(r_ax + 1) * blk2] if rs else want
            g2 = bool(((y.cpu().float() - ref).abs() <= 2e-2 * ref.abs() + 2e-2).all())
            if not g2:
                print(f"[rank {rank}] b200_ffn_fwd_tp reduce_scatter={rs} chunks={ch}: FAIL", flush=True)
            good &= g2
    torch.cuda.synchronize()
    mc.destroy()
    print(f"[rank {rank}] C-ABI mesh collectives + b200_ffn_fwd_tp on mesh {shape}: {'OK' if good else 'FAIL'}", flush=True)
    ok &= good
    # the streaming host-buffer step bench.py times as `e2e`, in both shardings: several overlapping calls, every returned
    # pinned buffer against the oracle rows this rank owns
    for mode in ("tp", "fsdp"):
        cfg = dict(batch=2, seq=T // 2, d_model=d, d_ff=f, parallelism=mode)
        fb = layers.FFNBench(cfg, "gelu_tanh", rank, world, dev)
        outs = [fb.step_e2e() for _ in range(5)]
        fb.e2e_drain()
        torch.cuda.synchronize()
        blk = T // world
        ref = want[rank * blk:(rank + 1) * blk]
        good = all(bool(((o.float() - ref).abs() <= 2e-2 * ref.abs() + 2e-2).all()) for o in outs[-2:])
        good &= outs[0].data_ptr() != outs[1].data_ptr()
        print(f"[rank {rank}] streaming host-buffer step ({mode}, {fb.parallelism_name()}): {'OK' if good else 'FAIL'}", flush=True)
        ok &= good
        del fb
    t = torch.tensor([1 if ok else 0], device=dev)
    dist.all_reduce(t, op=dist.ReduceOp.MIN)
    if rank == 0:
        print("RESULT", "PASS" if t.item() == 1 else "FAIL", flush=True)
    dist.barrier()
    dist.destroy_process_group()
    sys.exit(0 if t.item() == 1 else 1)


if __name__ == "__main__":
    main()
