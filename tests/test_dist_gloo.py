"""The one-process-per-device path on CPU: world_size 2, gloo backend, per-shard compute routed to
the oracle-backed checker.  Covers the collectives of DistributedRuntime and that every rank
computes the same plan."""
import os
import sys

import pytest
import torch
import torch.multiprocessing as mp

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _worker(rank, world, port, results):
    sys.path.insert(0, ROOT)
    os.environ.update(RANK=str(rank), WORLD_SIZE=str(world), LOCAL_RANK=str(rank), MASTER_ADDR="127.0.0.1",
                      MASTER_PORT=str(port))
    import learning_jax_sharding_b200 as ljs
    from oracle import reference_math as rm
    from tests import _cpu_checker
    ljs.runtime.set_local_compute_for_testing(_cpu_checker)
    rt = ljs.init_distributed("gloo")
    P = ljs.PartitionSpec
    ok = {}
    # (1) contraction-sharded fp32 matmul on a 1-D mesh: partial products + all-reduce
    mesh1 = ljs.Mesh(ljs.create_device_mesh((2,)), ("m",))
    x, w = rm.synthetic((8, 32), 1), rm.synthetic((32, 12), 2)
    a = ljs.device_put(x, ljs.NamedSharding(mesh1, P(None, "m")))
    b = ljs.device_put(w, ljs.NamedSharding(mesh1, P("m", None)))
    assert len(a.addressable_shards) == 1 and tuple(a.device_buffers[0].shape) == (8, 16)
    c = ljs.lax.dot(a, b)
    ok["dot"] = torch.allclose(c.device_buffers[0], rm.dot_f32(x, w), rtol=1e-5, atol=1e-5)
    # (2) FFN tensor-parallel over 'model' (all-reduce and reduce-scatter) and FSDP over 'data'
    xh = rm.synthetic((4, 8, 32), 3)
    w1, b1 = rm.synthetic((32, 64), 4, "weight"), rm.synthetic((64,), 5, "bias")
    w2, b2 = rm.synthetic((64, 32), 6, "weight"), rm.synthetic((32,), 7, "bias")
    want = rm.ffn_forward(xh, w1, b1, w2, b2, "gelu_tanh").float()
    tp = ljs.Mesh(ljs.create_device_mesh((1, 2)), ("data", "model"))
    ns = lambda s: ljs.NamedSharding(tp, s)
    args = (ljs.device_put(w1, ns(P(None, "model"))), ljs.device_put(b1, ns(P("model"))),
            ljs.device_put(w2, ns(P("model", None))), ljs.device_put(b2, ns(P())))
    y = ljs.nn.ffn(ljs.device_put(xh, ns(P("data"))), *args)
    ok["ffn_tp"] = torch.allclose(y.device_buffers[0].float(), want, rtol=2e-2, atol=2e-2)
    y = ljs.nn.ffn(ljs.device_put(xh, ns(P("data"))), *args, scatter_dim=1)
    ok["ffn_tp_rs"] = torch.allclose(y.device_buffers[0].float(), want[:, 4 * rank:4 * rank + 4], rtol=2e-2, atol=2e-2)
    fs = ljs.Mesh(ljs.create_device_mesh((2, 1)), ("data", "model"))
    nf = lambda s: ljs.NamedSharding(fs, s)
    y = ljs.nn.ffn(ljs.device_put(xh, nf(P("data"))), ljs.device_put(w1, nf(P("data", None))), ljs.device_put(b1, nf(P())),
                   ljs.device_put(w2, nf(P("data", None))), ljs.device_put(b2, nf(P())))
    ok["ffn_fsdp"] = torch.allclose(y.device_buffers[0].float(), want[2 * rank:2 * rank + 2], rtol=2e-2, atol=2e-2)
    # (3) head-sharded attention module on a (1,2) mesh; all-to-all resharding
    model = ljs.nn.FlaxAttention(32, heads=4, dim_head=8, layout="heads")
    host = model.init_host(seed=3)
    rules = (("batch", "data"), ("heads", "model"))
    xs = ljs.NamedSharding(tp, P("data", "model"))
    with tp, ljs.nn.axis_rules(rules):
        yy = model(model.shard_params(host, tp, rules), ljs.device_put(xh, xs))
    want_a = rm.mha_forward(xh, host, 4, 8).float()
    ok["attn_heads"] = torch.allclose(yy.device_buffers[0].float(), want_a[:, 4 * rank:4 * rank + 4], rtol=2e-2, atol=2e-2)
    # (3b) the LITERAL annotation set (activations split along S over 'model'): ring attention over two processes --
    # point-to-point send / receive of the K, V blocks -- causal and not
    rules_l = (("batch", "data"), ("embed", "model"), ("hidden", "model"))
    for causal in (False, True):
        model_l = ljs.nn.FlaxAttention(32, heads=4, dim_head=8, causal=causal)
        rt.collective_log.clear()
        with tp, ljs.nn.axis_rules(rules_l):
            yl = model_l(model_l.shard_params(host, tp, rules_l), ljs.device_put(xh, xs))
        want_l = rm.mha_forward(xh, host, 4, 8, causal).float()
        ok[f"attn_literal_ring_causal{int(causal)}"] = (
            torch.allclose(yl.device_buffers[0].float(), want_l[:, 4 * rank:4 * rank + 4], rtol=2e-2, atol=2e-2)
            and "ring_shift" in [n for n, _, _ in rt.collective_log])
    t = torch.arange(4 * 8 * 6, dtype=torch.float32).reshape(4, 8, 6)
    z = ljs.reshard(ljs.device_put(t, ns(P(None, "model"))), ns(P(None, None, "model")))
    ok["all_to_all"] = torch.equal(z.device_buffers[0], t[:, :, 3 * rank:3 * rank + 3])
    ok["global"] = torch.equal(z.global_array(), t)
    # (4) train_step over the data axis (one process per device): Dense module, then the attention module (heads layout)
    xd, wd = rm.synthetic((4, 8, 32), 11), rm.synthetic((32, 16), 12, "weight")
    st = ljs.train.TrainState.create(apply_fn=None, params={"kernel": ljs.device_put(wd.clone(), nf(P("data", None)))},
                                     tx=ljs.train.adam(0.001))
    st = ljs.train.dense_train_step(st, ljs.device_put(xd, nf(P("data"))))
    k2, _, _, _ = rm.dense_train_step(xd, wd, torch.zeros_like(wd), torch.zeros_like(wd), 1)
    ok["dense_train"] = torch.allclose(st.params["kernel"].device_buffers[0], k2[16 * rank:16 * rank + 16], rtol=0, atol=2e-5)
    dp = ljs.Mesh(ljs.create_device_mesh((2, 1)), ("data", "model"))
    with dp, ljs.nn.axis_rules(rules):
        params = model.shard_params(host, dp, rules)
        grads = ljs.train.attention_sum_loss_grads(model, params, ljs.device_put(xh, ljs.NamedSharding(dp, P("data", "model"))))
    want_g = rm.mha_sum_loss_grads(xh, host, 4, 8)
    ok["attn_grads"] = all(
        torch.allclose(grads[n].device_buffers[0], want_g[n], rtol=3e-2, atol=3e-2 * want_g[n].abs().max().item())
        for n in ("to_q", "to_k", "to_v", "to_out_0.kernel", "to_out_0.bias"))
    # (5) the same train_step gradients under the LITERAL annotation set on a (1,2) mesh: S split over two processes
    model_l = ljs.nn.FlaxAttention(32, heads=4, dim_head=8)
    with tp, ljs.nn.axis_rules(rules_l):
        params_l = model_l.shard_params(host, tp, rules_l)
        grads_l = ljs.train.attention_sum_loss_grads(model_l, params_l, ljs.device_put(xh, xs))
    ok["attn_grads_literal"] = all(
        torch.allclose(grads_l[n].global_array(), want_g[n], rtol=3e-2, atol=3e-2 * want_g[n].abs().max().item())
        for n in ("to_q", "to_k", "to_v", "to_out_0.kernel", "to_out_0.bias"))
    results[rank] = ok
    rt.barrier()
    torch.distributed.destroy_process_group()


def test_world_size_2_gloo():
    world = 2
    port = 29500 + (os.getpid() % 400)
    ctx = mp.get_context("spawn")
    with ctx.Manager() as mgr:
        results = mgr.dict()
        procs = [ctx.Process(target=_worker, args=(r, world, port, results)) for r in range(world)]
        for p in procs:
            p.start()
        for p in procs:
            p.join(timeout=240)
        for p in procs:
            if p.is_alive():
                p.terminate()
                pytest.fail("gloo worker timed out")
            assert p.exitcode == 0, f"gloo worker exited with {p.exitcode}"
        for r in range(world):
            assert all(results[r].values()), f"rank {r}: {dict(results[r])}"
