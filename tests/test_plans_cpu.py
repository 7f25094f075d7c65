"""Host-layer plans on a GPU-less box: virtual devices on cpu, per-shard compute routed to the
oracle-backed checker (tests/_cpu_checker.py).  What is tested here is the PLUMBING: which slices
move where, which groups reduce, and that every sharded plan reproduces the unsharded result
(sharding invariance, SURVEY.md §8c tier T2)."""
import pytest
import torch

from oracle import reference_math as rm


def _ab():
    g = torch.Generator().manual_seed(0)
    return torch.randn(4, 16, generator=g), torch.randn(16, 4, generator=g)


def _layouts(s):
    return {
        "case1a": (s.replicate(axis=0, keepdims=True), s.reshape(4, 2).replicate(axis=1, keepdims=True)),
        "case1b": (s.replicate(axis=0, keepdims=True), s.replicate(axis=1, keepdims=True)),
        "case2": (s, s.replicate(axis=1, keepdims=True)),
        "case3": (s, s),
        "case4": (s.replicate(axis=1, keepdims=True), s.replicate(axis=0, keepdims=True)),
    }


EXPECTED = {"case1a": ("partial_allreduce", "all_reduce"), "case1b": ("partial_allreduce", "all_reduce"),
            "case2": ("partial_allreduce", "all_reduce"), "case3": ("gather", "exchange"), "case4": ("local", None)}


@pytest.mark.parametrize("case", ["case1a", "case1b", "case2", "case3", "case4"])
def test_sharded_dot_equals_unsharded_product(ljs, cpu_checker, case):
    rt = ljs.init_virtual(8, "cpu")
    s = ljs.PositionalSharding(ljs.create_device_mesh((2, 4)))
    A, B = _ab()
    sa, sb = _layouts(s)[case]
    a, b = ljs.device_put(A, sa), ljs.device_put(B, sb)
    strategy, collective = EXPECTED[case]
    assert ljs.lax.plan_dot(a, b)["strategy"] == strategy
    rt.collective_log.clear()
    c = ljs.lax.dot(a, b)
    names = [n for n, _, _ in rt.collective_log]
    want = rm.dot_f32(A, B)
    torch.testing.assert_close(c.global_array(), want, rtol=1e-5, atol=1e-6)
    # every device's shard is the matching block of the unsharded product
    for sh in c.addressable_shards:
        torch.testing.assert_close(sh.data, want[sh.index], rtol=1e-5, atol=1e-6)
    if collective is None:
        assert names == []
    else:
        assert collective in names


def test_1d_mesh_of_8_matmul_allreduce(ljs, cpu_checker):
    """BASELINE.json configs[0]: x@W on a 1-D mesh of 8, contraction dim sharded on both."""
    rt = ljs.init_virtual(8, "cpu")
    mesh = ljs.Mesh(ljs.create_device_mesh((8,)), ("m",))
    P = ljs.PartitionSpec
    x, w = rm.synthetic((32, 64), 1), rm.synthetic((64, 24), 2)
    a = ljs.device_put(x, ljs.NamedSharding(mesh, P(None, "m")))
    b = ljs.device_put(w, ljs.NamedSharding(mesh, P("m", None)))
    plan = ljs.lax.plan_dot(a, b)
    assert plan["strategy"] == "partial_allreduce" and plan["reduce_groups"] == [list(range(8))]
    rt.collective_log.clear()
    c = ljs.lax.dot(a, b)
    assert [n for n, _, _ in rt.collective_log] == ["all_reduce"]      # aligned K blocks: no exchange
    torch.testing.assert_close(c.global_array(), rm.dot_f32(x, w), rtol=1e-5, atol=1e-5)


def _attention_setup(ljs, layout, mesh_shape=(2, 2), B=4, S=32, M=64, heads=4, dim_head=16, causal=False):
    ljs.init_virtual(mesh_shape[0] * mesh_shape[1], "cpu")
    mesh = ljs.Mesh(ljs.create_device_mesh(mesh_shape), ("data", "model"))
    model = ljs.nn.FlaxAttention(M, heads=heads, dim_head=dim_head, causal=causal, layout=layout)
    host = model.init_host(seed=1)
    host["to_out_0"] = (host["to_out_0"][0], rm.synthetic((M,), 9, "bias"))
    if layout == "literal":
        rules = (("batch", "data"), ("embed", "model"), ("hidden", "model"))    # case6_attention.py:183-187
    else:
        rules = (("batch", "data"), ("heads", "model"), ("hidden", "model"))    # SURVEY.md §0.1 modified set
    x_host = rm.synthetic((B, S, M), 5)
    x_sharding = ljs.NamedSharding(mesh, ljs.PartitionSpec("data", "model"))   # case6_attention.py:161
    return mesh, model, host, rules, x_host, x_sharding


@pytest.mark.parametrize("layout,causal", [("literal", False), ("heads", False), ("heads", True), ("literal", True)])
def test_attention_module_matches_oracle(ljs, cpu_checker, layout, causal):
    mesh, model, host, rules, x_host, x_sharding = _attention_setup(ljs, layout, causal=causal)
    rt = ljs.get_runtime()
    params = model.shard_params(host, mesh, rules)
    x = ljs.device_put(x_host, x_sharding)
    apply_fn = ljs.nn.jit(lambda p, xx: model(p, xx), in_shardings=(None, x_sharding), out_shardings=x_sharding)
    rt.collective_log.clear()
    with mesh, ljs.nn.axis_rules(rules):
        y = apply_fn(params, x)                                              # case6_attention.py:234-237
    names = [n for n, _, _ in rt.collective_log]
    want = rm.mha_forward(x_host, host, model.heads, model.dim_head, causal)
    assert y.shape == (4, 32, 64)
    assert tuple(y.device_buffers[0].shape) == (2, 16, 64)                   # out_shardings = x_sharding
    got = y.global_array().float()
    torch.testing.assert_close(got, want.float(), rtol=2e-2, atol=2e-2)
    if layout == "literal":
        # weights gathered, K / V blocks passed round the ring of the S-sharding group (never gathered), M->S re-split
        assert "all_gather" in names and "all_to_all" in names and "ring_shift" in names
    else:
        assert "reduce_scatter" in names                                     # row-parallel out-projection


def test_case5_fsdp_dense(ljs, cpu_checker):
    """case5_attention_dense.py: one bf16 Dense whose kernel is P('data', None) (FSDP-style)."""
    rt = ljs.init_virtual(4, "cpu")
    mesh = ljs.Mesh(ljs.create_device_mesh((2, 2)), ("data", "model"))
    P = ljs.PartitionSpec
    rules = (("batch", "data"), ("embed", "data"), ("hidden", "model"))      # case5_attention_dense.py:109-112
    x_host, w_host = rm.synthetic((4, 32, 64), 1), rm.synthetic((64, 48), 2, "weight")
    x = ljs.device_put(x_host, ljs.NamedSharding(mesh, P("data", "model")))
    w = ljs.device_put(w_host, ljs.nn.logical_to_mesh_sharding(("embed", "kv"), mesh, rules))
    assert tuple(w.device_buffers[0].shape) == (32, 48)
    rt.collective_log.clear()
    y = ljs.nn.dense(x, w)
    assert [n for n, _, _ in rt.collective_log] == ["all_gather"]           # weight gathered over 'data'
    assert tuple(y.device_buffers[0].shape) == (2, 16, 48)
    torch.testing.assert_close(y.global_array().float(), rm.dense_bf16(x_host, w_host).float(), rtol=2e-2, atol=2e-2)


@pytest.mark.parametrize("mode", ["tp_allreduce", "tp_reduce_scatter", "fsdp"])
@pytest.mark.parametrize("act", ["gelu_tanh", "relu"])
def test_ffn_plans(ljs, cpu_checker, mode, act):
    rt = ljs.init_virtual(4, "cpu")
    P = ljs.PartitionSpec
    T_b, T_s, dm, dff = 4, 16, 32, 64
    xh = rm.synthetic((T_b, T_s, dm), 1)
    w1, b1 = rm.synthetic((dm, dff), 2, "weight"), rm.synthetic((dff,), 3, "bias")
    w2, b2 = rm.synthetic((dff, dm), 4, "weight"), rm.synthetic((dm,), 5, "bias")
    if mode == "fsdp":
        mesh = ljs.Mesh(ljs.create_device_mesh((4, 1)), ("data", "model"))
        xs, w1s, w2s, b1s = P("data"), P("data", None), P("data", None), P()
    else:
        mesh = ljs.Mesh(ljs.create_device_mesh((1, 4)), ("data", "model"))
        xs, w1s, w2s, b1s = P("data"), P(None, "model"), P("model", None), P("model")
    ns = lambda s: ljs.NamedSharding(mesh, s)
    x = ljs.device_put(xh, ns(xs))
    args = (ljs.device_put(w1, ns(w1s)), ljs.device_put(b1, ns(b1s)), ljs.device_put(w2, ns(w2s)),
            ljs.device_put(b2, ns(P())))
    rt.collective_log.clear()
    y = ljs.nn.ffn(x, *args, act=act, scatter_dim=1 if mode == "tp_reduce_scatter" else None)
    names = [n for n, _, _ in rt.collective_log]
    assert names == {"tp_allreduce": ["all_reduce"], "tp_reduce_scatter": ["reduce_scatter"],
                     "fsdp": ["all_gather", "all_gather"]}[mode]
    if mode == "tp_reduce_scatter":
        assert tuple(y.device_buffers[0].shape) == (4, 4, 32)
    want = rm.ffn_forward(xh, w1, b1, w2, b2, act)
    torch.testing.assert_close(y.global_array().float(), want.float(), rtol=2e-2, atol=2e-2)


def test_reshard_patterns(ljs):
    rt = ljs.init_virtual(8, "cpu")
    mesh = ljs.Mesh(ljs.create_device_mesh((2, 4)), ("data", "model"))
    P = ljs.PartitionSpec
    t = torch.arange(8 * 16 * 4, dtype=torch.float32).reshape(8, 16, 4)
    x = ljs.device_put(t, ljs.NamedSharding(mesh, P("data", "model")))
    for spec, expect in [(P("data"), "all_gather"), (P("data", None, "model"), "all_to_all"),
                         (P(None, "model"), "all_gather"), (P("model", "data"), "exchange"), (P("data", "model"), None)]:
        rt.collective_log.clear()
        y = ljs.reshard(x, ljs.NamedSharding(mesh, spec))
        names = [n for n, _, _ in rt.collective_log]
        assert torch.equal(y.global_array(), t)
        for sh in y.addressable_shards:
            assert torch.equal(sh.data, t[sh.index])
        assert names == ([expect] if expect else [])
    # replicated -> sharded is local slicing
    r = ljs.device_put(t, ljs.NamedSharding(mesh, None))
    rt.collective_log.clear()
    y = ljs.reshard(r, ljs.NamedSharding(mesh, P("data", "model")))
    assert rt.collective_log == [] and tuple(y.device_buffers[0].shape) == (4, 4, 4)


def test_product_path_refuses_cpu_tensors(ljs):
    """No CPU fallback: without the test-only checker, per-shard compute on host buffers raises."""
    ljs.init_virtual(8, "cpu")
    s = ljs.PositionalSharding(ljs.create_device_mesh((2, 4)))
    A, B = _ab()
    a = ljs.device_put(A, s.replicate(axis=1, keepdims=True))
    b = ljs.device_put(B, s.replicate(axis=0, keepdims=True))
    with pytest.raises(ljs.capi.B200Error):
        ljs.lax.dot(a, b)


@pytest.mark.parametrize("causal", [False, True])
@pytest.mark.parametrize("mesh_shape", [(1, 4), (2, 2), (2, 4)])
def test_ring_attention_equals_gathered_attention(ljs, cpu_checker, causal, mesh_shape):
    """q, k, v split along S over 'model' (case6_attention.py:114-116): the ring plan (blocks travel, per-block results
    merged through their log-sum-exp) gives what gathering K and V along S gives, moves W - 1 blocks per device instead of
    gathering, and skips the blocks a causal mask hides completely."""
    n = mesh_shape[0] * mesh_shape[1]
    ljs.init_virtual(n, "cpu")
    rt = ljs.get_runtime()
    mesh = ljs.Mesh(ljs.create_device_mesh(mesh_shape), ("data", "model"))
    sh = ljs.NamedSharding(mesh, ljs.PartitionSpec("data", "model", None, None))
    q, k, v = (rm.synthetic((4, 64, 2, 16), 70 + i).to(torch.bfloat16) for i in range(3))
    qs, ks, vs = (ljs.device_put(t, sh) for t in (q, k, v))
    rt.collective_log.clear()
    ring = ljs.nn.dot_product_attention(qs, ks, vs, causal=causal)                     # default for this layout
    names = [nm for nm, _, _ in rt.collective_log]
    assert names.count("ring_shift") == mesh_shape[1] - 1 and "all_gather" not in names
    rt.collective_log.clear()
    gathered = ljs.nn.dot_product_attention(qs, ks, vs, causal=causal, ring=False)
    assert "all_gather" in [nm for nm, _, _ in rt.collective_log]
    want = rm.attention_core(q, k, v, causal=causal).float()
    torch.testing.assert_close(gathered.global_array().float(), want, rtol=2e-2, atol=2e-2)
    torch.testing.assert_close(ring.global_array().float(), want, rtol=2e-2, atol=2e-2)
    assert ring.sharding.spec == qs.sharding.spec
    with pytest.raises(ValueError, match="ring attention needs"):
        ljs.nn.dot_product_attention(ljs.device_put(q, ljs.NamedSharding(mesh, ljs.PartitionSpec("data"))), ks, vs, ring=True)
