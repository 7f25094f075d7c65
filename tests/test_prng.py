"""jax.random restated (oracle/jax_prng.py), the host-side key arithmetic (learning_jax_sharding_b200.random) and -- on
a GPU -- the born-sharded generator kernel (b200_random_fill).

This is the one place where parity is anchored on values that are NOT this repo's own output: the threefry2x32
known-answer vectors of the Random123 distribution (the ones jax's tests/random_test.py::testThreefry2x32 asserts), and
values printed in jax's public documentation for PRNGKey(0): random.split, random.uniform, random.normal."""
import numpy as np
import pytest
import torch

from oracle import jax_prng as jp

KAT = [((0x0, 0x0), (0x0, 0x0), (0x6b200159, 0x99ba4efe)),
       ((0xffffffff, 0xffffffff), (0xffffffff, 0xffffffff), (0x1cb996fc, 0xbb002be7)),
       ((0x13198a2e, 0x03707344), (0x243f6a88, 0x85a308d3), (0xc4923a9c, 0x483df7a0))]


def test_oracle_threefry_known_answers():
    for key, ctr, want in KAT:
        y0, y1 = jp.threefry2x32(np.array(key, dtype=np.uint32), np.array([ctr[0]], dtype=np.uint32),
                                 np.array([ctr[1]], dtype=np.uint32))
        assert (int(y0[0]), int(y1[0])) == want


def test_oracle_matches_documented_jax_outputs():
    key = jp.PRNGKey(0)
    assert jp.split(key).tolist() == [[4146024105, 967050713], [2718843009, 1272950319]]      # jax.random.split docs
    assert abs(float(jp.uniform(key, ())) - 0.41845703) < 5e-9                               # random.uniform(key)
    assert abs(float(jp.normal(key, (1,))[0]) - (-0.20584226)) < 2e-8                         # random.normal(key, (1,))
    np.testing.assert_allclose(jp.normal(key, (3,)), [1.8160863, -0.48262316, 0.33988908], rtol=0, atol=2e-7)
    assert abs(float(jp.normal(jp.PRNGKey(42), ())) - (-0.18471177)) < 2e-8                   # jax's PRNG tutorial


def test_oracle_same_key_gives_reshaped_draws():
    """case1a.py:16-18 draws A (4,16) and B (16,4) from the SAME key: B is A's stream in another shape."""
    a, b = jp.normal(jp.PRNGKey(0), (4, 16)), jp.normal(jp.PRNGKey(0), (16, 4))
    assert np.array_equal(a.reshape(-1), b.reshape(-1))


def test_oracle_lecun_normal_statistics():
    w = jp.lecun_normal(jp.PRNGKey(1), (640, 512))
    assert abs(float(w.std()) * 640 ** 0.5 - 1.0) < 0.01                      # variance 1 / fan_in
    assert float(np.abs(w).max()) * 640 ** 0.5 * jp.LECUN_TRUNC_STD < 2.0     # truncated at two (unscaled) sigmas
    assert abs(float(w.mean())) * 640 ** 0.5 < 0.01


def test_host_key_arithmetic_matches_oracle(ljs):
    r = ljs.random
    for key, ctr, want in KAT:
        assert r._threefry(key, *ctr) == want
    assert r.split(r.PRNGKey(0)) == [(4146024105, 967050713), (2718843009, 1272950319)]
    for seed in (0, 1, 7, 2 ** 31 + 5):
        for num in (1, 2, 3, 5):
            assert [list(k) for k in r.split(r.PRNGKey(seed), num)] == jp.split(jp.PRNGKey(seed), num).tolist()
        assert list(r.fold_in(r.PRNGKey(seed), 99)) == jp.fold_in(jp.PRNGKey(seed), 99).tolist()
        assert list(r.path_key(r.PRNGKey(seed), "to_q/kernel")) == jp.path_key(jp.PRNGKey(seed), "to_q/kernel").tolist()


def test_state_sharding_derivation_literal_rules(ljs):
    """get_partition_spec -> logical_to_mesh_sharding over the parameter tree with the script's rules
    (case6_attention.py:183-191): q/k/v kernels P('model', None), to_out_0 kernel P(None, 'model'), bias replicated."""
    ljs.init_virtual(4, "cpu")
    mesh = ljs.Mesh(ljs.create_device_mesh((2, 2)), ("data", "model"))
    rules = (("batch", "data"), ("embed", "model"), ("hidden", "model"))
    model = ljs.nn.FlaxAttention(640, heads=8, dim_head=64)
    sh = ljs.nn.logical_to_mesh_sharding(ljs.nn.get_partition_spec(model), mesh, rules)
    assert sh["to_q"].spec == ljs.PartitionSpec("model", None)
    assert sh["to_out_0"][0].spec == ljs.PartitionSpec(None, "model")
    assert sh["to_out_0"][1].spec == ljs.PartitionSpec()
    idx = sh["to_q"].devices_indices_map((640, 512))
    assert {(s[0].start, s[0].stop) for s in idx.values()} == {(0, 320), (320, 640)}      # case6_attention.py:200-204


# ------------------------------------------------------------------------------------------------ GPU: the kernel
def _ulps(a, b):
    a, b = np.asarray(a, dtype=np.float32), np.asarray(b, dtype=np.float32)
    return np.abs(a.view(np.int32).astype(np.int64) - b.view(np.int32).astype(np.int64))


@pytest.mark.gpu
@pytest.mark.parametrize("shape", [(1,), (3,), (4, 16), (16, 4), (7, 5, 3), (640, 512), (2, 3, 4, 5), (1000003,)])
def test_kernel_bits_are_jax_bits(ljs, shape):
    """Integer stage, bit-exact: even and odd sizes (the odd ones exercise the zero pad word of the count vector)."""
    ljs.init_virtual(1, "cuda:0")
    r = ljs.random
    for seed in (0, 42):
        got = r.bits(r.PRNGKey(seed), shape).device_buffers[0].cpu().numpy().view(np.uint32)
        want = jp.random_bits(jp.PRNGKey(seed), int(np.prod(shape))).reshape(shape)
        assert np.array_equal(got, want)


@pytest.mark.gpu
def test_kernel_known_answers(ljs):
    ljs.init_virtual(1, "cuda:0")
    r = ljs.random
    key = r.PRNGKey(0)
    # one word: the count vector [0] is padded to [0, 0], so the block is threefry(key = (0, 0), counter = (0, 0))
    assert r.bits(key, (1,)).device_buffers[0].cpu().numpy().view(np.uint32).tolist() == [0x6b200159]
    # four words: split(PRNGKey(0)) as documented
    assert r.bits(key, (4,)).device_buffers[0].cpu().numpy().view(np.uint32).tolist() == [4146024105, 967050713, 2718843009,
                                                                                         1272950319]
    assert abs(float(r.uniform(key, (1,)).device_buffers[0].cpu()) - 0.41845703) < 5e-9
    assert abs(float(r.normal(key, (1,)).device_buffers[0].cpu()) - (-0.20584226)) < 2e-8
    np.testing.assert_allclose(r.normal(key, (3,)).device_buffers[0].cpu().numpy(), [1.8160863, -0.48262316, 0.33988908],
                               rtol=0, atol=2e-7)
    # case1a.py:16-18: A and B from the same key
    a, b = r.normal(key, (4, 16)).device_buffers[0], r.normal(key, (16, 4)).device_buffers[0]
    assert torch.equal(a.reshape(-1), b.reshape(-1))


@pytest.mark.gpu
def test_kernel_float_stages_match_oracle(ljs):
    ljs.init_virtual(1, "cuda:0")
    r = ljs.random
    k = r.PRNGKey(5)
    n = 1 << 20
    got = r.uniform(k, (n,), -3.0, 5.0).device_buffers[0].cpu().numpy()
    assert np.array_equal(got, jp.uniform(jp.PRNGKey(5), (n,), -3.0, 5.0))             # mul + add of exact operands
    got = r.normal(k, (n,)).device_buffers[0].cpu().numpy()
    want = jp.normal(jp.PRNGKey(5), (n,))
    assert _ulps(got, want).max() <= 4 and (_ulps(got, want) == 0).mean() > 0.9
    assert abs(got.mean()) < 5e-3 and abs(got.std() - 1.0) < 5e-3
    got = r.lecun_normal(k, (640, 512)).device_buffers[0].cpu().numpy()
    want = jp.lecun_normal(jp.PRNGKey(5), (640, 512))
    assert _ulps(got, want).max() <= 4


@pytest.mark.gpu
def test_parameters_are_born_sharded(ljs):
    """model.init under out_shardings (case6_attention.py:171-196) on a 2x2 mesh of virtual devices: every device's shard
    equals the same slice of the unsharded draw -- literal rules and the head-sharded rules -- and nothing was generated
    on the host: one kernel launch per (parameter, device)."""
    ljs.init_virtual(4, "cuda:0")
    mesh = ljs.Mesh(ljs.create_device_mesh((2, 2)), ("data", "model"))
    model = ljs.nn.FlaxAttention(640, heads=8, dim_head=64)
    key = ljs.random.PRNGKey(1)
    for rules in ((("batch", "data"), ("embed", "model"), ("hidden", "model")),
                  (("batch", "data"), ("heads", "model"), ("hidden", "model"))):
        sh = ljs.nn.logical_to_mesh_sharding(ljs.nn.get_partition_spec(model), mesh, rules)
        n0 = ljs.capi.launch_count()
        params = model.init(key, sh)
        assert ljs.capi.launch_count() - n0 == 4 * 4
        for name in ("to_q", "to_k", "to_v", "to_out_0"):
            p = params[name][0] if name == "to_out_0" else params[name]
            want = jp.lecun_normal(jp.path_key(jp.PRNGKey(1), name + "/kernel"), p.shape)
            for s in p.addressable_shards:
                assert s.data.is_cuda and _ulps(s.data.cpu().numpy(), want[s.index]).max() <= 4
        assert float(params["to_out_0"][1].global_array().abs().sum()) == 0.0
    assert tuple(params["to_q"].device_buffers[0].shape) == (640, 256)      # heads rules: columns over 'model'
    # the module runs on the parameters it was born with
    x = ljs.device_put(torch.randn(8, 256, 640, generator=torch.Generator().manual_seed(0)),
                       ljs.NamedSharding(mesh, ljs.PartitionSpec("data", "model")))
    model_h = ljs.nn.FlaxAttention(640, heads=8, dim_head=64, layout="heads")
    with mesh, ljs.nn.axis_rules(rules):
        y = model_h(params, x)
    assert y.shape == (8, 256, 640) and torch.isfinite(y.global_array().float()).all()


@pytest.mark.gpu
def test_random_fill_rejects_bad_arguments(ljs):
    out = torch.empty((4, 4), dtype=torch.float32, device="cuda")
    with pytest.raises(ljs.capi.B200Error, match="does not lie inside"):
        ljs.capi.random_fill(out, (0, 0), (4, 6), (0, 4), ljs.capi.RNG_NORMAL)
    with pytest.raises(ljs.capi.B200Error):
        ljs.capi.random_fill(out, (0, 0), (4, 4), (0, 0), ljs.capi.RNG_BITS)       # bits are written as 32-bit integers
