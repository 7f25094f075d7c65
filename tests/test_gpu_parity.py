"""Parity tests proper (-m gpu): the CUDA path, called through the C ABI (libb200sl.so via ctypes),
against the CPU oracle on the same seeded inputs and against the committed golden fixtures.
Tolerances are the ones BASELINE.json's north_star states: fp32 rtol 1e-5; bf16 rtol 2e-2 against the
fp32-accumulate oracle tier.  Where outputs are O(1) (GEMM / Dense / FFN at small K) the rtol is paired with a flat
atol 2e-2; the attention kernels, whose outputs shrink with the number of keys, use bounds that scale with the expected
row (tests/_tol.py): |got - T1| <= 2e-2 |T1| + 4 * 2^-8 * rms_row, and |got - f64| <= |T1 - f64| + 2 bf16 ulp."""
import os

import pytest
import torch

from oracle import reference_math as rm
from tests import _tol
from tests.golden import make_golden as mg

pytestmark = pytest.mark.gpu

GOLD = torch.load(os.path.join(os.path.dirname(__file__), "golden", "oracle_outputs.pt"))
BF16_TOL = dict(rtol=2e-2, atol=2e-2)


def _dev(t, dtype=None):
    t = t.cuda()
    return t.to(dtype) if dtype is not None else t


def _check_attention(o, q, k, v, causal, what, blocked=0):
    """Both tiers of tests/_tol.py for an attention forward output (CPU tensors)."""
    if blocked:
        t1 = rm.attention_core_blocked(q, k, v, causal=causal, q_block=blocked)
        f64 = rm.attention_core_blocked(q, k, v, causal=causal, q_block=blocked, f64=True)
    else:
        t1 = rm.attention_core(q, k, v, causal=causal)
        f64 = rm.attention_core_f64(q.to(torch.bfloat16).double(), k.to(torch.bfloat16).double(), v.to(torch.bfloat16).double(),
                                    causal=causal)
    _tol.assert_close_scaled(o.float(), t1.float(), rtol=2e-2, c=4.0, what=what)
    _tol.assert_no_further_from_f64(o.float(), t1.float(), f64, ulps=2.0, what=what)


# ---------------------------------------------------------------------------- fp32 matmuls
@pytest.mark.parametrize("case", ["case1a", "case1b", "case2", "case3", "case4"])
def test_literal_sharded_dot_fp32(ljs, case):
    """The five literal layouts on 8 virtual devices backed by cuda:0; per-shard products through
    b200_gemm_f32; result equals the unsharded fp32 product at rtol 1e-5."""
    ljs.init_virtual(8, "cuda:0")
    s = ljs.PositionalSharding(ljs.create_device_mesh((2, 4)))
    layouts = {
        "case1a": (s.replicate(axis=0, keepdims=True), s.reshape(4, 2).replicate(axis=1, keepdims=True)),
        "case1b": (s.replicate(axis=0, keepdims=True), s.replicate(axis=1, keepdims=True)),
        "case2": (s, s.replicate(axis=1, keepdims=True)),
        "case3": (s, s),
        "case4": (s.replicate(axis=1, keepdims=True), s.replicate(axis=0, keepdims=True)),
    }
    A, B = mg.dot_inputs()
    sa, sb = layouts[case]
    n0 = ljs.capi.launch_count()
    C = ljs.lax.dot(ljs.device_put(A, sa), ljs.device_put(B, sb))
    assert ljs.capi.launch_count() - n0 == 8            # one C-ABI launch per virtual device
    torch.testing.assert_close(C.global_array(), GOLD["dot_4x16x4"], rtol=1e-5, atol=1e-6)
    for sh in C.addressable_shards:
        assert sh.data.is_cuda
        torch.testing.assert_close(sh.data.cpu(), GOLD["dot_4x16x4"][sh.index], rtol=1e-5, atol=1e-6)


def test_batched_einsum_fp32(ljs):
    ljs.init_virtual(1, "cuda:0")
    g = torch.Generator().manual_seed(0)
    a, b = torch.randn(8, 4, 16, generator=g), torch.randn(8, 16, 4, generator=g)
    c = ljs.lax.einsum("ABC,ACD->ABD", ljs.device_put(a), ljs.device_put(b))      # case4_gspmd_ff.py:30
    torch.testing.assert_close(c.device_buffers[0].cpu(), GOLD["einsum_8x4x16x4"], rtol=1e-5, atol=1e-6)


@pytest.mark.parametrize("shape", [(1, 1, 1, 1), (3, 5, 7, 2), (257, 129, 77, 1), (512, 384, 1024, 1), (64, 64, 4096, 2)])
def test_gemm_f32_shapes(ljs, shape):
    M, N, K, Bt = shape
    a, b = rm.synthetic((Bt, M, K), 1), rm.synthetic((Bt, K, N), 2)
    got = ljs.capi.gemm_f32(_dev(a), _dev(b)).cpu()
    want = (a.double() @ b.double())
    torch.testing.assert_close(got.double(), want, rtol=1e-5, atol=1e-5 * K ** 0.5)
    # accumulate flag adds into C
    c0 = rm.synthetic((Bt, M, N), 3)
    got2 = ljs.capi.gemm_f32(_dev(a), _dev(b), out=_dev(c0).clone(), accumulate=True).cpu()
    torch.testing.assert_close(got2.double(), want + c0.double(), rtol=1e-5, atol=1e-5 * K ** 0.5)


# ------------------------------------------------------------------------------ bf16 GEMM
@pytest.mark.parametrize("g,bn", [(0, 0), (1, 256), (1, 128), (1, 64), (2, 256), (2, 128)])
@pytest.mark.parametrize("layout", ["kn", "nk"])
def test_gemm_bf16_tile_configs(ljs, g, bn, layout):
    capi = ljs.capi
    tune = {"cta_group": g, "block_n": bn}          # per-call overrides: the library holds no tuning state
    for (M, N, K) in [(128, 256, 64), (384, 640, 192), (1000, 264, 136), (2048, 512, 640)]:
        a = rm.synthetic((M, K), 10).to(torch.bfloat16)
        w = rm.synthetic((K, N), 11, "weight").to(torch.bfloat16)
        bmat = w if layout == "kn" else w.t().contiguous()
        got = capi.gemm_bf16(_dev(a), _dev(bmat), b_layout=capi.B_KN if layout == "kn" else capi.B_NK, tune=tune).cpu()
        want = rm.dense_bf16(a, w)
        torch.testing.assert_close(got.float(), want.float(), **BF16_TOL)
        # the fp32-accumulate check: at most one bf16 ulp from the once-rounded f32 product
        exact = (a.float() @ w.float())
        ulp = exact.abs().clamp_min(2.0 ** -10) * 2.0 ** -7
        assert ((got.float() - exact).abs() <= ulp).all()


@pytest.mark.parametrize("epi", ["bias", "bias_relu", "bias_gelu_tanh", "relu", "gelu_tanh", "f32"])
def test_gemm_bf16_epilogues(ljs, epi):
    capi = ljs.capi
    M, N, K = 512, 768, 256
    a = rm.synthetic((M, K), 20).to(torch.bfloat16)
    w = rm.synthetic((K, N), 21, "weight").to(torch.bfloat16)
    b = (rm.synthetic((N,), 22, "bias") * 25).to(torch.bfloat16)
    if epi == "f32":
        got = capi.gemm_bf16(_dev(a), _dev(w), out_dtype=torch.float32).cpu()
        torch.testing.assert_close(got, a.float() @ w.float(), rtol=1e-4, atol=1e-4)
        return
    code = {"bias": capi.EPI_BIAS, "bias_relu": capi.EPI_BIAS_RELU, "bias_gelu_tanh": capi.EPI_BIAS_GELU_TANH,
            "relu": capi.EPI_RELU, "gelu_tanh": capi.EPI_GELU_TANH}[epi]
    got = capi.gemm_bf16(_dev(a), _dev(w), bias=_dev(b) if "bias" in epi else None, epilogue=code).cpu().float()
    h = rm.dense_bf16(a, w, b if "bias" in epi else None).float()
    if "relu" in epi:
        h = torch.relu(h)
    if "gelu" in epi:
        h = rm.gelu_tanh(h).to(torch.bfloat16).float()
    torch.testing.assert_close(got, h, **BF16_TOL)


def test_gemm_bf16_strided_views_and_rowblocks(ljs):
    capi = ljs.capi
    a = rm.synthetic((1024, 256), 30).to(torch.bfloat16)
    w = rm.synthetic((256, 512), 31, "weight").to(torch.bfloat16)
    ad, wd = _dev(a), _dev(w)
    # column slice of B (ldb > N) into a column slice of C (ldc > N)
    out = torch.zeros((1024, 512), dtype=torch.bfloat16, device="cuda")
    capi.gemm_bf16(ad, wd[:, 128:384], out=out[:, 128:384])
    want = rm.dense_bf16(a, w[:, 128:384])
    torch.testing.assert_close(out[:, 128:384].cpu().float(), want.float(), **BF16_TOL)
    assert out[:, :128].abs().sum().item() == 0 and out[:, 384:].abs().sum().item() == 0
    # row-blocked A: rows [128,256) of each of 4 blocks of 256 rows
    got = capi.gemm_bf16(ad[128:], wd, a_row_blocks=(512, 128, 256)).cpu()
    rows = torch.cat([a[256 * r + 128:256 * r + 256] for r in range(4)])
    torch.testing.assert_close(got.float(), rm.dense_bf16(rows, w).float(), **BF16_TOL)


def test_gemm_bf16_linearity_at_config_size(ljs):
    """BASELINE.json FFN shape (T=16384, d_model 4096 -> d_ff 16384): too big for the CPU oracle in
    seconds, so check size-independent properties: rows are independent (a sampled row block equals
    the oracle on that block) and the product is linear in A."""
    capi = ljs.capi
    T, K, N = 16384, 4096, 16384
    g = torch.Generator(device="cuda").manual_seed(1)
    a = torch.randn(T, K, device="cuda", generator=g).to(torch.bfloat16)
    w = (torch.randn(K, N, device="cuda", generator=g) / K ** 0.5).to(torch.bfloat16)
    y = capi.gemm_bf16(a, w, out_dtype=torch.float32)
    rows = torch.tensor([0, 127, 128, 5000, 8191, 8192, 16383])
    want = a[rows.cuda()].cpu().float() @ w.cpu().float()
    torch.testing.assert_close(y[rows.cuda()].cpu(), want, rtol=1e-3, atol=1e-3)
    y2 = capi.gemm_bf16((a.float() * 2).to(torch.bfloat16), w, out_dtype=torch.float32)   # exact doubling in bf16
    assert torch.equal(y2, y * 2)


@pytest.mark.parametrize("act", ["gelu_tanh", "relu"])
def test_ffn_config_size_sampled_rows(ljs, act):
    """The whole FFN at BASELINE.json configs[1] (T = 16384, d_model 4096, d_ff 16384, bias): y = act(x W1 + b1) W2 + b2
    through b200_ffn_fwd, against the oracle on 64 seeded rows (rows of an FFN are independent).  Covers GEMM2 at
    K = 16384 and the fused bias + activation epilogue at N = 16384, which the small-shape tests do not reach."""
    capi = ljs.capi
    T, d, f = 16384, 4096, 16384
    x = rm.synthetic((T, d), 0).to(torch.bfloat16)
    w1, b1 = rm.synthetic((d, f), 1, "weight").to(torch.bfloat16), rm.synthetic((f,), 2, "bias").to(torch.bfloat16)
    w2, b2 = rm.synthetic((f, d), 3, "weight").to(torch.bfloat16), rm.synthetic((d,), 4, "bias").to(torch.bfloat16)
    code = capi.EPI_BIAS_GELU_TANH if act == "gelu_tanh" else capi.EPI_BIAS_RELU
    h = torch.empty((T, f), dtype=torch.bfloat16, device="cuda")
    y = capi.ffn_fwd(_dev(x), _dev(w1), _dev(b1), _dev(w2), _dev(b2), act=code, h=h)
    rows = torch.cat([torch.tensor([0, 127, 128, 255, 256, 8191, 8192, 16383]),
                      torch.randint(0, T, (56,), generator=torch.Generator().manual_seed(7))])
    want = rm.ffn_forward(x[rows], w1, b1, w2, b2, act)
    got = y[rows.cuda()].cpu()
    _tol.assert_close_scaled(got.float(), want.float(), rtol=2e-2, c=4.0, what=f"FFN {act} config size")
    f64 = rm.ffn_forward_f64(x[rows].double(), w1.double(), b1.double(), w2.double(), b2.double(), act)
    _tol.assert_no_further_from_f64(got.float(), want.float(), f64, ulps=2.0, what=f"FFN {act} config size")
    # the intermediate too: h = act(x W1 + b1) on the same rows, bf16 (the fused epilogue at N = 16384)
    hw = rm.dense_bf16(x[rows], w1, b1).float()
    hw = torch.relu(hw) if act == "relu" else rm.gelu_tanh(hw).to(torch.bfloat16).float()
    _tol.assert_close_scaled(h[rows.cuda()].cpu().float(), hw, rtol=2e-2, c=4.0, what=f"FFN {act} hidden")


# ------------------------------------------------------------------------------------ FFN
@pytest.mark.parametrize("act", ["gelu", "relu"])
def test_ffn_against_golden(ljs, act):
    capi = ljs.capi
    x, w1, b1, w2, b2 = mg.ffn_small_inputs()
    bf = lambda t: _dev(t, torch.bfloat16)
    code = capi.EPI_BIAS_GELU_TANH if act == "gelu" else capi.EPI_BIAS_RELU
    y = capi.ffn_fwd(bf(x.reshape(-1, 256)), bf(w1), bf(b1), bf(w2), bf(b2), act=code).cpu().reshape(4, 64, 256)
    torch.testing.assert_close(y.float(), GOLD[f"ffn_{act}"].float(), **BF16_TOL)


def test_ffn_sharded_layers_virtual_mesh(ljs):
    """TP (all-reduce and reduce-scatter) and FSDP plans on 4 virtual devices backed by the GPU."""
    ljs.init_virtual(4, "cuda:0")
    P = ljs.PartitionSpec
    x, w1, b1, w2, b2 = mg.ffn_small_inputs()
    want = GOLD["ffn_gelu"].float()
    tp = ljs.Mesh(ljs.create_device_mesh((1, 4)), ("data", "model"))
    ns = lambda s: ljs.NamedSharding(tp, s)
    args = (ljs.device_put(w1, ns(P(None, "model"))), ljs.device_put(b1, ns(P("model"))),
            ljs.device_put(w2, ns(P("model", None))), ljs.device_put(b2, ns(P())))
    y = ljs.nn.ffn(ljs.device_put(x, ns(P("data"))), *args)
    torch.testing.assert_close(y.global_array().float(), want, **BF16_TOL)
    y = ljs.nn.ffn(ljs.device_put(x, ns(P("data"))), *args, scatter_dim=1)
    assert tuple(y.device_buffers[0].shape) == (4, 16, 256)
    torch.testing.assert_close(y.global_array().float(), want, **BF16_TOL)
    fs = ljs.Mesh(ljs.create_device_mesh((4, 1)), ("data", "model"))
    nf = lambda s: ljs.NamedSharding(fs, s)
    y = ljs.nn.ffn(ljs.device_put(x, nf(P("data"))), ljs.device_put(w1, nf(P("data", None))), ljs.device_put(b1, nf(P())),
                   ljs.device_put(w2, nf(P("data", None))), ljs.device_put(b2, nf(P())))
    torch.testing.assert_close(y.global_array().float(), want, **BF16_TOL)


# ------------------------------------------------------------------------------ attention
@pytest.mark.parametrize("d", [64, 128])
@pytest.mark.parametrize("causal", [False, True])
def test_attention_against_golden(ljs, d, causal):
    q, k, v = mg.attention_inputs(d)
    o = ljs.capi.attention_fwd(_dev(q), _dev(k), _dev(v), causal=causal).cpu()
    _tol.assert_close_scaled(o.float(), GOLD[f"attn_d{d}" + ("_causal" if causal else "")].float(), rtol=2e-2, c=4.0,
                             what=f"attention golden d{d}")
    _check_attention(o, q, k, v, causal, f"attention golden d{d}")


@pytest.mark.parametrize("shape", [(1, 1, 1, 1, 64), (1, 5, 300, 2, 64), (2, 128, 128, 3, 128), (1, 257, 257, 2, 128),
                                   (2, 100, 1000, 2, 64), (1, 1024, 1024, 4, 128), (1, 384, 640, 2, 128),
                                   (1, 1300, 1300, 1, 64)])
@pytest.mark.parametrize("causal", [False, True])
@pytest.mark.parametrize("variant", [0, 1, 8, 9])
def test_attention_ragged_shapes(ljs, shape, causal, variant):
    """Every kernel of the library (variant is a per-call argument: 0 = library choice, 1 = one query tile per CTA,
    8 = persistent, 9 = persistent CTA pairs) on ragged Sq / Sk, both tolerance tiers."""
    B, Sq, Sk, H, D = shape
    if variant == 9 and D != 128:
        pytest.skip("variant 9 (CTA pairs) is head dim 128 only")
    q = rm.synthetic((B, Sq, H, D), 40).to(torch.bfloat16)
    k = rm.synthetic((B, Sk, H, D), 41).to(torch.bfloat16)
    v = rm.synthetic((B, Sk, H, D), 42).to(torch.bfloat16)
    o = ljs.capi.attention_fwd(_dev(q), _dev(k), _dev(v), causal=causal, variant=variant).cpu()
    _check_attention(o, q, k, v, causal, f"attention v{variant} {shape} causal={causal}")


@pytest.mark.parametrize("variant", [8, 9])
@pytest.mark.parametrize("causal", [False, True])
@pytest.mark.parametrize("ramp_top", [24.0, 2.0, 1.0])   # every row rescales in every tile / 90-99 % of rows / alternate tiles
def test_attention_growing_row_max(ljs, variant, causal, ramp_top):
    """Logits that grow along the key axis: the running max rises by more than the lazy-rescale threshold (2^8) several
    times per row, in even and in odd KV tiles, so O is rescaled in TMEM by either softmax group."""
    B, S, H, D = 1, 1024, 2, 128
    q = rm.synthetic((B, S, H, D), 60).to(torch.bfloat16)
    k = rm.synthetic((B, S, H, D), 61).to(torch.bfloat16)
    v = rm.synthetic((B, S, H, D), 62).to(torch.bfloat16)
    ramp = torch.linspace(0.0, ramp_top, S).view(1, S, 1, 1)
    q = (q.float().abs() + 0.5).to(torch.bfloat16)                     # positive queries ...
    k = (k.float().abs() * ramp + ramp).to(torch.bfloat16)             # ... against keys whose size grows with position
    o = ljs.capi.attention_fwd(_dev(q), _dev(k), _dev(v), causal=causal, variant=variant).cpu()
    _check_attention(o, q, k, v, causal, f"attention v{variant} ramp {ramp_top} causal={causal}")


def test_attention_properties(ljs):
    capi = ljs.capi
    q = rm.synthetic((2, 512, 4, 128), 50).to(torch.bfloat16)
    k = rm.synthetic((2, 512, 4, 128), 51).to(torch.bfloat16)
    # uniform V: softmax rows sum to one, so O == V
    ones = torch.full((2, 512, 4, 128), 0.5, dtype=torch.bfloat16)
    o = capi.attention_fwd(_dev(q), _dev(k), _dev(ones)).cpu()
    torch.testing.assert_close(o.float(), ones.float(), rtol=2 ** -7, atol=0)
    # causal row 0 attends to key 0 only
    v = rm.synthetic((2, 512, 4, 128), 52).to(torch.bfloat16)
    o = capi.attention_fwd(_dev(q), _dev(k), _dev(v), causal=True).cpu()
    assert torch.equal(o[:, 0], v[:, 0])
    # large logits: the running max keeps exp in range
    o = capi.attention_fwd(_dev(q * 16), _dev(k * 16), _dev(v)).cpu()
    _check_attention(o, q * 16, k * 16, v, False, "attention large logits")
    # strided input: q/k/v as slices of one fused projection buffer [B, S, 3*H*D]
    qkv = torch.cat([q.reshape(2, 512, -1), k.reshape(2, 512, -1), v.reshape(2, 512, -1)], dim=-1).cuda()
    qs, ks, vs = (qkv[..., i * 512:(i + 1) * 512].view(2, 512, 4, 128) for i in range(3))
    o2 = capi.attention_fwd(qs, ks, vs).cpu()
    _check_attention(o2, q, k, v, False, "attention strided views")


def test_attention_rejects_bad_variant(ljs):
    q = torch.zeros((1, 256, 1, 64), dtype=torch.bfloat16, device="cuda")
    with pytest.raises(ljs.capi.B200Error, match="variant"):
        ljs.capi.attention_fwd(q, q, q, variant=5)          # the archived experiments are not in the library
    with pytest.raises(ljs.capi.B200Error, match="head dim 128"):
        ljs.capi.attention_fwd(q, q, q, variant=9)


@pytest.mark.parametrize("cfg", ["case5", "case6"])
def test_attention_config_size_sample(ljs, cfg):
    """BASELINE.json attention shapes -- case5: B8, S4096, 32 heads x 128, non-causal (and causal); case6: B16, S8192,
    32 x 128, causal, on the heads one GPU of the 2x4 mesh holds (batch 8, 8 heads).  The oracle runs on a seeded sample
    of (batch, head) pairs: heads and batch items are independent.  Both tolerance tiers."""
    capi = ljs.capi
    if cfg == "case5":
        B, S, H, D, modes, pairs = 8, 4096, 32, 128, (False, True), [(0, 0), (3, 17), (7, 31)]
    else:
        B, S, H, D, modes, pairs = 8, 8192, 8, 128, (True,), [(0, 0), (5, 3), (7, 7)]
    g = torch.Generator(device="cuda").manual_seed(3)
    q, k, v = (torch.randn(B, S, H, D, device="cuda", generator=g).to(torch.bfloat16) for _ in range(3))
    for causal in modes:
        o = capi.attention_fwd(q, k, v, causal=causal)
        for (b, h) in pairs:
            sl = (slice(b, b + 1), slice(None), slice(h, h + 1))
            _check_attention(o[sl].cpu(), q[sl].cpu(), k[sl].cpu(), v[sl].cpu(), causal,
                             f"{cfg} config size (b={b}, h={h}, causal={causal})", blocked=1024)


# ------------------------------------------------------------------- literal case5 / case6
def test_case6_literal_module_against_golden(ljs):
    """FlaxAttention on a 2x2 mesh (4 virtual devices on cuda:0) with the reference's rules and
    shapes (B8, S256, M640, 8x64) -- case6_attention.py:149-237."""
    ljs.init_virtual(4, "cuda:0")
    mesh = ljs.Mesh(ljs.create_device_mesh((2, 2)), ("data", "model"))
    x_host, host = mg.case6_literal_inputs()
    rules = (("batch", "data"), ("embed", "model"), ("hidden", "model"))
    model = ljs.nn.FlaxAttention(640, heads=8, dim_head=64)
    x_sharding = ljs.NamedSharding(mesh, ljs.PartitionSpec("data", "model"))
    params = model.shard_params(host, mesh, rules)
    assert tuple(params["to_q"].device_buffers[0].shape) == (320, 512)
    x = ljs.device_put(x_host, x_sharding)
    assert tuple(x.device_buffers[0].shape) == (4, 128, 640)
    apply_fn = ljs.nn.jit(lambda p, xx: model(p, xx), in_shardings=(None, x_sharding), out_shardings=x_sharding)
    with mesh, ljs.nn.axis_rules(rules):
        y = apply_fn(params, x)
    assert tuple(y.device_buffers[0].shape) == (4, 128, 640)
    torch.testing.assert_close(y.global_array()[::2, ::4].float(), GOLD["case6_literal_y"].float(), **BF16_TOL)
    # head-sharded annotation set gives the same values
    rules_h = (("batch", "data"), ("heads", "model"), ("hidden", "model"))
    model_h = ljs.nn.FlaxAttention(640, heads=8, dim_head=64, layout="heads")
    with mesh, ljs.nn.axis_rules(rules_h):
        yh = model_h(model_h.shard_params(host, mesh, rules_h), x)
    torch.testing.assert_close(yh.global_array()[::2, ::4].float(), GOLD["case6_literal_y"].float(), **BF16_TOL)


@pytest.mark.parametrize("causal", [False, True])
@pytest.mark.parametrize("d", [64, 128])
def test_ring_attention_virtual_mesh(ljs, causal, d):
    """Ring attention on 4 virtual devices of one GPU (S split 4-way over 'model'): b200_attention_fwd_lse per key/value
    block + b200_attention_combine, against the oracle's unsharded attention, both tolerance tiers; and the merge kernel
    alone against its formula."""
    ljs.init_virtual(4, "cuda:0")
    mesh = ljs.Mesh(ljs.create_device_mesh((1, 4)), ("data", "model"))
    sh = ljs.NamedSharding(mesh, ljs.PartitionSpec("data", "model", None, None))
    B, S, H = 2, 1024, 2
    q, k, v = (rm.synthetic((B, S, H, d), 110 + i).to(torch.bfloat16) for i in range(3))
    n0 = ljs.capi.launch_count()
    o = ljs.nn.dot_product_attention(ljs.device_put(q, sh), ljs.device_put(k, sh), ljs.device_put(v, sh), causal=causal)
    blocks = 10 if causal else 16                      # (query block, key block) pairs that are not fully masked
    assert ljs.capi.launch_count() - n0 == 2 * blocks + 4            # attention + merge per pair, one cast per device
    assert tuple(o.device_buffers[0].shape) == (B, S // 4, H, d)
    _check_attention(o.global_array(), q, k, v, causal, f"ring attention d{d} causal={causal}")


def test_attention_combine_formula(ljs):
    capi = ljs.capi
    B, S, H, D = 2, 300, 3, 128
    o1, o2 = (rm.synthetic((B, S, H, D), 120 + i).to(torch.bfloat16).cuda() for i in range(2))
    l1, l2 = (rm.synthetic((B, H, S), 122 + i).cuda() * 3 for i in range(2))
    acc = torch.full((B, S, H, D), float("nan"), device="cuda")
    lse = torch.full((B, H, S), float("nan"), device="cuda")
    capi.attention_combine(acc, lse, o1, l1, first=True)
    assert torch.equal(acc, o1.float()) and torch.equal(lse, l1)
    capi.attention_combine(acc, lse, o2, l2, first=False)
    ln = torch.log2(torch.exp2(l1.double()) + torch.exp2(l2.double()))
    w1 = torch.exp2(l1.double() - ln).permute(0, 2, 1).unsqueeze(-1)
    w2 = torch.exp2(l2.double() - ln).permute(0, 2, 1).unsqueeze(-1)
    torch.testing.assert_close(lse.double(), ln, rtol=0, atol=1e-5)
    torch.testing.assert_close(acc.double(), o1.double() * w1 + o2.double() * w2, rtol=1e-5, atol=1e-5)


def test_case5_literal_dense_against_golden(ljs):
    ljs.init_virtual(4, "cuda:0")
    mesh = ljs.Mesh(ljs.create_device_mesh((2, 2)), ("data", "model"))
    x_host, host = mg.case6_literal_inputs()
    rules = (("batch", "data"), ("embed", "data"), ("hidden", "model"))      # case5_attention_dense.py:109-112
    w = ljs.device_put(host["to_q"], ljs.nn.logical_to_mesh_sharding(("embed", "kv"), mesh, rules))
    x = ljs.device_put(x_host, ljs.NamedSharding(mesh, ljs.PartitionSpec("data", "model")))
    y = ljs.nn.dense(x, w)
    assert tuple(y.device_buffers[0].shape) == (4, 128, 512)                 # case5_attention_dense.py:146 comment
    torch.testing.assert_close(y.global_array()[::2, ::4].float(), GOLD["case6_literal_q"].float(), **BF16_TOL)


def test_casts_and_partials(ljs):
    capi = ljs.capi
    x = rm.synthetic((3, 1000, 37), 60)
    assert torch.equal(capi.cast_f32_to_bf16(_dev(x)).cpu(), x.to(torch.bfloat16))
    assert torch.equal(capi.cast_bf16_to_f32(_dev(x.to(torch.bfloat16))).cpu(), x.to(torch.bfloat16).float())
    parts = rm.synthetic((4, 64, 128), 61)
    bias = rm.synthetic((128,), 62, "bias").to(torch.bfloat16)
    got = capi.sum_partials_to_bf16(_dev(parts), bias=_dev(bias)).cpu()
    want = ((parts.sum(0).to(torch.bfloat16).float() + bias.float())).to(torch.bfloat16)
    torch.testing.assert_close(got.float(), want.float(), rtol=1e-2, atol=1e-2)


def test_gemm_row_block_scatter_and_partial_sum(ljs):
    """b200_gemm_bf16_ex's row-block scatter (the compute half of the fused reduce-scatter) with the
    'peer' slots being separate local buffers, and b200_sum_bf16_partials (the local half)."""
    capi = ljs.capi
    M, N, K, W = 1024, 512, 384, 4
    R = M // W
    a = rm.synthetic((M, K), 70).to(torch.bfloat16)
    w = rm.synthetic((K, N), 71, "weight").to(torch.bfloat16)
    want = rm.dense_bf16(a, w).float()
    slots = [torch.zeros((R, N), dtype=torch.bfloat16, device="cuda") for _ in range(W)]
    ptrs = torch.tensor([s.data_ptr() for s in slots], dtype=torch.int64, device="cuda")
    first = None
    for rot, interleave in ((0, False), (1, False), (3, False), (0, True), (2, True)):
        for s in slots:
            s.zero_()
        capi.gemm_bf16_scatter(_dev(a), _dev(w), ptrs, R, rot, N, interleave=interleave)
        torch.cuda.synchronize()
        # physical block b (rows [bR, (b+1)R) of A) lands in slot b whatever the rotation and the tile order
        for b in range(W):
            torch.testing.assert_close(slots[b].cpu().float(), want[b * R:(b + 1) * R], **BF16_TOL)
        got = torch.cat([s.cpu() for s in slots])
        first = got if first is None else first
        assert torch.equal(got, first)                   # the tile order changes no bit
    # local reduction of W bf16 partials (+ bias)
    parts = rm.synthetic((W, R, N), 72).to(torch.bfloat16)
    bias = rm.synthetic((N,), 73, "bias").to(torch.bfloat16)
    pd = _dev(parts)
    out = torch.empty((R, N), dtype=torch.bfloat16, device="cuda")
    capi.sum_bf16_partials(pd.data_ptr(), R * N, W, _dev(bias), N, out)
    ref = (parts.float().sum(0).to(torch.bfloat16).float() + bias.float()).to(torch.bfloat16)
    torch.testing.assert_close(out.cpu().float(), ref.float(), rtol=1e-2, atol=1e-2)


def test_gemm_scatter_two_level_view_and_interleave(ljs):
    """Reduce-scatter ALONG SEQ of a [batch, seq] activation (b200_gemm_args.a_blocks_inner): owner j's slot receives its
    seq block of every batch item, [batch, seq / W, N]; with and without the interleaved tile order; in two chunks over the
    batch items."""
    capi = ljs.capi
    Bn, S, N, K, W = 4, 1024, 256, 192, 4
    rpb = S // W
    a = rm.synthetic((Bn * S, K), 75).to(torch.bfloat16)
    w = rm.synthetic((K, N), 76, "weight").to(torch.bfloat16)
    want = rm.dense_bf16(a, w).float().reshape(Bn, W, rpb, N)
    ad, wd = _dev(a), _dev(w)
    for interleave in (False, True):
        slots = [torch.zeros((Bn * rpb, N), dtype=torch.bfloat16, device="cuda") for _ in range(W)]
        ptrs = torch.tensor([s.data_ptr() for s in slots], dtype=torch.int64, device="cuda")
        capi.gemm_bf16_scatter(ad, wd, ptrs, Bn * rpb, 1, N, M=Bn * S, a_view=(rpb, Bn, S, rpb), interleave=interleave)
        torch.cuda.synchronize()
        for j in range(W):
            torch.testing.assert_close(slots[j].cpu().float().reshape(Bn, rpb, N), want[:, j], **BF16_TOL)
        # two chunks of two batch items each: chunk c writes rows [c * 2 * rpb, (c + 1) * 2 * rpb) of every slot
        slots2 = [torch.zeros((Bn * rpb, N), dtype=torch.bfloat16, device="cuda") for _ in range(W)]
        for c in range(2):
            p2 = torch.tensor([s.data_ptr() + c * 2 * rpb * N * 2 for s in slots2], dtype=torch.int64, device="cuda")
            capi.gemm_bf16_scatter(ad[c * 2 * S:], wd, p2, 2 * rpb, 3, N, M=W * 2 * rpb, a_view=(rpb, 2, S, rpb),
                                   interleave=interleave)
        torch.cuda.synchronize()
        for j in range(W):
            assert torch.equal(slots2[j], slots[j])


def test_gemm_progress_signals_and_progressive_sum(ljs):
    """The pushing GEMM's per-raster-group progress flags (b200_gemm_args.done_*) and the reduction that follows them
    (b200_peer_progressive_sum), with the 'owners' being local buffers: two senders' partial products of 8192 rows (two
    raster groups), four owners, interleaved tile order; flags and counters end at their expected values, and two steps in a
    row (counters are never reset) give the sum of the partials."""
    capi = ljs.capi
    M, N, K, W, S = 8192, 512, 256, 4, 2
    R = M // W
    groups, rows_per_group = M // 4096, (16 // W) * 256
    w = rm.synthetic((K, N), 131, "weight").to(torch.bfloat16)
    bias = rm.synthetic((N,), 132, "bias").to(torch.bfloat16)
    a = [rm.synthetic((M, K), 133 + s).to(torch.bfloat16) for s in range(S)]
    flags = torch.zeros((W, S, groups), dtype=torch.int32, device="cuda")            # one flag array per owner
    flag_ptrs = torch.tensor([flags[o].data_ptr() for o in range(W)], dtype=torch.int64, device="cuda")
    counters = [torch.zeros(groups, dtype=torch.int32, device="cuda") for _ in range(S)]
    slots = torch.zeros((W, S, R, N), dtype=torch.bfloat16, device="cuda")           # owner o, sender s
    tune = {"cta_group": 2, "block_n": 256, "group_m": 16}
    for epoch in (1, 2):
        outs = [torch.full((R, N), float("nan"), dtype=torch.bfloat16, device="cuda") for _ in range(W)]
        ad, wd, bd = [_dev(t) for t in a], _dev(w), _dev(bias)
        ptrs = [torch.tensor([slots[o, s].data_ptr() for o in range(W)], dtype=torch.int64, device="cuda") for s in range(S)]
        torch.cuda.synchronize()
        main = torch.cuda.current_stream()
        # as in the product: sender 0's GEMM first, then -- behind the LAST pushing GEMM on the same stream, as its
        # programmatic dependent -- the owners' reductions, which poll the flags beside it
        for s in range(S):
            capi.gemm_bf16_scatter(ad[s], wd, ptrs[s], R, (s + 1) % W, N, interleave=True, tune=tune,
                                   progress=(counters[s], flag_ptrs, s, groups, epoch))
        for o in range(W):
            capi.peer_progressive_sum(flags[o].data_ptr(), epoch, S, groups, rows_per_group, slots[o].data_ptr(), R * N,
                                      bd, N, outs[o], 8, main)
        torch.cuda.synchronize()
        assert (flags == epoch).all()
        tiles_n = N // 256
        for s in range(S):
            assert counters[s].tolist() == [epoch * 16 * tiles_n * 2] * groups       # (tile, CTA of the pair) per group
        parts = [rm.dense_bf16(a[s], w).float() for s in range(S)]
        want = (sum(parts).to(torch.bfloat16).float() + bias.float()).to(torch.bfloat16).float()
        for o in range(W):
            torch.testing.assert_close(outs[o].cpu().float(), want[o * R:(o + 1) * R], rtol=1e-2, atol=1e-2)


@pytest.mark.parametrize("causal", [False, True])
def test_sharded_mha_layer_one_gpu(ljs, causal):
    """layers.ShardedMHA (fused qkv projection -> attention on strided views -> out projection) against the
    oracle's FlaxAttention restatement."""
    from learning_jax_sharding_b200 import layers
    B, S, M, H, D = 2, 384, 512, 4, 128
    x = rm.synthetic((B, S, M), 80)
    params = {"to_q": rm.synthetic((M, H * D), 81, "weight"), "to_k": rm.synthetic((M, H * D), 82, "weight"),
              "to_v": rm.synthetic((M, H * D), 83, "weight"),
              "to_out_0": (rm.synthetic((H * D, M), 84, "weight"), rm.synthetic((M,), 85, "bias"))}
    bf = lambda t: _dev(t, torch.bfloat16)
    layer = layers.ShardedMHA(bf(params["to_q"]), bf(params["to_k"]), bf(params["to_v"]), bf(params["to_out_0"][0]),
                              bf(params["to_out_0"][1]), H, D, B, S, causal)
    y = layer.forward(bf(x)).cpu().reshape(B, S, M)
    want = rm.mha_forward(x, params, H, D, causal)
    torch.testing.assert_close(y.float(), want.float(), **BF16_TOL)


def test_host_buffer_ffn_step_matches_oracle(ljs):
    """The end-to-end call bench.py times as `e2e` (x in pinned host memory -> chunked H2D | GEMMs | D2H -> y in
    pinned host memory) returns what the oracle computes from the same seeded tensors."""
    from learning_jax_sharding_b200 import layers
    cfg = dict(batch=2, seq=1024, d_model=512, d_ff=2048, parallelism="tp")
    bench = layers.FFNBench(cfg, "gelu_tanh", rank=0, world=1, dev=torch.device("cuda", 0))
    y_host = bench.step_e2e()
    bench.e2e_drain()
    torch.cuda.synchronize()
    x = rm.synthetic((2048, 512), 0)
    w1, b1 = rm.synthetic((512, 2048), 1, "weight"), rm.synthetic((2048,), 2, "bias")
    w2, b2 = rm.synthetic((2048, 512), 3, "weight"), rm.synthetic((512,), 4, "bias")
    want = rm.ffn_forward(x, w1, b1, w2, b2, "gelu_tanh")
    torch.testing.assert_close(y_host.float(), want.float(), **BF16_TOL)
    # and the resident-input step gives the same tensor
    y_dev = bench.step_resident()
    torch.cuda.synchronize()
    torch.testing.assert_close(y_dev.cpu().float(), y_host.float(), rtol=0, atol=0)
    # streaming: back-to-back calls alternate between two sets of staging buffers and overlap; every result is complete
    # after the drain and identical
    first = y_host.clone()
    outs = [bench.step_e2e() for _ in range(5)]
    bench.e2e_drain()
    torch.cuda.synchronize()
    assert outs[0].data_ptr() != outs[1].data_ptr() and outs[0].data_ptr() == outs[2].data_ptr()
    for o in outs[-2:]:
        assert torch.equal(o, first)


# ------------------------------------------------------------------------------ train_step of the Dense module (8f row 1)
@pytest.mark.parametrize("shape", [(64, 128, 128), (2048, 640, 512), (1000, 264, 72), (8192, 1024, 2048), (4096, 4096, 512)])
def test_weight_gradient_gemm_tn(ljs, shape):
    """dW = x^T . dy with x stored [T, K_in] (b200_gemm_bf16_ex, a_layout = B200_A_KM) against f32-accumulate math."""
    T, M, N = shape
    x = rm.synthetic((T, M), 70).to(torch.bfloat16)
    dy = rm.synthetic((T, N), 71).to(torch.bfloat16)
    want = x.float().t() @ dy.float()
    got32 = ljs.capi.gemm_bf16_tn(_dev(x), _dev(dy), out_dtype=torch.float32).cpu()
    torch.testing.assert_close(got32, want, rtol=1e-4, atol=1e-3 * T ** 0.5)
    got16 = ljs.capi.gemm_bf16_tn(_dev(x), _dev(dy)).cpu()
    torch.testing.assert_close(got16.float(), want.to(torch.bfloat16).float(), **BF16_TOL)


def test_weight_gradient_gemm_tn_rejects_bad_arguments(ljs):
    x = _dev(rm.synthetic((64, 100), 72).to(torch.bfloat16))          # M = 100 is not a multiple of 8
    dy = _dev(rm.synthetic((64, 64), 73).to(torch.bfloat16))
    with pytest.raises(ljs.capi.B200Error):
        ljs.capi.gemm_bf16_tn(x, dy)
    with pytest.raises(ljs.capi.B200Error):
        ljs.capi.gemm_bf16_tn(_dev(rm.synthetic((32, 64), 74).to(torch.bfloat16)), dy)   # contraction mismatch


@pytest.mark.parametrize("n", [8, 1000, 640 * 512, 1 << 22])
@pytest.mark.parametrize("grad_dtype", [torch.float32, torch.bfloat16])
def test_adam_step_matches_oracle(ljs, n, grad_dtype):
    p = rm.synthetic((n,), 80)
    g = rm.synthetic((n,), 81).to(grad_dtype)
    mu, nu = torch.zeros(n), torch.zeros(n)
    dp, dmu, dnu = _dev(p.clone()), _dev(mu.clone()), _dev(nu.clone())
    for step in (1, 2, 3):
        p, mu, nu = rm.adam_update(p, g, mu, nu, step)
        ljs.capi.adam_step(dp, _dev(g), dmu, dnu, step=step)
        torch.testing.assert_close(dp.cpu(), p, rtol=0, atol=3e-7)       # |p| ~ 1: a few f32 ulps of the update
        torch.testing.assert_close(dmu.cpu(), mu, rtol=2e-6, atol=1e-9)
        torch.testing.assert_close(dnu.cpu(), nu, rtol=2e-6, atol=1e-12)


@pytest.mark.parametrize("wspec", [("embed", "kv"), ("embed", "hidden")])
def test_dense_train_step_literal_case5(ljs, wspec):
    """case5_attention_dense.py at its literal size (x (8,256,640), kernel (640,512), 2x2 mesh, the script's rules):
    two train steps on four virtual devices against the unsharded oracle."""
    ljs.init_virtual(4, "cuda:0")
    mesh = ljs.Mesh(ljs.create_device_mesh((2, 2)), ("data", "model"))
    P = ljs.PartitionSpec
    rules = (("batch", "data"), ("embed", "data"), ("hidden", "model"))
    x_host, w_host = rm.synthetic((8, 256, 640), 1), rm.synthetic((640, 512), 2, "weight")
    x = ljs.device_put(x_host, ljs.NamedSharding(mesh, P("data", "model")))
    w = ljs.device_put(w_host.clone(), ljs.nn.logical_to_mesh_sharding(wspec, mesh, rules))
    state = ljs.train.TrainState.create(apply_fn=None, params={"kernel": w}, tx=ljs.train.adam(learning_rate=0.001))
    g = ljs.train.dense_sum_loss_grad(x, w)
    g_want, _ = rm.dense_sum_loss_grads(x_host, w_host)
    torch.testing.assert_close(g.global_array().cpu(), g_want, rtol=2 ** -7, atol=0)
    k, mu, nu = w_host.clone(), torch.zeros_like(w_host), torch.zeros_like(w_host)
    for step in (1, 2):
        state = ljs.train.dense_train_step(state, x)
        k, mu, nu, _ = rm.dense_train_step(x_host, k, mu, nu, step)
        torch.testing.assert_close(state.params["kernel"].global_array().cpu(), k, rtol=0, atol=2e-5)
    # forward with the updated parameters still matches the oracle's forward
    y = ljs.nn.dense(x, state.params["kernel"])
    torch.testing.assert_close(y.global_array().cpu().float(), rm.dense_bf16(x_host, k).float(), **BF16_TOL)


# ------------------------------------------------------------------------------ attention backward (8f row 1)
def _check_grads(got, want_bf16, want_f64, terms, what):
    """Both tiers of tests/_tol.py for one gradient tensor; rtol 2e-2 (north_star), absolute term scaled by the row
    (its rms, or the root-sum-square of its terms where the sum cancels)."""
    _tol.assert_close_scaled(got.float(), want_bf16.float(), rtol=2e-2, c=4.0, what=what, terms=terms)
    _tol.assert_no_further_from_f64(got.float(), want_bf16.float(), want_f64, ulps=2.0, what=what, terms=terms)


@pytest.mark.parametrize("shape", [(1, 128, 128, 1, 64), (1, 128, 128, 1, 128), (2, 256, 256, 2, 128), (1, 200, 333, 2, 64),
                                   (1, 384, 640, 2, 128), (2, 1024, 1024, 2, 128)])
@pytest.mark.parametrize("causal", [False, True])
def test_attention_backward_matches_oracle(ljs, shape, causal):
    """dQ, dK, dV (b200_attention_bwd) and the forward's lse2 (b200_attention_fwd_lse) against autodiff of the f64
    formula; cotangent and inputs seeded, shapes ragged in Sq / Sk."""
    B, Sq, Sk, H, D = shape
    q = rm.synthetic((B, Sq, H, D), 90).to(torch.bfloat16)
    k = rm.synthetic((B, Sk, H, D), 91).to(torch.bfloat16)
    v = rm.synthetic((B, Sk, H, D), 92).to(torch.bfloat16)
    do = rm.synthetic((B, Sq, H, D), 93).to(torch.bfloat16)
    dq_w, dk_w, dv_w, lse_w = rm.attention_core_grads(q, k, v, do, causal=causal)
    dq_e, dk_e, dv_e, _ = rm.attention_core_grads(q, k, v, do, causal=causal, rounded=False)
    sq, sk, sv = rm.attention_core_grad_term_scales(q, k, v, do, causal=causal)
    o, lse = ljs.capi.attention_fwd_lse(_dev(q), _dev(k), _dev(v), causal=causal)
    _check_attention(o.cpu(), q, k, v, causal, f"attention fwd_lse {shape}")
    torch.testing.assert_close(lse.cpu(), lse_w, rtol=0, atol=4e-3)      # log2 units: a 0.3 % error of the denominator
    dq, dk, dv = ljs.capi.attention_bwd(_dev(q), _dev(k), _dev(v), o, _dev(do), lse, causal=causal)
    _check_grads(dv.cpu(), dv_w, dv_e, sv, f"dV {shape} causal={causal}")
    _check_grads(dk.cpu(), dk_w, dk_e, sk, f"dK {shape} causal={causal}")
    _check_grads(dq.cpu(), dq_w, dq_e, sq, f"dQ {shape} causal={causal}")


def test_attention_backward_properties(ljs):
    """Size-independent checks at a larger shape: uniform V makes dQ and dK vanish (softmax rows sum to one, so O does not
    depend on the scores), and dV's column sums equal dO's column sums per (batch, head)."""
    B, S, H, D = 2, 2048, 4, 128
    q, k = (torch.randn(B, S, H, D, generator=torch.Generator().manual_seed(i)).to(torch.bfloat16).cuda() for i in (1, 2))
    do = torch.randn(B, S, H, D, generator=torch.Generator().manual_seed(3)).to(torch.bfloat16).cuda()
    v = torch.full((B, S, H, D), 0.5, dtype=torch.bfloat16, device="cuda")
    o, lse = ljs.capi.attention_fwd_lse(q, k, v)
    dq, dk, dv = ljs.capi.attention_bwd(q, k, v, o, do, lse)
    # with random V at this shape |dQ|, |dK| ~ 0.04-0.09 (rms); here they must vanish up to bf16 rounding of dS
    assert dq.float().abs().max().item() < 2e-3 and dk.float().abs().max().item() < 2e-3
    torch.testing.assert_close(dv.float().sum(1), do.float().sum(1), rtol=2e-2, atol=0.5)


@pytest.mark.parametrize("causal", [False, True])
def test_attention_backward_config_size_sample(ljs, causal):
    """dQ, dK, dV at a config-size sequence (S = 4096, head dim 128) against autodiff of the f64 formula on sampled
    (batch, head) pairs; both tolerance tiers."""
    B, S, H, D = 2, 4096, 4, 128
    g = torch.Generator(device="cuda").manual_seed(11)
    q, k, v, do = (torch.randn(B, S, H, D, device="cuda", generator=g).to(torch.bfloat16) for _ in range(4))
    o, lse = ljs.capi.attention_fwd_lse(q, k, v, causal=causal)
    dq, dk, dv = ljs.capi.attention_bwd(q, k, v, o, do, lse, causal=causal)
    for (b, h) in [(0, 0), (1, 3)]:
        sl = (slice(b, b + 1), slice(None), slice(h, h + 1))
        args = (q[sl].cpu(), k[sl].cpu(), v[sl].cpu(), do[sl].cpu())
        dq_w, dk_w, dv_w, _ = rm.attention_core_grads(*args, causal=causal)
        dq_e, dk_e, dv_e, _ = rm.attention_core_grads(*args, causal=causal, rounded=False)
        sq, sk, sv = rm.attention_core_grad_term_scales(*args, causal=causal)
        _check_grads(dv[sl].cpu(), dv_w, dv_e, sv, f"dV S4096 (b={b}, h={h}) causal={causal}")
        _check_grads(dk[sl].cpu(), dk_w, dk_e, sk, f"dK S4096 (b={b}, h={h}) causal={causal}")
        _check_grads(dq[sl].cpu(), dq_w, dq_e, sq, f"dQ S4096 (b={b}, h={h}) causal={causal}")


@pytest.mark.parametrize("causal", [False, True])
@pytest.mark.parametrize("layout", ["heads", "literal"])
def test_attention_module_train_step_literal_size(ljs, causal, layout):
    """train_step of the attention module (case6_attention.py:206-220) at the script's literal size (x (8,256,640), 8 heads x
    64) on four virtual devices of one GPU, in the head-sharded layout and under the script's LITERAL annotation set (S split
    over 'model', kernels split on embed): gradients in the parameters' shardings against autodiff of the unsharded f64
    formula, then one Adam step through b200_adam_step."""
    ljs.init_virtual(4, "cuda:0")
    mesh = ljs.Mesh(ljs.create_device_mesh((2, 2)), ("data", "model"))
    model = ljs.nn.FlaxAttention(640, heads=8, dim_head=64, causal=causal, layout=layout)
    host = model.init_host(seed=1)
    host["to_out_0"] = (host["to_out_0"][0], rm.synthetic((640,), 9, "bias"))
    rules = (("batch", "data"), ("heads" if layout == "heads" else "embed", "model"), ("hidden", "model"))
    x_host = rm.synthetic((8, 256, 640), 5)
    x = ljs.device_put(x_host, ljs.NamedSharding(mesh, ljs.PartitionSpec("data", "model")))
    params = model.shard_params(host, mesh, rules)
    want = rm.mha_sum_loss_grads(x_host, host, 8, 64, causal)
    with mesh, ljs.nn.axis_rules(rules):
        grads = ljs.train.attention_sum_loss_grads(model, params, x)
        for name, g in grads.items():
            scale = want[name].abs().max().item()
            torch.testing.assert_close(g.global_array().cpu(), want[name], rtol=3e-2, atol=3e-2 * scale,
                                       msg=lambda m, n=name: n + ": " + m)
        state = ljs.train.attention_train_state(model, params, ljs.train.adam(learning_rate=0.001))
        state = ljs.train.attention_train_step(model, state, x)
        y = model(ljs.train.attention_params_of(state), x)                    # the updated parameters still run forward
    assert state.step == 1 and y.shape == (8, 256, 640)
    moved = (state.params["to_q"].global_array().cpu() - host["to_q"]).abs()
    assert 0.5e-3 < moved.mean().item() <= 1.01e-3                            # first Adam step: |update| = lr almost everywhere
