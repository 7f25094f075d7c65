"""b200_ffn_fwd_static_scratch: the FFN entry point for integrators who cannot allocate device memory
around the call (a handler under a framework that already owns nearly all of HBM, or a binding
whose operand list has no scratch slot).  It must give what b200_ffn_fwd gives with a caller
workspace, call after call, although every call reuses the same built-in 32 MiB array.

Collected last on purpose (file name): a failure here must not hide any other test under -x."""
import ctypes

import pytest
import torch

pytestmark = pytest.mark.gpu

SCRATCH_BYTES = 32 << 20


def _bind(lib):
    vp, i32 = ctypes.c_void_p, ctypes.c_int
    lib.b200_ffn_fwd_static_scratch.argtypes = [vp, vp, vp, vp, vp, vp, i32, i32, i32, i32, i32, vp]
    lib.b200_ffn_fwd_static_scratch.restype = i32
    return lib.b200_ffn_fwd_static_scratch


def _inputs(T, d_model, F, seed):
    g = torch.Generator().manual_seed(seed)
    bf = lambda t: t.to(torch.bfloat16).cuda()
    x = bf(torch.randn(T, d_model, generator=g))
    w1 = bf(torch.randn(d_model, F, generator=g) / d_model ** 0.5)
    b1 = bf(torch.randn(F, generator=g) * 0.02)
    w2 = bf(torch.randn(F, d_model, generator=g) / F ** 0.5)
    b2 = bf(torch.randn(d_model, generator=g) * 0.02)
    return x, w1, b1, w2, b2


def test_static_scratch_ffn_matches_workspace_ffn_twice_in_a_row(ljs):
    from oracle import reference_math as rm
    capi = ljs.capi
    fn = _bind(capi.load())
    stream = ctypes.c_void_p(torch.cuda.current_stream().cuda_stream)
    # two calls back to back with different shapes and activations: the second one finds the first
    # one's h still sitting in the shared array and must not be affected by it
    for call, (T, d_model, F, act, seed) in enumerate([(512, 256, 1024, capi.EPI_BIAS_GELU_TANH, 1),
                                                       (384, 512, 2048, capi.EPI_BIAS_RELU, 2)]):
        x, w1, b1, w2, b2 = _inputs(T, d_model, F, seed)
        want = capi.ffn_fwd(x, w1, b1, w2, b2, act=act)                 # existing entry point, caller workspace
        got = torch.full((T, d_model), float("nan"), dtype=torch.bfloat16, device="cuda")
        torch.cuda.synchronize()
        held = torch.cuda.memory_allocated()
        free0 = torch.cuda.mem_get_info()[0]
        rc = fn(x.data_ptr(), w1.data_ptr(), b1.data_ptr(), w2.data_ptr(), b2.data_ptr(), got.data_ptr(),
                capi.BF16, T, d_model, F, act, stream)
        assert rc == 0, capi.load().b200_last_error().decode()
        torch.cuda.synchronize()
        # nothing was allocated on the way: neither through torch's allocator nor behind its back
        assert torch.cuda.memory_allocated() == held
        if call > 0:      # (the first call on a device is the documented set-up call: the driver may load the library's
            assert torch.cuda.mem_get_info()[0] >= free0      #  image, built-in array included, at that point)
        # same two launches on the same inputs as b200_ffn_fwd: the same bits, not merely close
        assert torch.equal(got, want)
        # and both agree with the CPU oracle of the layer
        cpu = lambda t: t.cpu().float()
        ref = rm.ffn_forward(cpu(x), cpu(w1), cpu(b1), cpu(w2), cpu(b2), "gelu_tanh" if act == capi.EPI_BIAS_GELU_TANH else "relu")
        torch.testing.assert_close(got.cpu().float(), ref.float(), rtol=2e-2, atol=2e-2)


def test_static_scratch_ffn_size_limit_is_inclusive(ljs):
    """A request whose intermediate is exactly 32 MiB runs; one 8-column step beyond it is refused before any pointer is
    touched, with both sizes in the message."""
    capi = ljs.capi
    fn = _bind(capi.load())
    T, d_model, F = 1024, 128, 16384
    assert 2 * T * F == SCRATCH_BYTES
    x, w1, b1, w2, b2 = _inputs(T, d_model, F, 4)
    y = torch.empty((T, d_model), dtype=torch.bfloat16, device="cuda")
    stream = ctypes.c_void_p(torch.cuda.current_stream().cuda_stream)
    rc = fn(x.data_ptr(), w1.data_ptr(), b1.data_ptr(), w2.data_ptr(), b2.data_ptr(), y.data_ptr(), capi.BF16, T, d_model,
            F, capi.EPI_BIAS_GELU_TANH, stream)
    assert rc == 0, capi.load().b200_last_error().decode()
    torch.cuda.synchronize()
    assert torch.equal(y, capi.ffn_fwd(x, w1, b1, w2, b2, act=capi.EPI_BIAS_GELU_TANH))
    rc = fn(x.data_ptr(), w1.data_ptr(), b1.data_ptr(), w2.data_ptr(), b2.data_ptr(), y.data_ptr(), capi.BF16, T, d_model,
            F + 8, capi.EPI_BIAS_GELU_TANH, stream)
    assert rc == 1
    msg = capi.load().b200_last_error().decode()
    assert "built-in scratch" in msg and str(2 * T * (F + 8)) in msg and str(SCRATCH_BYTES) in msg


def test_static_scratch_ffn_rejects_oversized_requests(ljs):
    capi = ljs.capi
    fn = _bind(capi.load())
    T, d_model, F = 2048, 256, 16384                                   # h would need 64 MiB > 32 MiB
    assert 2 * T * F > SCRATCH_BYTES
    x, w1, b1, w2, b2 = _inputs(8, d_model, 8, 3)                      # never dereferenced: rejected up front
    y = torch.empty((8, d_model), dtype=torch.bfloat16, device="cuda")
    rc = fn(x.data_ptr(), w1.data_ptr(), b1.data_ptr(), w2.data_ptr(), b2.data_ptr(), y.data_ptr(), capi.BF16, T,
            d_model, F, capi.EPI_BIAS_GELU_TANH, None)
    assert rc == 1
    msg = capi.load().b200_last_error().decode()
    assert "built-in scratch" in msg and str(2 * T * F) in msg
