// XLA typed-FFI handlers over the C ABI ("each of (1)-(4) additionally gets an XLA-FFI-typed
// wrapper behind #if __has_include").  jaxlib's headers (xla/ffi/api/ffi.h) are NOT in this image, so in this build the
// translation unit is empty; with them on the include path (a maintainer adds -I$(python -c "import jaxlib, os;
// print(os.path.join(os.path.dirname(jaxlib.__file__), 'include'))") to B200_EXTRA_NVCC_FLAGS) the library also exports
//   B200Dense, B200DotF32, B200Attention, B200Ffn, B200FfnTp, B200AllReduce, B200ReduceScatter, B200AllGather, B200AllToAll
// for jax.ffi.register_ffi_target(name, jax.ffi.pycapsule(lib.<Symbol>), platform="CUDA") -- INTEGRATION.md sections 3, 4.
// UNVERIFIED against a real jaxlib.  What IS checked here: tools/check_ffi_compiles.sh (run by __graft_entry__.build())
// compiles this file against tools/ffi_stub/xla/ffi/api/ffi.h, a stand-in that declares the part of the public typed-FFI
// surface used below and static_asserts that every implementation takes the argument list its binding describes.
// They replace the per-device lowering of nn.Dense (case6_attention.py:61-90), of the attention core (:121-135), of the
// FFN the module comments describe (:36-40) and of the psum after a contraction-sharded dot (case1a.py:49).
#if defined(__has_include)
#if __has_include("xla/ffi/api/ffi.h")
#define B200_HAVE_XLA_FFI 1
#endif
#endif

#ifdef B200_HAVE_XLA_FFI
#include <cuda_runtime.h>

#include "../../include/b200_sharded.h"
#include "xla/ffi/api/ffi.h"

namespace ffi = xla::ffi;

static ffi::Error status_of(int rc) {
  if (rc == B200_OK) return ffi::Error::Success();
  return ffi::Error(rc == B200_ERR_INVALID_ARGUMENT ? ffi::ErrorCode::kInvalidArgument : ffi::ErrorCode::kInternal,
                    b200_last_error());
}

// y[..., N] = epilogue(x[..., K] . w[K, N] + bias[N]); epilogue = a b200_epilogue value
static ffi::Error DenseImpl(cudaStream_t stream, ffi::Buffer<ffi::BF16> x, ffi::Buffer<ffi::BF16> w,
                            ffi::Buffer<ffi::BF16> bias, ffi::ResultBuffer<ffi::BF16> y, int64_t epilogue) {
  auto xd = x.dimensions();
  auto wd = w.dimensions();
  const int64_t K = xd.back(), M = x.element_count() / K, N = wd[1];
  return status_of(b200_gemm_bf16(x.untyped_data(), K, w.untyped_data(), N, B200_B_KN,
                                  bias.element_count() ? bias.untyped_data() : nullptr, y->untyped_data(), N, B200_BF16,
                                  static_cast<int>(M), static_cast<int>(N), static_cast<int>(K), static_cast<int>(epilogue),
                                  stream));
}
XLA_FFI_DEFINE_HANDLER_SYMBOL(B200Dense, DenseImpl,
                              ffi::Ffi::Bind()
                                  .Ctx<ffi::PlatformStream<cudaStream_t>>()
                                  .Arg<ffi::Buffer<ffi::BF16>>()
                                  .Arg<ffi::Buffer<ffi::BF16>>()
                                  .Arg<ffi::Buffer<ffi::BF16>>()
                                  .Ret<ffi::Buffer<ffi::BF16>>()
                                  .Attr<int64_t>("epilogue"));

// q [B,Sq,H,D], k, v [B,Sk,H,D] contiguous ("b s n h"), o like q
static ffi::Error AttentionImpl(cudaStream_t stream, ffi::Buffer<ffi::BF16> q, ffi::Buffer<ffi::BF16> k,
                                ffi::Buffer<ffi::BF16> v, ffi::ResultBuffer<ffi::BF16> o, float scale, int64_t causal) {
  auto qd = q.dimensions();
  auto kd = k.dimensions();
  const int64_t B = qd[0], Sq = qd[1], H = qd[2], D = qd[3], Sk = kd[1];
  return status_of(b200_attention_fwd(q.untyped_data(), k.untyped_data(), v.untyped_data(), o->untyped_data(),
                                      static_cast<int>(B), static_cast<int>(Sq), static_cast<int>(Sk), static_cast<int>(H),
                                      static_cast<int>(D), Sq * H * D, H * D, D, Sk * H * D, H * D, D, Sk * H * D, H * D, D,
                                      Sq * H * D, H * D, D, scale, static_cast<int>(causal), stream));
}
XLA_FFI_DEFINE_HANDLER_SYMBOL(B200Attention, AttentionImpl,
                              ffi::Ffi::Bind()
                                  .Ctx<ffi::PlatformStream<cudaStream_t>>()
                                  .Arg<ffi::Buffer<ffi::BF16>>()
                                  .Arg<ffi::Buffer<ffi::BF16>>()
                                  .Arg<ffi::Buffer<ffi::BF16>>()
                                  .Ret<ffi::Buffer<ffi::BF16>>()
                                  .Attr<float>("scale")
                                  .Attr<int64_t>("causal"));

// One device's share of the FFN; h is declared as a second (scratch) result so that XLA allocates it.
static ffi::Error FfnImpl(cudaStream_t stream, ffi::Buffer<ffi::BF16> x, ffi::Buffer<ffi::BF16> w1,
                          ffi::Buffer<ffi::BF16> b1, ffi::Buffer<ffi::BF16> w2, ffi::Buffer<ffi::BF16> b2,
                          ffi::ResultBuffer<ffi::BF16> y, ffi::ResultBuffer<ffi::BF16> h, int64_t act) {
  auto xd = x.dimensions();
  auto w1d = w1.dimensions();
  const int64_t d_model = xd.back(), T = x.element_count() / d_model, F = w1d[1];
  return status_of(b200_ffn_fwd(x.untyped_data(), w1.untyped_data(), b1.element_count() ? b1.untyped_data() : nullptr,
                                w2.untyped_data(), b2.element_count() ? b2.untyped_data() : nullptr, h->untyped_data(),
                                y->untyped_data(), B200_BF16, static_cast<int>(T), static_cast<int>(d_model),
                                static_cast<int>(F), static_cast<int>(act), stream));
}
XLA_FFI_DEFINE_HANDLER_SYMBOL(B200Ffn, FfnImpl,
                              ffi::Ffi::Bind()
                                  .Ctx<ffi::PlatformStream<cudaStream_t>>()
                                  .Arg<ffi::Buffer<ffi::BF16>>()
                                  .Arg<ffi::Buffer<ffi::BF16>>()
                                  .Arg<ffi::Buffer<ffi::BF16>>()
                                  .Arg<ffi::Buffer<ffi::BF16>>()
                                  .Arg<ffi::Buffer<ffi::BF16>>()
                                  .Ret<ffi::Buffer<ffi::BF16>>()
                                  .Ret<ffi::Buffer<ffi::BF16>>()
                                  .Attr<int64_t>("act"));

// jax.lax.dot on one device's f32 shards (case1a.py:49 and siblings): c [M,N] = a [M,K] . b [K,N], true fp32;
// accumulate != 0 adds into c (summing K-partials locally)
static ffi::Error DotF32Impl(cudaStream_t stream, ffi::Buffer<ffi::F32> a, ffi::Buffer<ffi::F32> b,
                             ffi::ResultBuffer<ffi::F32> c, int64_t accumulate) {
  auto ad = a.dimensions();
  auto bd = b.dimensions();
  const int64_t K = ad.back(), M = a.element_count() / K, N = bd.back();
  return status_of(b200_gemm_f32(static_cast<const float*>(a.untyped_data()), K, M * K,
                                 static_cast<const float*>(b.untyped_data()), N, K * N,
                                 static_cast<float*>(c->untyped_data()), N, M * N, 1, static_cast<int>(M),
                                 static_cast<int>(N), static_cast<int>(K), static_cast<int>(accumulate), stream));
}
XLA_FFI_DEFINE_HANDLER_SYMBOL(B200DotF32, DotF32Impl,
                              ffi::Ffi::Bind()
                                  .Ctx<ffi::PlatformStream<cudaStream_t>>()
                                  .Arg<ffi::Buffer<ffi::F32>>()
                                  .Arg<ffi::Buffer<ffi::F32>>()
                                  .Ret<ffi::Buffer<ffi::F32>>()
                                  .Attr<int64_t>("accumulate"));

// Collectives over one mesh axis; `comm` is the b200_comm* of that axis, passed as an integer attribute by the binder.
// Element type f32 or bf16, taken from the buffer.
static int dtype_of(const ffi::AnyBuffer& b) { return b.element_type() == ffi::BF16 ? B200_BF16 : B200_F32; }

static ffi::Error AllReduceImpl(cudaStream_t stream, ffi::AnyBuffer x, ffi::AnyResultBuffer y, int64_t comm) {
  return status_of(b200_all_reduce(reinterpret_cast<b200_comm*>(comm), x.untyped_data(), y->untyped_data(),
                                   static_cast<int64_t>(x.element_count()), dtype_of(x), stream));
}
XLA_FFI_DEFINE_HANDLER_SYMBOL(B200AllReduce, AllReduceImpl,
                              ffi::Ffi::Bind()
                                  .Ctx<ffi::PlatformStream<cudaStream_t>>()
                                  .Arg<ffi::AnyBuffer>()
                                  .Ret<ffi::AnyBuffer>()
                                  .Attr<int64_t>("comm"));

// psum_scatter: x = axis_size blocks along dim 0, y = the sum of everybody's block [axis_rank]
static ffi::Error ReduceScatterImpl(cudaStream_t stream, ffi::AnyBuffer x, ffi::AnyResultBuffer y, int64_t comm) {
  return status_of(b200_reduce_scatter(reinterpret_cast<b200_comm*>(comm), x.untyped_data(), y->untyped_data(),
                                       static_cast<int64_t>(y->element_count()), dtype_of(x), stream));
}
XLA_FFI_DEFINE_HANDLER_SYMBOL(B200ReduceScatter, ReduceScatterImpl,
                              ffi::Ffi::Bind()
                                  .Ctx<ffi::PlatformStream<cudaStream_t>>()
                                  .Arg<ffi::AnyBuffer>()
                                  .Ret<ffi::AnyBuffer>()
                                  .Attr<int64_t>("comm"));

// all_gather along dim 0 in axis order (FSDP weight gather; K / V gather along S of the literal case6 layout)
static ffi::Error AllGatherImpl(cudaStream_t stream, ffi::AnyBuffer x, ffi::AnyResultBuffer y, int64_t comm) {
  return status_of(b200_all_gather(reinterpret_cast<b200_comm*>(comm), x.untyped_data(), y->untyped_data(),
                                   static_cast<int64_t>(x.element_count()), dtype_of(x), stream));
}
XLA_FFI_DEFINE_HANDLER_SYMBOL(B200AllGather, AllGatherImpl,
                              ffi::Ffi::Bind()
                                  .Ctx<ffi::PlatformStream<cudaStream_t>>()
                                  .Arg<ffi::AnyBuffer>()
                                  .Ret<ffi::AnyBuffer>()
                                  .Attr<int64_t>("comm"));

// all_to_all: block p of x (dim 0 split in axis_size blocks) goes to rank p (the M-sharded -> S-sharded re-split of
// case6_attention.py:141)
static ffi::Error AllToAllImpl(cudaStream_t stream, ffi::AnyBuffer x, ffi::AnyResultBuffer y, int64_t comm) {
  b200_comm* c = reinterpret_cast<b200_comm*>(comm);
  const int n = b200_comm_size(c);
  if (n <= 0) return ffi::Error(ffi::ErrorCode::kInvalidArgument, "B200AllToAll: bad communicator");
  return status_of(b200_all_to_all(c, x.untyped_data(), y->untyped_data(), static_cast<int64_t>(x.element_count()) / n,
                                   dtype_of(x), stream));
}
XLA_FFI_DEFINE_HANDLER_SYMBOL(B200AllToAll, AllToAllImpl,
                              ffi::Ffi::Bind()
                                  .Ctx<ffi::PlatformStream<cudaStream_t>>()
                                  .Arg<ffi::AnyBuffer>()
                                  .Ret<ffi::AnyBuffer>()
                                  .Attr<int64_t>("comm"));

// The tensor-parallel FFN with its reduction inside the call (b200_ffn_fwd_tp); h and the f32 partial are declared as
// extra results so that XLA allocates the scratch.  reduce_mode: 0 = psum (y [T, d_model]), 1 = psum_scatter (y [T/size, d_model]).
static ffi::Error FfnTpImpl(cudaStream_t stream, ffi::Buffer<ffi::BF16> x, ffi::Buffer<ffi::BF16> w1,
                            ffi::Buffer<ffi::BF16> b1, ffi::Buffer<ffi::BF16> w2, ffi::Buffer<ffi::BF16> b2,
                            ffi::ResultBuffer<ffi::BF16> y, ffi::ResultBuffer<ffi::BF16> h, ffi::ResultBuffer<ffi::F32> partial,
                            int64_t comm, int64_t reduce_mode, int64_t act, int64_t chunks) {
  auto xd = x.dimensions();
  auto w1d = w1.dimensions();
  const int64_t d_model = xd.back(), T = x.element_count() / d_model, F = w1d[1];
  return status_of(b200_ffn_fwd_tp(reinterpret_cast<b200_comm*>(comm), x.untyped_data(), w1.untyped_data(),
                                   b1.element_count() ? b1.untyped_data() : nullptr, w2.untyped_data(),
                                   b2.element_count() ? b2.untyped_data() : nullptr, h->untyped_data(),
                                   static_cast<float*>(partial->untyped_data()), y->untyped_data(),
                                   static_cast<int>(reduce_mode), static_cast<int>(T), static_cast<int>(d_model),
                                   static_cast<int>(F), static_cast<int>(act), static_cast<int>(chunks), stream));
}
XLA_FFI_DEFINE_HANDLER_SYMBOL(B200FfnTp, FfnTpImpl,
                              ffi::Ffi::Bind()
                                  .Ctx<ffi::PlatformStream<cudaStream_t>>()
                                  .Arg<ffi::Buffer<ffi::BF16>>()
                                  .Arg<ffi::Buffer<ffi::BF16>>()
                                  .Arg<ffi::Buffer<ffi::BF16>>()
                                  .Arg<ffi::Buffer<ffi::BF16>>()
                                  .Arg<ffi::Buffer<ffi::BF16>>()
                                  .Ret<ffi::Buffer<ffi::BF16>>()
                                  .Ret<ffi::Buffer<ffi::BF16>>()
                                  .Ret<ffi::Buffer<ffi::F32>>()
                                  .Attr<int64_t>("comm")
                                  .Attr<int64_t>("reduce_mode")
                                  .Attr<int64_t>("act")
                                  .Attr<int64_t>("chunks"));
#endif  // B200_HAVE_XLA_FFI
