// XLA typed-FFI handlers over the C ABI ("each of (1)-(4) additionally gets an XLA-FFI-typed
// wrapper behind #if __has_include").  jaxlib's headers (xla/ffi/api/ffi.h) are NOT in this image, so in this build the
// translation unit is empty; with them on the include path (a maintainer adds -I$(python -c "import jaxlib, os;
// print(os.path.join(os.path.dirname(jaxlib.__file__), 'include'))") to B200_EXTRA_NVCC_FLAGS) the library also exports
//   B200Dense, B200Attention, B200Ffn, B200AllReduce
// for jax.ffi.register_ffi_target(name, jax.ffi.pycapsule(lib.<Symbol>), platform="CUDA") -- INTEGRATION.md sections 3, 4.
// UNVERIFIED against a real jaxlib: written to the public FFI API, never compiled here.
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

// psum over one mesh axis; `comm` is the b200_comm* of that axis, passed as an integer attribute by the binder
static ffi::Error AllReduceImpl(cudaStream_t stream, ffi::Buffer<ffi::F32> x, ffi::ResultBuffer<ffi::F32> y, int64_t comm) {
  return status_of(b200_all_reduce(reinterpret_cast<b200_comm*>(comm), x.untyped_data(), y->untyped_data(),
                                   static_cast<int64_t>(x.element_count()), B200_F32, stream));
}
XLA_FFI_DEFINE_HANDLER_SYMBOL(B200AllReduce, AllReduceImpl,
                              ffi::Ffi::Bind()
                                  .Ctx<ffi::PlatformStream<cudaStream_t>>()
                                  .Arg<ffi::Buffer<ffi::F32>>()
                                  .Ret<ffi::Buffer<ffi::F32>>()
                                  .Attr<int64_t>("comm"));
#endif  // B200_HAVE_XLA_FFI
