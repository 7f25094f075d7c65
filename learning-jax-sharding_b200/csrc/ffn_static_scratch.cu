// Feed-forward block for callers that cannot hand the library a workspace.
//
// b200_ffn_fwd needs an intermediate buffer h [T, F] between its two GEMMs and expects the caller
// to allocate it.  Some hosts of this library cannot do that at the point of the call: a handler
// running inside a framework that has already reserved (nearly) all of HBM for its own allocator --
// XLA's preallocating BFC allocator is the usual case for the jax scripts this library serves --
// finds cudaMalloc failing or stalling the device, and a custom call whose signature was declared
// without a scratch operand has nowhere to receive one.  For them the entry point below keeps h in
// a fixed array that is part of the library image: nothing is allocated or freed, ever.
#include <mutex>

#include "host_common.h"

// B200_FFN_STATIC_SCRATCH_BYTES (b200_sharded.h): 32 MiB per GPU that loads the library, held until
// the process exits.

// In host code this name is only a handle for the runtime, not an address a kernel can use;
// the real device address comes from cudaGetSymbolAddress below.
static __device__ __align__(256) unsigned char g_ffn_static_scratch[B200_FFN_STATIC_SCRATCH_BYTES];

namespace {

constexpr int kMaxDevices = 64;
std::mutex g_setup_mutex;
void* g_scratch_addr[kMaxDevices] = {nullptr};
bool g_scratch_ready[kMaxDevices] = {false};

// Device address of the scratch array on the CURRENT device; resolved once per device.
int static_scratch_for_current_device(void** out) {
  int dev = -1;
  B200_CUDA_CHECK(cudaGetDevice(&dev));
  if (dev < 0 || dev >= kMaxDevices)
    return b200::fail(B200_ERR_UNSUPPORTED, "b200_ffn_fwd_static_scratch: device ordinal %d out of range", dev);
  std::lock_guard<std::mutex> lock(g_setup_mutex);
  if (!g_scratch_ready[dev]) {
    void* ptr = nullptr;
    cudaError_t e = cudaGetSymbolAddress(&ptr, g_ffn_static_scratch);
    if (e != cudaSuccess || ptr == nullptr)
      return b200::fail(B200_ERR_CUDA, "b200_ffn_fwd_static_scratch: cudaGetSymbolAddress failed on device %d: %s",
                        dev, cudaGetErrorString(e));
    g_scratch_addr[dev] = ptr;
    g_scratch_ready[dev] = true;
  }
  *out = g_scratch_addr[dev];
  return B200_OK;
}

}  // namespace

extern "C" int b200_ffn_fwd_static_scratch(const void* x, const void* W1, const void* b1, const void* W2,
                                           const void* b2, void* y, int y_dtype, int T, int d_model, int F, int act,
                                           void* stream) {
  B200_REQUIRE(x && W1 && W2 && y, "b200_ffn_fwd_static_scratch: null operand pointer");
  B200_REQUIRE(T > 0 && d_model > 0 && F > 0, "b200_ffn_fwd_static_scratch: T, d_model, F must be positive");
  const unsigned long long need = 2ull * static_cast<unsigned long long>(T) * static_cast<unsigned long long>(F);
  if (need > B200_FFN_STATIC_SCRATCH_BYTES)
    return b200::fail(B200_ERR_INVALID_ARGUMENT,
                      "b200_ffn_fwd_static_scratch: the intermediate h [T=%d, F=%d] bf16 needs %llu bytes but the "
                      "built-in scratch holds %llu; split the tokens into smaller calls or use b200_ffn_fwd with a "
                      "workspace of your own",
                      T, F, need, static_cast<unsigned long long>(B200_FFN_STATIC_SCRATCH_BYTES));
  void* h = nullptr;
  int rc = static_scratch_for_current_device(&h);
  if (rc != B200_OK) return rc;
  // Nothing in h has to start at zero: the first GEMM writes every element [T, F] the second one reads,
  // so whatever an earlier call left behind is simply overwritten.
  const bool act_bias = act == B200_EPI_BIAS_RELU || act == B200_EPI_BIAS_GELU_TANH;
  B200_REQUIRE(act == B200_EPI_BIAS_RELU || act == B200_EPI_BIAS_GELU_TANH || act == B200_EPI_RELU ||
                   act == B200_EPI_GELU_TANH,
               "b200_ffn_fwd_static_scratch: act must be one of the RELU / GELU_TANH epilogues (got %d)", act);
  B200_REQUIRE(!act_bias || b1, "b200_ffn_fwd_static_scratch: activation %d needs b1", act);
  B200_REQUIRE(!(y_dtype == B200_F32 && b2),
               "b200_ffn_fwd_static_scratch: f32 partial output takes no b2 (add it after the reduction)");
  rc = b200_gemm_bf16(x, d_model, W1, F, B200_B_KN, act_bias ? b1 : nullptr, h, F, B200_BF16, T, F, d_model, act,
                      stream);
  if (rc != B200_OK) return rc;
  return b200_gemm_bf16(h, F, W2, d_model, B200_B_KN, b2, y, d_model, y_dtype, T, d_model, F,
                        b2 ? B200_EPI_BIAS : B200_EPI_NONE, stream);
}
