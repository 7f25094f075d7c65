// optax.adam(learning_rate) as the reference's train_step applies it (case5_attention_dense.py:107,130-139;
// case6_attention.py:181,206-220: TrainState.apply_gradients -> scale_by_adam -> scale(-lr) -> apply_updates), one
// HBM-bound pass per parameter tensor: 16 B of reads (param, mu, nu, grad) and 12 B of writes per element in f32.
//   mu' = b1 mu + (1 - b1) g;  nu' = b2 nu + (1 - b2) g^2;  param' = param - lr * (mu' / (1 - b1^t)) / (sqrt(nu' / (1 - b2^t)) + eps)
// 16-byte vector accesses, grid = a multiple of the SM count, grid-stride loop.  Nothing here is taken from the reference
// (it has no native code); the arithmetic order follows optax's published update rule.
#include <math.h>

#include "host_common.h"
#include "sm100_ptx.cuh"

namespace b200 {

struct AdamConsts {
  float lr, b1, b2, eps, one_minus_b1, one_minus_b2, bc1, bc2;   // bc = 1 - b^t
};

__device__ __forceinline__ void adam_one(float& p, float g, float& m, float& v, const AdamConsts& c) {
  m = c.b1 * m + c.one_minus_b1 * g;
  v = c.b2 * v + c.one_minus_b2 * (g * g);
  const float mh = m / c.bc1;
  const float vh = v / c.bc2;
  p = p - c.lr * (mh / (sqrtf(vh) + c.eps));
}

template <bool kGradBf16>
__global__ void __launch_bounds__(256) adam_step_kernel(float* __restrict__ param, const void* __restrict__ grad,
                                                       float* __restrict__ mu, float* __restrict__ nu, int64_t n,
                                                       const AdamConsts c) {
  const int64_t nvec = n / 4;
  const int64_t stride = static_cast<int64_t>(gridDim.x) * blockDim.x;
  for (int64_t i = static_cast<int64_t>(blockIdx.x) * blockDim.x + threadIdx.x; i < nvec; i += stride) {
    float4 p = reinterpret_cast<float4*>(param)[i];
    float4 m = reinterpret_cast<float4*>(mu)[i];
    float4 v = reinterpret_cast<float4*>(nu)[i];
    float4 g;
    if (kGradBf16) {
      const uint2 raw = reinterpret_cast<const uint2*>(grad)[i];
      const float2 g0 = __bfloat1622float2(*reinterpret_cast<const __nv_bfloat162*>(&raw.x));
      const float2 g1 = __bfloat1622float2(*reinterpret_cast<const __nv_bfloat162*>(&raw.y));
      g = make_float4(g0.x, g0.y, g1.x, g1.y);
    } else {
      g = reinterpret_cast<const float4*>(grad)[i];
    }
    adam_one(p.x, g.x, m.x, v.x, c);
    adam_one(p.y, g.y, m.y, v.y, c);
    adam_one(p.z, g.z, m.z, v.z, c);
    adam_one(p.w, g.w, m.w, v.w, c);
    reinterpret_cast<float4*>(param)[i] = p;
    reinterpret_cast<float4*>(mu)[i] = m;
    reinterpret_cast<float4*>(nu)[i] = v;
  }
  if (blockIdx.x == 0) {
    for (int64_t i = nvec * 4 + threadIdx.x; i < n; i += blockDim.x) {
      const float g = kGradBf16 ? __bfloat162float(reinterpret_cast<const __nv_bfloat16*>(grad)[i])
                                : reinterpret_cast<const float*>(grad)[i];
      adam_one(param[i], g, mu[i], nu[i], c);
    }
  }
}

static int grid_for(int64_t work_items) {
  const int sms = num_sms();
  int64_t blocks = (work_items + 255) / 256;
  const int64_t cap = static_cast<int64_t>(sms) * 8;   // 8 resident 256-thread blocks per SM
  if (blocks > cap) blocks = cap;
  if (blocks < 1) blocks = 1;
  return static_cast<int>(blocks);
}

}  // namespace b200

using namespace b200;

extern "C" int b200_adam_step(float* param, const void* grad, int grad_dtype, float* mu, float* nu, int64_t n, float lr,
                              float b1, float b2, float eps, int step, void* stream) {
  B200_REQUIRE(param && grad && mu && nu && n >= 0, "b200_adam_step: bad arguments");
  B200_REQUIRE(grad_dtype == B200_BF16 || grad_dtype == B200_F32, "b200_adam_step: grad_dtype must be bf16 or f32");
  B200_REQUIRE(step >= 1, "b200_adam_step: step is the 1-based update count (got %d)", step);
  B200_REQUIRE(b1 >= 0.f && b1 < 1.f && b2 >= 0.f && b2 < 1.f, "b200_adam_step: betas must be in [0, 1)");
  B200_REQUIRE((reinterpret_cast<uintptr_t>(param) | reinterpret_cast<uintptr_t>(grad) | reinterpret_cast<uintptr_t>(mu) |
                reinterpret_cast<uintptr_t>(nu)) % 16 == 0,
               "b200_adam_step: buffers must be 16-byte aligned");
  if (n == 0) return B200_OK;
  AdamConsts c;
  c.lr = lr;
  c.b1 = b1;
  c.b2 = b2;
  c.eps = eps;
  c.one_minus_b1 = 1.0f - b1;
  c.one_minus_b2 = 1.0f - b2;
  c.bc1 = 1.0f - powf(b1, static_cast<float>(step));
  c.bc2 = 1.0f - powf(b2, static_cast<float>(step));
  cudaStream_t st = reinterpret_cast<cudaStream_t>(stream);
  if (grad_dtype == B200_BF16) adam_step_kernel<true><<<grid_for(n / 4 + 1), 256, 0, st>>>(param, grad, mu, nu, n, c);
  else adam_step_kernel<false><<<grid_for(n / 4 + 1), 256, 0, st>>>(param, grad, mu, nu, n, c);
  B200_CUDA_CHECK(cudaGetLastError());
  launch_counter().fetch_add(1);
  return B200_OK;
}
