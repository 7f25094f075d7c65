// HBM-bound helpers around the GEMMs: dtype casts (flax promote_dtype casts f32 params and
// activations to bf16 on every nn.Dense call, reference case6_attention.py:61-90) and the
// finishing pass of a row-parallel reduction (sum of f32 partials -> bf16 (+bias)).
// 16-byte vector accesses, grid sized to a multiple of the SM count, grid-stride loops.
#include "host_common.h"
#include "sm100_ptx.cuh"

namespace b200 {

__global__ void __launch_bounds__(256) cast_f32_to_bf16_kernel(const float* __restrict__ src,
                                                              __nv_bfloat16* __restrict__ dst, int64_t n) {
  const int64_t nvec = n / 8;
  const int64_t stride = static_cast<int64_t>(gridDim.x) * blockDim.x;
  for (int64_t i = static_cast<int64_t>(blockIdx.x) * blockDim.x + threadIdx.x; i < nvec; i += stride) {
    const float4 a = reinterpret_cast<const float4*>(src)[2 * i];
    const float4 b = reinterpret_cast<const float4*>(src)[2 * i + 1];
    uint4 o;
    o.x = pack_bf16x2(a.x, a.y);
    o.y = pack_bf16x2(a.z, a.w);
    o.z = pack_bf16x2(b.x, b.y);
    o.w = pack_bf16x2(b.z, b.w);
    reinterpret_cast<uint4*>(dst)[i] = o;
  }
  if (blockIdx.x == 0) {
    for (int64_t i = nvec * 8 + threadIdx.x; i < n; i += blockDim.x) dst[i] = __float2bfloat16_rn(src[i]);
  }
}

__global__ void __launch_bounds__(256) cast_bf16_to_f32_kernel(const __nv_bfloat16* __restrict__ src,
                                                              float* __restrict__ dst, int64_t n) {
  const int64_t nvec = n / 8;
  const int64_t stride = static_cast<int64_t>(gridDim.x) * blockDim.x;
  for (int64_t i = static_cast<int64_t>(blockIdx.x) * blockDim.x + threadIdx.x; i < nvec; i += stride) {
    const uint4 v = reinterpret_cast<const uint4*>(src)[i];
    const __nv_bfloat162* h = reinterpret_cast<const __nv_bfloat162*>(&v);
    const float2 f0 = __bfloat1622float2(h[0]), f1 = __bfloat1622float2(h[1]);
    const float2 f2 = __bfloat1622float2(h[2]), f3 = __bfloat1622float2(h[3]);
    reinterpret_cast<float4*>(dst)[2 * i] = make_float4(f0.x, f0.y, f1.x, f1.y);
    reinterpret_cast<float4*>(dst)[2 * i + 1] = make_float4(f2.x, f2.y, f3.x, f3.y);
  }
  if (blockIdx.x == 0) {
    for (int64_t i = nvec * 8 + threadIdx.x; i < n; i += blockDim.x) dst[i] = __bfloat162float(src[i]);
  }
}

// n and cols are multiples of 4; dst 8-byte aligned
__global__ void __launch_bounds__(256) sum_partials_kernel(const float* __restrict__ src, int64_t src_stride,
                                                          int nsrc, const __nv_bfloat16* __restrict__ bias,
                                                          int64_t cols, __nv_bfloat16* __restrict__ dst,
                                                          int64_t n) {
  const int64_t nvec = n / 4;
  const int64_t stride = static_cast<int64_t>(gridDim.x) * blockDim.x;
  for (int64_t i = static_cast<int64_t>(blockIdx.x) * blockDim.x + threadIdx.x; i < nvec; i += stride) {
    float4 acc = reinterpret_cast<const float4*>(src)[i];
    for (int r = 1; r < nsrc; ++r) {
      const float4 v = reinterpret_cast<const float4*>(src + r * src_stride)[i];
      acc.x += v.x;
      acc.y += v.y;
      acc.z += v.z;
      acc.w += v.w;
    }
    float o[4] = {bf16_round(acc.x), bf16_round(acc.y), bf16_round(acc.z), bf16_round(acc.w)};
    if (bias) {
      const int64_t c = (i * 4) % cols;
#pragma unroll
      for (int t = 0; t < 4; ++t) o[t] = o[t] + __bfloat162float(bias[c + t]);
    }
    uint2 out;
    out.x = pack_bf16x2(o[0], o[1]);
    out.y = pack_bf16x2(o[2], o[3]);
    reinterpret_cast<uint2*>(dst)[i] = out;
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


// Blockwise ("ring") attention, merge step: fold the result of one more key/value block into the running result of a
// query block.  o_blk [rows, D] bf16 is that block's softmax-normalised output, lse_blk [rows] its log2-sum-exp
// (b200_attention_fwd_lse); acc [rows, D] f32 / lse_acc [rows] are the running pair:
//   lse' = log2(2^lse_acc + 2^lse_blk),  acc' = acc * 2^(lse_acc - lse') + o_blk * 2^(lse_blk - lse').
// first != 0 starts the running pair from the block.  Row r of acc is (b, s, h) of the [B, S, H, D] tensor; the lse
// arrays are [B, H, S] as the attention kernels write them.  One warp per row, 16-byte accesses.
template <int D>
__global__ void __launch_bounds__(256) attention_combine_kernel(float* __restrict__ acc, float* __restrict__ lse_acc,
                                                               const __nv_bfloat16* __restrict__ o_blk,
                                                               const float* __restrict__ lse_blk, int64_t rows, int S,
                                                               int H, int first) {
  const int lane = threadIdx.x & 31;
  const int64_t warps = (static_cast<int64_t>(gridDim.x) * blockDim.x) >> 5;
  for (int64_t r = (static_cast<int64_t>(blockIdx.x) * blockDim.x + threadIdx.x) >> 5; r < rows; r += warps) {
    const int h = static_cast<int>(r % H);
    const int64_t bs = r / H;
    const int s = static_cast<int>(bs % S);
    const int64_t b = bs / S;
    const int64_t li = (b * H + h) * S + s;
    const float lb = lse_blk[li];
    float w_old = 0.f, w_new = 1.f, l_new = lb;
    if (!first) {
      const float la = lse_acc[li];
      const float m = fmaxf(la, lb);
      l_new = m + log2f(exp2f(la - m) + exp2f(lb - m));
      w_old = exp2f(la - l_new);
      w_new = exp2f(lb - l_new);
    }
    __syncwarp();
    if (lane == 0) lse_acc[li] = l_new;
    constexpr int kPer = D / 32;                       // 2 or 4 elements per lane
    float* a = acc + r * D + lane * kPer;
    const __nv_bfloat16* o = o_blk + r * D + lane * kPer;
    if (kPer == 4) {
      const uint2 raw = *reinterpret_cast<const uint2*>(o);
      const float2 o01 = __bfloat1622float2(*reinterpret_cast<const __nv_bfloat162*>(&raw.x));
      const float2 o23 = __bfloat1622float2(*reinterpret_cast<const __nv_bfloat162*>(&raw.y));
      float4 v = first ? make_float4(0.f, 0.f, 0.f, 0.f) : *reinterpret_cast<const float4*>(a);
      v.x = v.x * w_old + o01.x * w_new;
      v.y = v.y * w_old + o01.y * w_new;
      v.z = v.z * w_old + o23.x * w_new;
      v.w = v.w * w_old + o23.y * w_new;
      *reinterpret_cast<float4*>(a) = v;
    } else {
      const float2 o01 = __bfloat1622float2(*reinterpret_cast<const __nv_bfloat162*>(o));
      float2 v = first ? make_float2(0.f, 0.f) : *reinterpret_cast<const float2*>(a);
      v.x = v.x * w_old + o01.x * w_new;
      v.y = v.y * w_old + o01.y * w_new;
      *reinterpret_cast<float2*>(a) = v;
    }
  }
}

}  // namespace b200

using namespace b200;

extern "C" int b200_cast_f32_to_bf16(const float* src, void* dst, int64_t n, void* stream) {
  B200_REQUIRE(src && dst && n >= 0, "b200_cast_f32_to_bf16: bad arguments");
  B200_REQUIRE(reinterpret_cast<uintptr_t>(src) % 16 == 0 && reinterpret_cast<uintptr_t>(dst) % 16 == 0,
               "b200_cast_f32_to_bf16: buffers must be 16-byte aligned");
  if (n == 0) return B200_OK;
  cast_f32_to_bf16_kernel<<<grid_for(n / 8 + 1), 256, 0, reinterpret_cast<cudaStream_t>(stream)>>>(
      src, reinterpret_cast<__nv_bfloat16*>(dst), n);
  B200_CUDA_CHECK(cudaGetLastError());
  launch_counter().fetch_add(1);
  return B200_OK;
}

extern "C" int b200_cast_bf16_to_f32(const void* src, float* dst, int64_t n, void* stream) {
  B200_REQUIRE(src && dst && n >= 0, "b200_cast_bf16_to_f32: bad arguments");
  B200_REQUIRE(reinterpret_cast<uintptr_t>(src) % 16 == 0 && reinterpret_cast<uintptr_t>(dst) % 16 == 0,
               "b200_cast_bf16_to_f32: buffers must be 16-byte aligned");
  if (n == 0) return B200_OK;
  cast_bf16_to_f32_kernel<<<grid_for(n / 8 + 1), 256, 0, reinterpret_cast<cudaStream_t>(stream)>>>(
      reinterpret_cast<const __nv_bfloat16*>(src), dst, n);
  B200_CUDA_CHECK(cudaGetLastError());
  launch_counter().fetch_add(1);
  return B200_OK;
}

extern "C" int b200_sum_partials_to_bf16(const float* src, int64_t src_stride, int nsrc, const void* bias,
                                         int64_t cols, void* dst, int64_t n, void* stream) {
  B200_REQUIRE(src && dst && nsrc >= 1 && n >= 0, "b200_sum_partials_to_bf16: bad arguments");
  B200_REQUIRE(n % 4 == 0 && src_stride % 4 == 0, "b200_sum_partials_to_bf16: n and src_stride must be multiples of 4");
  B200_REQUIRE(!bias || (cols > 0 && cols % 4 == 0 && n % cols == 0),
               "b200_sum_partials_to_bf16: cols must be a positive multiple of 4 dividing n when bias is given");
  B200_REQUIRE(reinterpret_cast<uintptr_t>(src) % 16 == 0 && reinterpret_cast<uintptr_t>(dst) % 8 == 0,
               "b200_sum_partials_to_bf16: misaligned buffers");
  if (n == 0) return B200_OK;
  sum_partials_kernel<<<grid_for(n / 4), 256, 0, reinterpret_cast<cudaStream_t>(stream)>>>(
      src, src_stride, nsrc, reinterpret_cast<const __nv_bfloat16*>(bias), cols > 0 ? cols : 4,
      reinterpret_cast<__nv_bfloat16*>(dst), n);
  B200_CUDA_CHECK(cudaGetLastError());
  launch_counter().fetch_add(1);
  return B200_OK;
}

extern "C" int b200_attention_combine(float* acc, float* lse_acc, const void* o_blk, const float* lse_blk, int B, int S, int H,
                                      int D, int first, void* stream) {
  B200_REQUIRE(acc && lse_acc && o_blk && lse_blk, "b200_attention_combine: null pointer");
  B200_REQUIRE(B > 0 && S > 0 && H > 0 && (D == 64 || D == 128), "b200_attention_combine: bad sizes (head dim 64 or 128)");
  B200_REQUIRE((reinterpret_cast<uintptr_t>(acc) | reinterpret_cast<uintptr_t>(o_blk)) % 16 == 0,
               "b200_attention_combine: acc and o_blk must be 16-byte aligned");
  const int64_t rows = static_cast<int64_t>(B) * S * H;
  int64_t blocks = (rows + 7) / 8;
  const int64_t cap = static_cast<int64_t>(num_sms()) * 8;
  if (blocks > cap) blocks = cap;
  cudaStream_t st = reinterpret_cast<cudaStream_t>(stream);
  if (D == 128)
    attention_combine_kernel<128><<<static_cast<int>(blocks), 256, 0, st>>>(acc, lse_acc, reinterpret_cast<const __nv_bfloat16*>(o_blk),
                                                                         lse_blk, rows, S, H, first);
  else
    attention_combine_kernel<64><<<static_cast<int>(blocks), 256, 0, st>>>(acc, lse_acc, reinterpret_cast<const __nv_bfloat16*>(o_blk),
                                                                        lse_blk, rows, S, H, first);
  B200_CUDA_CHECK(cudaGetLastError());
  launch_counter().fetch_add(1);
  return B200_OK;
}
