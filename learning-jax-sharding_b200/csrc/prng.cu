// Counter-based random initialisation, born sharded: every device fills ITS OWN slice of a parameter directly in HBM,
// with the values jax.random gives the full array -- no host generation, no device_put, no communication.
//
// Replaces, per device, what the reference gets from jax.random.normal(key, shape) (case1a.py:16-18;
// case6_attention.py:152) and from model.init under jit with out_shardings (case6_attention.py:171-196), where flax
// draws nn.initializers.lecun_normal() for every Dense kernel (case6_attention.py:57,86): threefry2x32 (20 rounds)
// over jax's counter layout -- counts = iota(n) padded to even length and split in halves, element i of the flat array
// takes word y0[i] (i < h) or y1[i - h] of threefry(key, (j, x1[j])) -- then bits -> uniform -> sqrt(2) erf_inv(u)
// (Giles' single-precision polynomial as XLA evaluates it), optionally truncated to (-2, 2) and scaled (lecun_normal).
// Because the generator is counter-based, element i depends on (key, i) only: a shard is a strided walk over i.
// Bound: integer ALU (about 110 integer operations per element against 4 bytes written).
#include <math.h>

#include "host_common.h"

namespace b200 {

struct RngParams {
  uint32_t k0, k1;
  int ndim;
  long long gstride[4];   // element strides of the GLOBAL row-major array
  long long start[4];     // first global index of the shard per dim
  long long sshape[4];    // shard extent per dim
  long long n_shard;      // elements of the shard
  long long n_global;     // elements of the whole array
  int dist;
  float lo, hi;           // uniform range (after the erf transform for the truncated normal)
  float clip_lo, clip_hi; // truncated normal: open-interval clip
  float scale;            // multiplies the draw (lecun_normal's stddev)
};

__device__ __forceinline__ uint32_t rotl32(uint32_t x, int r) { return (x << r) | (x >> (32 - r)); }

__device__ __forceinline__ void threefry2x32(uint32_t k0, uint32_t k1, uint32_t& x0, uint32_t& x1) {
  const uint32_t ks[3] = {k0, k1, k0 ^ k1 ^ 0x1BD11BDAu};
  x0 += ks[0];
  x1 += ks[1];
#pragma unroll
  for (int block = 0; block < 5; ++block) {
    const int r0 = (block & 1) ? 17 : 13, r1 = (block & 1) ? 29 : 15, r2 = (block & 1) ? 16 : 26, r3 = (block & 1) ? 24 : 6;
    x0 += x1; x1 = rotl32(x1, r0) ^ x0;
    x0 += x1; x1 = rotl32(x1, r1) ^ x0;
    x0 += x1; x1 = rotl32(x1, r2) ^ x0;
    x0 += x1; x1 = rotl32(x1, r3) ^ x0;
    x0 += ks[(block + 1) % 3];
    x1 += ks[(block + 2) % 3] + static_cast<uint32_t>(block + 1);
  }
}

// word i of jax's _threefry_random_bits_original(key, 32, n)
__device__ __forceinline__ uint32_t random_word(uint32_t k0, uint32_t k1, long long i, long long n) {
  const long long h = (n + 1) / 2;
  const long long j = i < h ? i : i - h;
  uint32_t x0 = static_cast<uint32_t>(j);
  uint32_t x1 = (h + j < n) ? static_cast<uint32_t>(h + j) : 0u;   // the pad word of an odd-length count vector is 0
  threefry2x32(k0, k1, x0, x1);
  return i < h ? x0 : x1;
}

// every operation rounded separately (no fma contraction), like the XLA / numpy evaluation the oracle restates
__device__ __forceinline__ float erf_inv_f32(float x) {
  float w = -log1pf(__fmul_rn(-x, x));
  const bool lt = w < 5.0f;
  w = lt ? __fadd_rn(w, -2.5f) : __fadd_rn(sqrtf(w), -3.0f);
  float p = lt ? 2.81022636e-08f : -0.000200214257f;
  p = __fadd_rn(lt ? 3.43273939e-07f : 0.000100950558f, __fmul_rn(p, w));
  p = __fadd_rn(lt ? -3.5233877e-06f : 0.00134934322f, __fmul_rn(p, w));
  p = __fadd_rn(lt ? -4.39150654e-06f : -0.00367342844f, __fmul_rn(p, w));
  p = __fadd_rn(lt ? 0.00021858087f : 0.00573950773f, __fmul_rn(p, w));
  p = __fadd_rn(lt ? -0.00125372503f : -0.0076224613f, __fmul_rn(p, w));
  p = __fadd_rn(lt ? -0.00417768164f : 0.00943887047f, __fmul_rn(p, w));
  p = __fadd_rn(lt ? 0.246640727f : 1.00167406f, __fmul_rn(p, w));
  p = __fadd_rn(lt ? 1.50140941f : 2.83297682f, __fmul_rn(p, w));
  if (fabsf(x) == 1.0f) return copysignf(INFINITY, x);
  return __fmul_rn(p, x);
}

__global__ void __launch_bounds__(256) random_fill_kernel(const RngParams p, void* __restrict__ out) {
  const long long stride = static_cast<long long>(gridDim.x) * blockDim.x;
  for (long long e = static_cast<long long>(blockIdx.x) * blockDim.x + threadIdx.x; e < p.n_shard; e += stride) {
    // shard-local linear index -> global linear index
    long long rem = e, gi = 0;
#pragma unroll
    for (int d = 3; d >= 0; --d) {
      if (d < p.ndim) {
        const long long c = rem % p.sshape[d];
        rem /= p.sshape[d];
        gi += (p.start[d] + c) * p.gstride[d];
      }
    }
    const uint32_t bits = random_word(p.k0, p.k1, gi, p.n_global);
    if (p.dist == B200_RNG_BITS) {
      reinterpret_cast<uint32_t*>(out)[e] = bits;
      continue;
    }
    const float f = __fadd_rn(__uint_as_float((bits >> 9) | 0x3F800000u), -1.0f);
    float u = fmaxf(p.lo, __fadd_rn(__fmul_rn(f, __fadd_rn(p.hi, -p.lo)), p.lo));
    float v = u;
    if (p.dist == B200_RNG_NORMAL || p.dist == B200_RNG_TRUNCATED_NORMAL) {
      v = __fmul_rn(1.41421356237309504880f, erf_inv_f32(u));
      if (p.dist == B200_RNG_TRUNCATED_NORMAL) v = fminf(fmaxf(v, p.clip_lo), p.clip_hi);
    }
    reinterpret_cast<float*>(out)[e] = __fmul_rn(v, p.scale);
  }
}

}  // namespace b200

using namespace b200;

extern "C" int b200_random_fill(const b200_random_args* a, void* out, void* stream) {
  B200_REQUIRE(a && out, "b200_random_fill: null argument");
  B200_REQUIRE(a->ndim >= 1 && a->ndim <= 4, "b200_random_fill: ndim must be 1..4 (got %d)", a->ndim);
  B200_REQUIRE(a->distribution >= B200_RNG_BITS && a->distribution <= B200_RNG_TRUNCATED_NORMAL,
               "b200_random_fill: bad distribution %d", a->distribution);
  RngParams p = {};
  p.k0 = a->key[0];
  p.k1 = a->key[1];
  p.ndim = a->ndim;
  long long n_global = 1, n_shard = 1;
  for (int d = a->ndim - 1; d >= 0; --d) {
    B200_REQUIRE(a->global_shape[d] > 0 && a->shard_shape[d] > 0 && a->shard_start[d] >= 0 &&
                     a->shard_start[d] + a->shard_shape[d] <= a->global_shape[d],
                 "b200_random_fill: shard [%lld, +%lld) does not lie inside dim %d of extent %lld",
                 (long long)a->shard_start[d], (long long)a->shard_shape[d], d, (long long)a->global_shape[d]);
    p.gstride[d] = n_global;
    p.start[d] = a->shard_start[d];
    p.sshape[d] = a->shard_shape[d];
    n_global *= a->global_shape[d];
    n_shard *= a->shard_shape[d];
  }
  // jax's count vector is uint32: arrays beyond 2^32 words use a different (multi-block) path that is not restated here
  B200_REQUIRE(n_global < (1ll << 32), "b200_random_fill: arrays of 2^32 or more elements are not supported");
  p.n_global = n_global;
  p.n_shard = n_shard;
  p.dist = a->distribution;
  p.scale = a->scale;
  if (a->distribution == B200_RNG_UNIFORM) {
    B200_REQUIRE(a->hi > a->lo, "b200_random_fill: uniform needs hi > lo");
    p.lo = a->lo;
    p.hi = a->hi;
  } else if (a->distribution == B200_RNG_NORMAL) {
    p.lo = nextafterf(-1.0f, 0.0f);
    p.hi = 1.0f;
  } else if (a->distribution == B200_RNG_TRUNCATED_NORMAL) {
    B200_REQUIRE(a->hi > a->lo, "b200_random_fill: truncated normal needs upper > lower");
    const float s2 = 1.41421356237309504880f;
    p.lo = static_cast<float>(erf(static_cast<double>(a->lo / s2)));
    p.hi = static_cast<float>(erf(static_cast<double>(a->hi / s2)));
    p.clip_lo = nextafterf(a->lo, INFINITY);
    p.clip_hi = nextafterf(a->hi, -INFINITY);
  }
  if (p.scale == 0.0f && a->distribution != B200_RNG_BITS) p.scale = 1.0f;
  long long blocks = (n_shard + 255) / 256;
  const long long cap = static_cast<long long>(num_sms()) * 16;
  if (blocks > cap) blocks = cap;
  random_fill_kernel<<<static_cast<int>(blocks), 256, 0, reinterpret_cast<cudaStream_t>(stream)>>>(p, out);
  B200_CUDA_CHECK(cudaGetLastError());
  launch_counter().fetch_add(1);
  return B200_OK;
}
