// Multi-head attention forward for sm_100a: O = softmax(scale * Q K^T [+causal]) V with an online
// softmax that never materialises the (b, n, f, t) score tensor.
//
// Math order follows the reference module (case6_attention.py:121-135): scores and softmax in
// f32, probabilities rounded to bf16 before P.V, f32 accumulation, bf16 result.  Layout is the
// reference's "b s n h" (head dim innermost), read straight out of the projection output.
//
// One CTA = one (batch, head, 128-query tile); KV is streamed in 128-key tiles:
//   warp 0      TMA producer  Q once, K and V tiles through two 2-stage rings (128B swizzle)
//   warp 1      MMA issuer    S = Q.K^T (tcgen05, f32 in TMEM, double-buffered), O += P.V
//   warp 2      TMEM allocator
//   warps 4..7  softmax       one query row per thread: tcgen05.ld S -> running max / exp2 / sum,
//                             bf16 P -> swizzled smem (A operand of P.V), lazy rescale of O in TMEM,
//                             final O / l -> global
// Q.K^T of tile j+1 is issued before P.V of tile j so the tensor pipe works while the softmax
// warps are in their exp pass.
#include <cuda.h>
#include <math.h>

#include "host_common.h"
#include "sm100_ptx.cuh"

namespace b200 {

constexpr int kAttnBM = 128;   // query rows per CTA (TMEM lanes)
constexpr int kAttnBN = 128;   // keys per KV tile
constexpr int kAttnThreads = 256;
constexpr float kRescaleThreshold = 8.0f;   // log2 domain: P <= 2^8 between lazy rescales

template <int D>
struct AttnCfg {
  static constexpr int kQBytes = kAttnBM * D * 2;
  static constexpr int kKVBytes = kAttnBN * D * 2;
  static constexpr int kPBytes = kAttnBM * kAttnBN * 2;
  static constexpr int kStages = 2;
  static constexpr int kSmemBytes = kQBytes + 2 * kStages * kKVBytes + kPBytes + 1024 + 256;
  static constexpr int kTmemCols = 512;
  static constexpr int kSCol0 = 0, kSCol1 = 128, kOCol = 256;
};

struct AttnParams {
  __nv_bfloat16* O;
  int B, Sq, Sk, H;
  int64_t o_stride_b, o_stride_s, o_stride_h;
  float scale_log2;   // scale * log2(e)
  int causal;
  int num_q_tiles;
};

template <int D>
__global__ void __launch_bounds__(kAttnThreads, 1)
attention_fwd_kernel(const __grid_constant__ CUtensorMap tmQ, const __grid_constant__ CUtensorMap tmK,
                     const __grid_constant__ CUtensorMap tmV, const AttnParams p) {
  using Cfg = AttnCfg<D>;
  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));
  uint8_t* sQ = smem;
  uint8_t* sK = sQ + Cfg::kQBytes;
  uint8_t* sV = sK + Cfg::kStages * Cfg::kKVBytes;
  uint8_t* sP = sV + Cfg::kStages * Cfg::kKVBytes;
  uint64_t* bars = reinterpret_cast<uint64_t*>(sP + Cfg::kPBytes);
  uint64_t* q_full = bars;            // 1
  uint64_t* k_full = bars + 1;        // 2
  uint64_t* k_empty = bars + 3;       // 2
  uint64_t* v_full = bars + 5;        // 2
  uint64_t* v_empty = bars + 7;       // 2
  uint64_t* s_full = bars + 9;        // 2
  uint64_t* s_empty = bars + 11;      // 2
  uint64_t* p_full = bars + 13;       // 1
  uint64_t* o_done = bars + 14;       // 1
  uint32_t* tmem_ptr_smem = reinterpret_cast<uint32_t*>(bars + 15);

  const int warp = threadIdx.x >> 5;
  const int lane = threadIdx.x & 31;
  // heaviest causal tiles (largest f0) are scheduled first
  const int q_tile = p.causal ? (p.num_q_tiles - 1 - blockIdx.x) : blockIdx.x;
  const int head = blockIdx.y;
  const int batch = blockIdx.z;
  const int f0 = q_tile * kAttnBM;
  const int off = p.Sk - p.Sq;   // causal: key t visible to query f iff t <= f + off
  int kv_len = p.Sk;
  if (p.causal) kv_len = min(p.Sk, f0 + kAttnBM + off);
  const int n_tiles = (kv_len + kAttnBN - 1) / kAttnBN;

  if (warp == 0 && lane == 0) {
    tma_prefetch_desc(&tmQ);
    tma_prefetch_desc(&tmK);
    tma_prefetch_desc(&tmV);
  }
  if (warp == 1 && lane == 0) {
    mbar_init(q_full, 1);
    for (int s = 0; s < 2; ++s) {
      mbar_init(&k_full[s], 1);
      mbar_init(&k_empty[s], 1);
      mbar_init(&v_full[s], 1);
      mbar_init(&v_empty[s], 1);
      mbar_init(&s_full[s], 1);
      mbar_init(&s_empty[s], 4);
    }
    mbar_init(p_full, 4);
    mbar_init(o_done, 1);
    fence_mbar_init();
  }
  if (warp == 2) tmem_alloc<Cfg::kTmemCols>(tmem_ptr_smem);
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = *tmem_ptr_smem;

  if (warp == 0 && lane == 0) {
    // ===================== TMA producer =====================
    mbar_arrive_expect_tx(q_full, Cfg::kQBytes);
#pragma unroll
    for (int c = 0; c < D / 64; ++c) tma_load_4d(sQ + c * (kAttnBM * 128), &tmQ, q_full, c * 64, head, f0, batch);
    for (int j = 0; j < n_tiles; ++j) {
      const int st = j & 1;
      const uint32_t ph = (j >> 1) & 1;
      const int t0 = j * kAttnBN;
      mbar_wait(&k_empty[st], ph ^ 1);
      mbar_arrive_expect_tx(&k_full[st], Cfg::kKVBytes);
#pragma unroll
      for (int c = 0; c < D / 64; ++c)
        tma_load_4d(sK + st * Cfg::kKVBytes + c * (kAttnBN * 128), &tmK, &k_full[st], c * 64, head, t0, batch);
      mbar_wait(&v_empty[st], ph ^ 1);
      mbar_arrive_expect_tx(&v_full[st], Cfg::kKVBytes);
#pragma unroll
      for (int c = 0; c < D / 64; ++c)
        tma_load_4d(sV + st * Cfg::kKVBytes + c * (kAttnBN * 128), &tmV, &v_full[st], c * 64, head, t0, batch);
    }
  } else if (warp == 1 && lane == 0) {
    // ===================== MMA issuer =====================
    constexpr uint32_t idesc_qk = make_idesc_bf16(kAttnBM, kAttnBN, false, false);
    constexpr uint32_t idesc_pv = make_idesc_bf16(kAttnBM, D, false, true);
    const uint32_t q_addr = smem_u32(sQ);
    const uint32_t p_addr = smem_u32(sP);
    auto issue_qk = [&](int j) {
      const int st = j & 1;
      const uint32_t ph = (j >> 1) & 1;
      mbar_wait(&k_full[st], ph);
      mbar_wait(&s_empty[st], ph ^ 1);          // softmax has finished reading S[st] of tile j-2
      tc_fence_after();
      const uint32_t k_addr = smem_u32(sK + st * Cfg::kKVBytes);
      const uint32_t s_tmem = tmem_base + (st ? Cfg::kSCol1 : Cfg::kSCol0);
#pragma unroll
      for (int kk = 0; kk < D / 16; ++kk) {
        const uint32_t o = (kk >> 2) * (kAttnBM * 128) + (kk & 3) * 32;
        mma_bf16_ss(s_tmem, make_smem_desc_sw128(q_addr + o, 16, 1024), make_smem_desc_sw128(k_addr + o, 16, 1024),
                    idesc_qk, kk > 0 ? 1u : 0u);
      }
      mma_commit(&s_full[st]);
      mma_commit(&k_empty[st]);
    };
    mbar_wait(q_full, 0);
    issue_qk(0);
    for (int j = 0; j < n_tiles; ++j) {
      if (j + 1 < n_tiles) issue_qk(j + 1);
      const int st = j & 1;
      const uint32_t ph = (j >> 1) & 1;
      mbar_wait(&v_full[st], ph);
      mbar_wait(p_full, j & 1);                 // P of tile j is in smem, O has been rescaled
      tc_fence_after();
      const uint32_t v_addr = smem_u32(sV + st * Cfg::kKVBytes);
#pragma unroll
      for (int kk = 0; kk < kAttnBN / 16; ++kk) {
        const uint64_t da = make_smem_desc_sw128(p_addr + (kk >> 2) * (kAttnBM * 128) + (kk & 3) * 32, 16, 1024);
        // V tile: [128 keys][D] with D contiguous = MN-major B; 16 key-rows per MMA, 64-wide D chunks 16 KB apart
        const uint64_t db = make_smem_desc_sw128(v_addr + kk * (16 * 128), kAttnBN * 128, 1024);
        mma_bf16_ss(tmem_base + Cfg::kOCol, da, db, idesc_pv, (j > 0 || kk > 0) ? 1u : 0u);
      }
      mma_commit(&v_empty[st]);
      mma_commit(o_done);
    }
  } else if (warp >= 4) {
    // ===================== softmax / correction / epilogue =====================
    const int q = warp & 3;
    const int row = q * 32 + lane;              // query row inside the tile = TMEM lane
    const int f = f0 + row;
    const uint32_t lane_base = tmem_base + (static_cast<uint32_t>(q * 32) << 16);
    float m_used = -INFINITY;                   // running max (log2 domain) the stored P / O are scaled by
    float l_sum = 0.f;
    for (int j = 0; j < n_tiles; ++j) {
      const int st = j & 1;
      const uint32_t ph = (j >> 1) & 1;
      const int t0 = j * kAttnBN;
      const uint32_t s_tmem = lane_base + (st ? Cfg::kSCol1 : Cfg::kSCol0);
      // keys of this tile visible to this row: t < limit
      int limit = p.Sk - t0;
      if (p.causal) limit = min(limit, f + off + 1 - t0);
      const bool need_mask = limit < kAttnBN;
      mbar_wait(&s_full[st], ph);
      tc_fence_after();
      // ---- pass 1: row max
      float mx = -INFINITY;
#pragma unroll 1
      for (int c = 0; c < kAttnBN / 32; ++c) {
        uint32_t r[32];
        tmem_ld_32x32(s_tmem + c * 32, r);
        tmem_wait_ld();
        if (need_mask) {
#pragma unroll
          for (int i = 0; i < 32; ++i) {
            const float v = (c * 32 + i < limit) ? __uint_as_float(r[i]) : -INFINITY;
            mx = fmaxf(mx, v);
          }
        } else {
#pragma unroll
          for (int i = 0; i < 32; ++i) mx = fmaxf(mx, __uint_as_float(r[i]));
        }
      }
      const float m_tile = mx * p.scale_log2;
      float factor = 1.0f;
      bool rescale = false;
      if (j == 0) {
        m_used = m_tile;                        // tile 0 always has a visible key for every stored row
      } else if (m_tile > m_used + kRescaleThreshold) {
        factor = exp2_approx(m_used - m_tile);
        m_used = m_tile;
        rescale = true;
      }
      // P buffer free and O stable: P.V of tile j-1 has completed
      if (j > 0) {
        mbar_wait(o_done, (j - 1) & 1);
        tc_fence_after();
      }
      if (__any_sync(0xffffffffu, rescale)) {
        l_sum *= factor;
#pragma unroll 1
        for (int c = 0; c < D / 32; ++c) {
          uint32_t r[32];
          tmem_ld_32x32(lane_base + Cfg::kOCol + c * 32, r);
          tmem_wait_ld();
#pragma unroll
          for (int i = 0; i < 32; ++i) r[i] = __float_as_uint(__uint_as_float(r[i]) * factor);
          tmem_st_32x32(lane_base + Cfg::kOCol + c * 32, r);
        }
        tmem_wait_st();
      }
      // ---- pass 2: p = exp2(s * scale_log2 - m_used) -> bf16 -> swizzled smem; row sum
      const float neg_m = -m_used;
#pragma unroll 1
      for (int c = 0; c < kAttnBN / 32; ++c) {
        uint32_t r[32];
        tmem_ld_32x32(s_tmem + c * 32, r);
        tmem_wait_ld();
        uint32_t packed[16];
#pragma unroll
        for (int i = 0; i < 32; i += 2) {
          float p0 = exp2_approx(fmaf(__uint_as_float(r[i]), p.scale_log2, neg_m));
          float p1 = exp2_approx(fmaf(__uint_as_float(r[i + 1]), p.scale_log2, neg_m));
          if (need_mask) {
            if (c * 32 + i >= limit) p0 = 0.f;
            if (c * 32 + i + 1 >= limit) p1 = 0.f;
          }
          __nv_bfloat162 pb = __floats2bfloat162_rn(p0, p1);
          packed[i >> 1] = *reinterpret_cast<uint32_t*>(&pb);
          const float2 pf = __bfloat1622float2(pb);
          l_sum += pf.x + pf.y;                 // the denominator sums the same rounded P that multiplies V
        }
        // keys [c*32, c*32+32) of this row: four 16-byte units of the K-major 128B-swizzled P tile
        const int chunk = (c * 32) >> 6;
#pragma unroll
        for (int u4 = 0; u4 < 4; ++u4) {
          const int unit = ((c * 32) & 63) / 8 + u4;
          uint8_t* dst = sP + chunk * (kAttnBM * 128) + row * 128 + ((unit ^ (row & 7)) << 4);
          *reinterpret_cast<uint4*>(dst) =
              make_uint4(packed[u4 * 4], packed[u4 * 4 + 1], packed[u4 * 4 + 2], packed[u4 * 4 + 3]);
        }
      }
      fence_async_smem();                       // P visible to the tensor core's async proxy
      tc_fence_before();
      __syncwarp();
      if (lane == 0) {
        mbar_arrive(&s_empty[st]);              // S[st] may be overwritten by Q.K^T of tile j+2
        mbar_arrive(p_full);                    // P.V of tile j may be issued
      }
    }
    // ---- epilogue: O / l -> bf16 -> global [b, f, head, :]
    mbar_wait(o_done, (n_tiles - 1) & 1);
    tc_fence_after();
    const float inv_l = 1.0f / l_sum;
    __nv_bfloat16* dst = p.O + batch * p.o_stride_b + static_cast<int64_t>(f) * p.o_stride_s + head * p.o_stride_h;
#pragma unroll 1
    for (int c = 0; c < D / 32; ++c) {
      uint32_t r[32];
      tmem_ld_32x32(lane_base + Cfg::kOCol + c * 32, r);
      tmem_wait_ld();
      if (f < p.Sq) {
#pragma unroll
        for (int i = 0; i < 32; i += 8) {
          uint4 o;
          o.x = pack_bf16x2(__uint_as_float(r[i]) * inv_l, __uint_as_float(r[i + 1]) * inv_l);
          o.y = pack_bf16x2(__uint_as_float(r[i + 2]) * inv_l, __uint_as_float(r[i + 3]) * inv_l);
          o.z = pack_bf16x2(__uint_as_float(r[i + 4]) * inv_l, __uint_as_float(r[i + 5]) * inv_l);
          o.w = pack_bf16x2(__uint_as_float(r[i + 6]) * inv_l, __uint_as_float(r[i + 7]) * inv_l);
          *reinterpret_cast<uint4*>(dst + c * 32 + i) = o;
        }
      }
    }
  }

  __syncwarp();
  tc_fence_before();
  __syncthreads();
  if (warp == 2) {
    tc_fence_after();
    tmem_dealloc<Cfg::kTmemCols>(tmem_base);
  }
}

int make_bsnh_map(CUtensorMap* tm, const void* base, int B, int S, int H, int D, int64_t sb, int64_t ss,
                         int64_t sh, int rows) {
  uint64_t dims[4] = {static_cast<uint64_t>(D), static_cast<uint64_t>(H), static_cast<uint64_t>(S),
                      static_cast<uint64_t>(B)};
  uint64_t strides[4] = {0, static_cast<uint64_t>(sh) * 2, static_cast<uint64_t>(ss) * 2,
                         static_cast<uint64_t>(sb) * 2};
  uint32_t box[4] = {64, 1, static_cast<uint32_t>(rows), 1};
  return make_tensor_map(tm, base, 2, 4, dims, strides, box, 3);
}

template <int D>
static int launch_attention(const void* Q, const void* K, const void* V, const AttnParams& p, int B, int Sq, int Sk,
                            int H, const int64_t* qs, const int64_t* ks, const int64_t* vs, cudaStream_t stream) {
  using Cfg = AttnCfg<D>;
  CUtensorMap tmQ, tmK, tmV;
  int rc = make_bsnh_map(&tmQ, Q, B, Sq, H, D, qs[0], qs[1], qs[2], kAttnBM);
  if (rc != B200_OK) return rc;
  rc = make_bsnh_map(&tmK, K, B, Sk, H, D, ks[0], ks[1], ks[2], kAttnBN);
  if (rc != B200_OK) return rc;
  rc = make_bsnh_map(&tmV, V, B, Sk, H, D, vs[0], vs[1], vs[2], kAttnBN);
  if (rc != B200_OK) return rc;
  auto kern = attention_fwd_kernel<D>;
  B200_CUDA_CHECK(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, Cfg::kSmemBytes));
  dim3 grid(p.num_q_tiles, H, B);
  kern<<<grid, kAttnThreads, Cfg::kSmemBytes, stream>>>(tmQ, tmK, tmV, p);
  B200_CUDA_CHECK(cudaGetLastError());
  launch_counter().fetch_add(1);
  return B200_OK;
}

int attention_fwd8_dispatch(const void* Q, const void* K, const void* V, void* O, int B, int Sq, int Sk, int H, int D,
                            const int64_t* strides, float scale, int causal, float* lse, cudaStream_t stream);
int attention_fwd9_dispatch(const void* Q, const void* K, const void* V, void* O, int B, int Sq, int Sk, int H, int D,
                            const int64_t* strides, float scale, int causal, float* lse, cudaStream_t stream);

}  // namespace b200

using namespace b200;

// Kernels (b200_attention_args.variant): 1 = one 128-row query tile per CTA, blocking schedule (this file);
// 8 = persistent, one CTA per SM, two softmax warp groups on alternate KV tiles (attention8.cu);
// 9 = 8 on CTA pairs: cta_group::2 MMAs, K / V tiles loaded once per 256 query rows, head dim 128 only (attention9.cu).
// Measured predecessors (two query tiles per CTA, 64-key tiles, P through shared memory, ...) are archived under
// tools/variants/ and are not part of the library.
extern "C" int b200_attention_fwd_ex(const b200_attention_args* a, void* stream) {
  B200_REQUIRE(a != nullptr, "b200_attention_fwd: null argument block");
  const void *Q = a->Q, *K = a->K, *V = a->V;
  void* O = a->O;
  const int B = a->B, Sq = a->Sq, Sk = a->Sk, H = a->H, D = a->D, causal = a->causal, variant = a->variant;
  B200_REQUIRE(Q && K && V && O, "b200_attention_fwd: null operand pointer");
  B200_REQUIRE(B > 0 && Sq > 0 && Sk > 0 && H > 0, "b200_attention_fwd: B, Sq, Sk, H must be positive");
  B200_REQUIRE(D == 64 || D == 128, "b200_attention_fwd: head dim must be 64 or 128 (got %d)", D);
  B200_REQUIRE(B <= 65535 && H <= 65535, "b200_attention_fwd: B and H must be <= 65535");
  B200_REQUIRE(!causal || Sk >= Sq, "b200_attention_fwd: causal attention needs Sk >= Sq");
  B200_REQUIRE(variant == 0 || variant == 1 || variant == 8 || variant == 9,
               "b200_attention_fwd: variant must be 0 (auto), 1, 8 or 9 (got %d)", variant);
  B200_REQUIRE(variant != 9 || D == 128, "b200_attention_fwd: variant 9 (CTA pairs) needs head dim 128");
  B200_REQUIRE(variant != 1 || a->lse2 == nullptr,
               "b200_attention_fwd: lse2 is written by the persistent kernels (variant 0, 8, 9)");
  const int64_t* strides = a->strides;
  for (int i = 0; i < 12; ++i)
    B200_REQUIRE(strides[i] > 0 && strides[i] % 8 == 0, "b200_attention_fwd: strides must be positive multiples of 8");
  B200_REQUIRE((reinterpret_cast<uintptr_t>(Q) | reinterpret_cast<uintptr_t>(K) | reinterpret_cast<uintptr_t>(V) |
                reinterpret_cast<uintptr_t>(O)) % 16 == 0,
               "b200_attention_fwd: operands must be 16-byte aligned");
  cudaStream_t st = reinterpret_cast<cudaStream_t>(stream);
  const float scale = a->scale;
  // auto: measured best per head dim (profiles/r01_attention_v7.md): D = 128 -> CTA pairs (1223 TFLOP/s vs 1176 without
  // pairs), D = 64 -> persistent single CTAs; a single query tile without an lse request stays on the small kernel
  int v = variant;
  if (v == 0) v = (Sq > kAttnBM || a->lse2 != nullptr) ? (D == 128 ? 9 : 8) : 1;
  if (v == 9) return attention_fwd9_dispatch(Q, K, V, O, B, Sq, Sk, H, D, strides, scale, causal, a->lse2, st);
  if (v == 8) return attention_fwd8_dispatch(Q, K, V, O, B, Sq, Sk, H, D, strides, scale, causal, a->lse2, st);
  AttnParams p = {};
  p.O = reinterpret_cast<__nv_bfloat16*>(O);
  p.B = B;
  p.Sq = Sq;
  p.Sk = Sk;
  p.H = H;
  p.o_stride_b = strides[9];
  p.o_stride_s = strides[10];
  p.o_stride_h = strides[11];
  p.scale_log2 = scale * 1.4426950408889634f;
  p.causal = causal ? 1 : 0;
  p.num_q_tiles = (Sq + kAttnBM - 1) / kAttnBM;
  if (D == 64) return launch_attention<64>(Q, K, V, p, B, Sq, Sk, H, strides, strides + 3, strides + 6, st);
  return launch_attention<128>(Q, K, V, p, B, Sq, Sk, H, strides, strides + 3, strides + 6, st);
}

static int attention_fwd_flat(const void* Q, const void* K, const void* V, void* O, float* lse, int B, int Sq, int Sk, int H,
                              int D, int64_t q_stride_b, int64_t q_stride_s, int64_t q_stride_h, int64_t k_stride_b,
                              int64_t k_stride_s, int64_t k_stride_h, int64_t v_stride_b, int64_t v_stride_s,
                              int64_t v_stride_h, int64_t o_stride_b, int64_t o_stride_s, int64_t o_stride_h, float scale,
                              int causal, void* stream) {
  b200_attention_args a = {};
  a.Q = Q; a.K = K; a.V = V; a.O = O; a.lse2 = lse;
  a.B = B; a.Sq = Sq; a.Sk = Sk; a.H = H; a.D = D;
  const int64_t s[12] = {q_stride_b, q_stride_s, q_stride_h, k_stride_b, k_stride_s, k_stride_h,
                         v_stride_b, v_stride_s, v_stride_h, o_stride_b, o_stride_s, o_stride_h};
  for (int i = 0; i < 12; ++i) a.strides[i] = s[i];
  a.scale = scale;
  a.causal = causal;
  return b200_attention_fwd_ex(&a, stream);
}

extern "C" int b200_attention_fwd(const void* Q, const void* K, const void* V, void* O, int B, int Sq, int Sk, int H,
                                  int D, int64_t q_stride_b, int64_t q_stride_s, int64_t q_stride_h,
                                  int64_t k_stride_b, int64_t k_stride_s, int64_t k_stride_h, int64_t v_stride_b,
                                  int64_t v_stride_s, int64_t v_stride_h, int64_t o_stride_b, int64_t o_stride_s,
                                  int64_t o_stride_h, float scale, int causal, void* stream) {
  return attention_fwd_flat(Q, K, V, O, nullptr, B, Sq, Sk, H, D, q_stride_b, q_stride_s, q_stride_h, k_stride_b, k_stride_s,
                            k_stride_h, v_stride_b, v_stride_s, v_stride_h, o_stride_b, o_stride_s, o_stride_h, scale, causal,
                            stream);
}

// b200_attention_fwd that also returns, per (batch, head, query row), lse2 = log2(sum_t exp2(log2(e) * scale * s[t])) in
// lse2[B, H, Sq] (f32): what the backward needs to rebuild the probabilities without the running max.
extern "C" int b200_attention_fwd_lse(const void* Q, const void* K, const void* V, void* O, float* lse2, int B, int Sq, int Sk,
                                      int H, int D, int64_t q_stride_b, int64_t q_stride_s, int64_t q_stride_h,
                                      int64_t k_stride_b, int64_t k_stride_s, int64_t k_stride_h, int64_t v_stride_b,
                                      int64_t v_stride_s, int64_t v_stride_h, int64_t o_stride_b, int64_t o_stride_s,
                                      int64_t o_stride_h, float scale, int causal, void* stream) {
  B200_REQUIRE(lse2, "b200_attention_fwd_lse: null lse2");
  return attention_fwd_flat(Q, K, V, O, lse2, B, Sq, Sk, H, D, q_stride_b, q_stride_s, q_stride_h, k_stride_b, k_stride_s,
                            k_stride_h, v_stride_b, v_stride_s, v_stride_h, o_stride_b, o_stride_s, o_stride_h, scale, causal,
                            stream);
}
