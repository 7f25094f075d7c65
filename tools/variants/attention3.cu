// Attention forward, two query tiles per CTA, 64-key KV tiles, DOUBLE-BUFFERED S (TMEM) and P (smem):
// the variant used for long sequences.  Same math and layout as attention.cu / attention2.cu
// (reference order case6_attention.py:121-135).
//
// Why 64-key tiles: with 128-key tiles the smem only holds three K/V tiles and ONE P tile per query tile
// (D = 128), so every softmax step had to wait for the previous P.V to drain before it could write P, and
// the two query tiles ended up contending instead of alternating (profiles/r01_ncu_attention_v2*).  At
// 64 keys a P tile is 16 KB: two per query tile fit, S is double-buffered in TMEM as well (4 x 64 columns),
// the K/V ring is six tiles deep, and a softmax warp only ever waits for the tensor pipe when it is truly
// behind: Q.K^T runs up to two tiles ahead, P.V up to two tiles behind.
//
//   warp 0        TMA producer: both Q tiles once, then K_0,K_1,V_0,K_2,V_1,... through one smem ring
//   warp 1        MMA issuer: polls which of {Q.K^T(q,j), P.V(q,j)} is ready and issues it
//   warp 2        TMEM allocator (S: 2 tiles x 2 buffers x 64 columns; O: 2 x D columns)
//   warps 4..7    softmax / correction / epilogue of query tile 0 (one row per thread)
//   warps 8..11   the same for query tile 1
#include <cuda.h>
#include <math.h>

#include "host_common.h"
#include "sm100_ptx.cuh"

namespace b200 {

constexpr int kA3BM = 128;      // rows per query tile
constexpr int kA3BN = 64;       // keys per KV tile
constexpr int kA3Threads = 384;
constexpr float kA3Rescale = 8.0f;

template <int D>
struct Attn3Cfg {
  static constexpr int kQBytes = kA3BM * D * 2;          // one query tile
  static constexpr int kKVBytes = kA3BN * D * 2;         // one K or V tile
  static constexpr int kPBytes = kA3BM * kA3BN * 2;      // one P tile (128 rows x 128 B)
  static constexpr int kBudget = 227 * 1024 - 1024 - 512 - 2 * kQBytes - 4 * kPBytes;
  static constexpr int kStages = kBudget / kKVBytes > 8 ? 8 : kBudget / kKVBytes;
  static constexpr int kSmemBytes = 2 * kQBytes + 4 * kPBytes + kStages * kKVBytes + 1024 + 512;
  static constexpr int kTmemCols = 512;
  static constexpr int kSCol = 0;       // S[q][b] at kSCol + 128 q + 64 b
  static constexpr int kOCol = 256;     // O[q] at kOCol + 128 q
};

struct Attn3Params {
  __nv_bfloat16* O;
  int Sq, Sk;
  int64_t o_stride_b, o_stride_s, o_stride_h;
  float scale_log2;
  int causal;
  int num_q_blocks;
};

__device__ __forceinline__ float max3f(float a, float b, float c) {
  float r;
  asm("max.f32 %0, %1, %2, %3;" : "=f"(r) : "f"(a), "f"(b), "f"(c));
  return r;
}

template <int D>
__global__ void __launch_bounds__(kA3Threads, 1)
attention_fwd3_kernel(const __grid_constant__ CUtensorMap tmQ, const __grid_constant__ CUtensorMap tmK,
                      const __grid_constant__ CUtensorMap tmV, const Attn3Params p) {
  using Cfg = Attn3Cfg<D>;
  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));
  uint8_t* sQ = smem;                                   // [2][128][D]
  uint8_t* sP = sQ + 2 * Cfg::kQBytes;                  // [2 tiles][2 buffers][128][64]
  uint8_t* sKV = sP + 4 * Cfg::kPBytes;                 // ring
  uint64_t* bars = reinterpret_cast<uint64_t*>(sKV + Cfg::kStages * Cfg::kKVBytes);
  uint64_t* q_full = bars;                              // 1
  uint64_t* kv_full = bars + 1;                         // kStages
  uint64_t* kv_empty = kv_full + Cfg::kStages;          // kStages (2 arrivals: one per query tile)
  uint64_t* s_full = kv_empty + Cfg::kStages;           // [q][b] = 4
  uint64_t* s_free = s_full + 4;                        // 4 (4 warp arrivals each)
  uint64_t* p_full = s_free + 4;                        // 4 (4 warp arrivals each)
  uint64_t* pv_done = p_full + 4;                       // 4
  uint32_t* tmem_ptr_smem = reinterpret_cast<uint32_t*>(pv_done + 4);

  const int warp = uniform_warp_idx();
  const int lane = threadIdx.x & 31;
  const int q_block = p.causal ? (p.num_q_blocks - 1 - blockIdx.x) : blockIdx.x;   // heaviest first
  const int head = blockIdx.y;
  const int batch = blockIdx.z;
  const int f_base = q_block * 2 * kA3BM;
  const int off = p.Sk - p.Sq;
  // KV tiles each query tile needs (0 if the tile lies beyond Sq)
  int n_tiles[2];
#pragma unroll
  for (int q = 0; q < 2; ++q) {
    const int f0 = f_base + q * kA3BM;
    int kv_len = p.Sk;
    if (p.causal) kv_len = min(p.Sk, f0 + kA3BM + off);
    n_tiles[q] = f0 < p.Sq ? (kv_len + kA3BN - 1) / kA3BN : 0;
  }
  const int n_max = max(n_tiles[0], n_tiles[1]);

  if (warp == 0 && lane == 0) {
    tma_prefetch_desc(&tmQ);
    tma_prefetch_desc(&tmK);
    tma_prefetch_desc(&tmV);
  }
  if (warp == 1 && lane == 0) {
    mbar_init(q_full, 1);
    for (int s = 0; s < Cfg::kStages; ++s) {
      mbar_init(&kv_full[s], 1);
      mbar_init(&kv_empty[s], 2);
    }
    for (int i = 0; i < 4; ++i) {
      mbar_init(&s_full[i], 1);
      mbar_init(&s_free[i], 4);
      mbar_init(&p_full[i], 4);
      mbar_init(&pv_done[i], 1);
    }
    fence_mbar_init();
  }
  if (warp == 2) tmem_alloc<Cfg::kTmemCols>(tmem_ptr_smem);
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = *tmem_ptr_smem;

  // register re-split: the control warp group needs few, each softmax thread keeps a 128-wide score row
  if (warp < 4) {
  asm volatile("setmaxnreg.dec.sync.aligned.u32 56;");
  if (warp == 0 && lane == 0) {
    // ===================== TMA producer =====================
    mbar_arrive_expect_tx(q_full, 2 * Cfg::kQBytes);
#pragma unroll
    for (int q = 0; q < 2; ++q)
#pragma unroll
      for (int c = 0; c < D / 64; ++c)
        tma_load_4d(sQ + q * Cfg::kQBytes + c * (kA3BM * 128), &tmQ, q_full, c * 64, head, f_base + q * kA3BM, batch);
    // One ring, K running a tile ahead of V:  K_0, K_1, V_0, K_2, V_1, K_3, V_2, ...  K_{j+1} is wanted early
    // (Q.K^T of the next tile is issued as soon as the softmax has read S), V_j late (after the softmax of
    // tile j); in this order, with three slots, K_{j+2} reuses V_{j-1}'s slot and V_j reuses K_j's.
    int item = 0;
    auto load = [&](const CUtensorMap* tm, int j) {
      const int st = item % Cfg::kStages;
      const uint32_t ph = (item / Cfg::kStages) & 1;
      ++item;
      mbar_wait(&kv_empty[st], ph ^ 1);
      mbar_arrive_expect_tx(&kv_full[st], Cfg::kKVBytes);
#pragma unroll
      for (int c = 0; c < D / 64; ++c)
        tma_load_4d(sKV + st * Cfg::kKVBytes + c * (kA3BN * 128), tm, &kv_full[st], c * 64, head, j * kA3BN, batch);
    };
    if (n_max > 0) load(&tmK, 0);
    for (int j = 0; j < n_max; ++j) {
      if (j + 1 < n_max) load(&tmK, j + 1);
      load(&tmV, j);
    }
  } else if (warp == 1) {
    // ===================== MMA issuer (whole warp polls, one elected lane issues: operands stay in uniform registers) =====================
    constexpr uint32_t idesc_qk = make_idesc_bf16(kA3BM, kA3BN, false, false);
    constexpr uint32_t idesc_pv = make_idesc_bf16(kA3BM, D, false, true);
    int next_qk[2] = {0, 0};     // next KV tile whose Q.K^T is to be issued, per query tile
    int next_pv[2] = {0, 0};
    const uint64_t q_desc0 = make_smem_desc_sw128(smem_u32(sQ), 16, 1024);
    const uint64_t p_desc0 = make_smem_desc_sw128(smem_u32(sP), 16, 1024);
    const uint64_t k_desc0 = make_smem_desc_sw128(smem_u32(sKV), 16, 1024);
    const uint64_t v_desc0 = make_smem_desc_sw128(smem_u32(sKV), kA3BN * 128, 1024);
    mbar_wait(q_full, 0);
    int remaining = 2 * (n_tiles[0] + n_tiles[1]);
    uint32_t idle = 0;
    while (remaining > 0) {
      bool progressed = false;
#pragma unroll
      for (int q = 0; q < 2; ++q) {
        // ---- S[q] = Q_q . K_j^T
        if (next_qk[q] < n_tiles[q]) {
          const int j = next_qk[q];
          const int i = j == 0 ? 0 : 2 * j - 1, st = i % Cfg::kStages;      // ring position of K_j
          const uint32_t ph = (i / Cfg::kStages) & 1;
          const int b = j & 1;                   // S buffer; free once the softmax has read tile j-2 out of it
          const bool s_ok = (j < 2) || mbar_test_wait(&s_free[q * 2 + b], ((j - 2) >> 1) & 1);
          if (s_ok && mbar_test_wait(&kv_full[st], ph)) {
            tc_fence_after();
            if (elect_one()) {
            const uint64_t qd = desc_advance(q_desc0, q * Cfg::kQBytes);
            const uint64_t kd = desc_advance(k_desc0, st * Cfg::kKVBytes);
            const uint32_t s_tmem = tmem_base + Cfg::kSCol + q * 128 + b * 64;
            mma_bf16_ss_c<false>(s_tmem, qd, kd, idesc_qk);
#pragma unroll
            for (int kk = 1; kk < D / 16; ++kk)
              mma_bf16_ss_c<true>(s_tmem, desc_advance(qd, (kk >> 2) * (kA3BM * 128) + (kk & 3) * 32),
                                  desc_advance(kd, (kk >> 2) * (kA3BN * 128) + (kk & 3) * 32), idesc_qk);
            mma_commit(&s_full[q * 2 + b]);
            mma_commit(&kv_empty[st]);
            if (j >= n_tiles[q ^ 1]) mbar_arrive(&kv_empty[st]);   // the other tile never reads K_j
            }
            __syncwarp();
            ++next_qk[q];
            --remaining;
            progressed = true;
          }
        }
        // ---- O[q] += P_q . V_j
        if (next_pv[q] < next_qk[q]) {
          const int j = next_pv[q];
          const int i = (j + 1 < n_max) ? 2 * j + 2 : 2 * j + 1, st = i % Cfg::kStages;   // ring position of V_j
          const uint32_t ph = (i / Cfg::kStages) & 1;
          const int b = j & 1;
          if (mbar_test_wait(&p_full[q * 2 + b], (j >> 1) & 1) && mbar_test_wait(&kv_full[st], ph)) {
            tc_fence_after();
            if (elect_one()) {
            const uint64_t pd = desc_advance(p_desc0, (q * 2 + b) * Cfg::kPBytes);
            // V tile: [64 keys][D], D contiguous = MN-major B; 16 key-rows per MMA, 64-wide D chunks 8 KB apart
            const uint64_t vd = desc_advance(v_desc0, st * Cfg::kKVBytes);
            const uint32_t o_tmem = tmem_base + Cfg::kOCol + q * 128;
            if (j == 0) mma_bf16_ss_c<false>(o_tmem, pd, vd, idesc_pv);
            else mma_bf16_ss_c<true>(o_tmem, pd, vd, idesc_pv);
#pragma unroll
            for (int kk = 1; kk < kA3BN / 16; ++kk)
              mma_bf16_ss_c<true>(o_tmem, desc_advance(pd, kk * 32), desc_advance(vd, kk * (16 * 128)), idesc_pv);
            mma_commit(&pv_done[q * 2 + b]);
            mma_commit(&kv_empty[st]);
            if (j >= n_tiles[q ^ 1]) mbar_arrive(&kv_empty[st]);
            }
            __syncwarp();
            ++next_pv[q];
            --remaining;
            progressed = true;
          }
        }
      }
#if B200_WATCHDOG
      if (progressed) {
        idle = 0;
      } else if (++idle > (1u << 26)) {
        printf("b200: attention MMA scheduler stuck block=(%d,%d,%d) qk=(%d,%d) pv=(%d,%d)\n", (int)blockIdx.x,
               (int)blockIdx.y, (int)blockIdx.z, next_qk[0], next_qk[1], next_pv[0], next_pv[1]);
        __trap();
      }
#endif
    }
  }
  } else {
    asm volatile("setmaxnreg.inc.sync.aligned.u32 208;");
    // ===================== softmax / correction / epilogue =====================
    const int q = (warp - 4) >> 2;               // query tile of this warp group
    const int quad = warp & 3;                   // TMEM lane quadrant
    const int row = quad * 32 + lane;
    const int f = f_base + q * kA3BM + row;
    const int nt = n_tiles[q];
    const uint32_t lane_base = tmem_base + (static_cast<uint32_t>(quad * 32) << 16);
    const uint32_t s_tmem_q = lane_base + Cfg::kSCol + q * 128;
    const uint32_t o_tmem = lane_base + Cfg::kOCol + q * 128;
    float m_used = -INFINITY;
    float l_sum = 0.f;
    const uint64_t scale2 = pack_f32x2(p.scale_log2, p.scale_log2);
    const uint32_t r7x16 = (row & 7) << 4;
    for (int j = 0; j < nt; ++j) {
      const int b = j & 1;                          // S / P buffer of this tile
      const uint32_t use_ph = (j >> 1) & 1;         // phase of this use of buffer b
      const int t0 = j * kA3BN;
      int limit = p.Sk - t0;
      if (p.causal) limit = min(limit, f + off + 1 - t0);
      const bool need_mask = limit < kA3BN;
      mbar_wait(&s_full[q * 2 + b], use_ph);
      tc_fence_after();
      // the 64-wide score row -> registers, then S[q][b] is free for Q.K^T of tile j+2
      uint32_t s[64];
      {
        uint32_t(&s0)[32] = *reinterpret_cast<uint32_t(*)[32]>(&s[0]);
        uint32_t(&s1)[32] = *reinterpret_cast<uint32_t(*)[32]>(&s[32]);
        tmem_ld_32x32(s_tmem_q + b * 64, s0);
        tmem_ld_32x32(s_tmem_q + b * 64 + 32, s1);
        tmem_wait_ld();
      }
      tc_fence_before();
      __syncwarp();
      if (lane == 0) mbar_arrive(&s_free[q * 2 + b]);
      if (__any_sync(0xffffffffu, need_mask)) {     // warp-uniform: only diagonal / tail tiles pay for masking
        asm volatile("" ::: "memory");               // keep this a branch, not predicated selects on every tile
#pragma unroll
        for (int i = 0; i < 64; ++i)
          if (i >= limit) s[i] = 0xff800000u;        // -inf
      }
      float mx4[4] = {-INFINITY, -INFINITY, -INFINITY, -INFINITY};
#pragma unroll
      for (int i = 0; i < 64; i += 8) {
#pragma unroll
        for (int c = 0; c < 4; ++c)
          mx4[c] = max3f(mx4[c], __uint_as_float(s[i + 2 * c]), __uint_as_float(s[i + 2 * c + 1]));
      }
      const float mx = fmaxf(fmaxf(mx4[0], mx4[1]), fmaxf(mx4[2], mx4[3]));
      const float m_tile = mx * p.scale_log2;
      float factor = 1.0f;
      bool rescale = false;
      if (j == 0) {
        m_used = m_tile;
      } else if (m_tile > m_used + kA3Rescale) {
        factor = exp2_approx(m_used - m_tile);
        m_used = m_tile;
        rescale = true;
      }
      if (__any_sync(0xffffffffu, rescale)) {
        // O[q] may only be touched once every P.V issued so far (tile j-1 is the last) has completed
        mbar_wait(&pv_done[q * 2 + ((j - 1) & 1)], ((j - 1) >> 1) & 1);
        tc_fence_after();
        l_sum *= factor;
#pragma unroll 1
        for (int c = 0; c < D / 32; ++c) {
          uint32_t r[32];
          tmem_ld_32x32(o_tmem + c * 32, r);
          tmem_wait_ld();
#pragma unroll
          for (int i = 0; i < 32; ++i) r[i] = __float_as_uint(__uint_as_float(r[i]) * factor);
          tmem_st_32x32(o_tmem + c * 32, r);
        }
        tmem_wait_st();
      } else if (j >= 2) {
        // P[q][b] is free once P.V of tile j-2 has read it
        mbar_wait(&pv_done[q * 2 + b], ((j - 2) >> 1) & 1);
      }
      // p = exp2(s * scale_log2 - m_used): packed f32x2 FMA / ADD, one MUFU per element, bf16x2 pack
      const uint64_t negm2 = pack_f32x2(-m_used, -m_used);
      uint64_t sum2[2] = {0ull, 0ull};
      uint8_t* prow = sP + (q * 2 + b) * Cfg::kPBytes + row * 128;
#pragma unroll
      for (int u = 0; u < 8; ++u) {             // 16-byte unit u = keys [8u, 8u+8)
        uint32_t pk[4];
#pragma unroll
        for (int e = 0; e < 4; ++e) {
          const uint64_t x2 = fma_f32x2(pack_f32x2(__uint_as_float(s[u * 8 + 2 * e]), __uint_as_float(s[u * 8 + 2 * e + 1])),
                                        scale2, negm2);
          float x0, x1;
          unpack_f32x2(x2, x0, x1);
          const float p0 = exp2_approx(x0), p1 = exp2_approx(x1);
          sum2[e & 1] = add_f32x2(sum2[e & 1], pack_f32x2(p0, p1));
          pk[e] = pack_bf16x2(p0, p1);
        }
        *reinterpret_cast<uint4*>(prow + ((u << 4) ^ r7x16)) = make_uint4(pk[0], pk[1], pk[2], pk[3]);
      }
      {
        float a0, a1, b0, b1;
        unpack_f32x2(sum2[0], a0, a1);
        unpack_f32x2(sum2[1], b0, b1);
        l_sum += (a0 + a1) + (b0 + b1);
      }
      fence_async_smem();
      tc_fence_before();
      __syncwarp();
      if (lane == 0) mbar_arrive(&p_full[q * 2 + b]);
    }
    if (nt > 0) {
      mbar_wait(&pv_done[q * 2 + ((nt - 1) & 1)], ((nt - 1) >> 1) & 1);
      tc_fence_after();
      const float inv_l = 1.0f / l_sum;
      __nv_bfloat16* dst = p.O + batch * p.o_stride_b + static_cast<int64_t>(f) * p.o_stride_s + head * p.o_stride_h;
#pragma unroll 1
      for (int c = 0; c < D / 32; ++c) {
        uint32_t r[32];
        tmem_ld_32x32(o_tmem + c * 32, r);
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
  }

  __syncwarp();
  tc_fence_before();
  __syncthreads();
  if (warp == 2) {
    tc_fence_after();
    tmem_dealloc<Cfg::kTmemCols>(tmem_base);
  }
}

int make_bsnh_map(CUtensorMap* tm, const void* base, int B, int S, int H, int D, int64_t sb, int64_t ss, int64_t sh,
                  int rows);

template <int D>
static int launch_attention3(const void* Q, const void* K, const void* V, const Attn3Params& p, int B, int Sq, int Sk,
                             int H, const int64_t* qs, const int64_t* ks, const int64_t* vs, cudaStream_t stream) {
  using Cfg = Attn3Cfg<D>;
  static_assert(Cfg::kStages >= 3, "KV ring needs at least three stages");
  CUtensorMap tmQ, tmK, tmV;
  int rc = make_bsnh_map(&tmQ, Q, B, Sq, H, D, qs[0], qs[1], qs[2], kA3BM);
  if (rc != B200_OK) return rc;
  rc = make_bsnh_map(&tmK, K, B, Sk, H, D, ks[0], ks[1], ks[2], kA3BN);
  if (rc != B200_OK) return rc;
  rc = make_bsnh_map(&tmV, V, B, Sk, H, D, vs[0], vs[1], vs[2], kA3BN);
  if (rc != B200_OK) return rc;
  auto kern = attention_fwd3_kernel<D>;
  B200_CUDA_CHECK(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, Cfg::kSmemBytes));
  dim3 grid(p.num_q_blocks, H, B);
  kern<<<grid, kA3Threads, Cfg::kSmemBytes, stream>>>(tmQ, tmK, tmV, p);
  B200_CUDA_CHECK(cudaGetLastError());
  launch_counter().fetch_add(1);
  return B200_OK;
}

int attention_fwd3_dispatch(const void* Q, const void* K, const void* V, void* O, int B, int Sq, int Sk, int H, int D,
                            const int64_t* strides, float scale, int causal, cudaStream_t stream) {
  Attn3Params p = {};
  p.O = reinterpret_cast<__nv_bfloat16*>(O);
  p.Sq = Sq;
  p.Sk = Sk;
  p.o_stride_b = strides[9];
  p.o_stride_s = strides[10];
  p.o_stride_h = strides[11];
  p.scale_log2 = scale * 1.4426950408889634f;
  p.causal = causal ? 1 : 0;
  p.num_q_blocks = (Sq + 2 * kA3BM - 1) / (2 * kA3BM);
  if (D == 64) return launch_attention3<64>(Q, K, V, p, B, Sq, Sk, H, strides, strides + 3, strides + 6, stream);
  return launch_attention3<128>(Q, K, V, p, B, Sq, Sk, H, strides, strides + 3, strides + 6, stream);
}

}  // namespace b200
