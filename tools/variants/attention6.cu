// Attention forward, v5 (attention5.cu: one 128-row query tile per CTA, rows split over two softmax warp
// groups, P in tensor memory) with Q ALSO in tensor memory: after the TMA has landed the query tile in smem the
// softmax warps copy it once into 64 TMEM columns, and Q.K^T takes it from there as the TMEM A operand.
// Same math and layout as attention.cu (reference order case6_attention.py:121-135).
//
// Why: with Q in smem a 128x128x16 Q.K^T MMA reads 8 KB of operands (Q 4 KB + K 4 KB) and runs at half rate on
// operand bandwidth (profiles/r01_smem_operand_bandwidth_model.md); v5's phase profile showed the tensor pipe
// saturated at ~2000 cycles per 128-key tile for that reason.  With Q in TMEM both MMAs read 4 KB from smem.
// TMEM budget: S[2] 2 x 128 + P 64 (single buffer) + Q D/2 + O D = 512 columns at D = 128.
//
//   warp 0        TMA producer: Q once, then K_0,K_1,V_0,K_2,V_1,... through one smem ring
//   warp 1        MMA issuer (whole warp polls, one elected lane issues); both MMAs take A from TMEM
//   warp 2        TMEM allocator
//   warps 4..7    softmax group A: keys [0,64) of every KV tile, O columns [0, D/2), Q columns [0, D/4)
//   warps 8..11   softmax group B: the other halves
#include <cuda.h>
#include <math.h>

#include "host_common.h"
#include "sm100_ptx.cuh"

namespace b200 {

#ifndef B200_ATTN_PROFILE
#define B200_ATTN_PROFILE 0   // 1: per-phase clock64 totals of one warp into Params::prof (tools/gpu_check.py attn_phases)
#endif

constexpr int kA6BM = 128;      // query rows per CTA
constexpr int kA6BN = 128;      // keys per KV tile
constexpr int kA6Threads = 384;
constexpr float kA6Rescale = 8.0f;
// Fraction of exp2 evaluations moved from MUFU to a degree-3 polynomial on the FMA pipe: every
// (1 << kA6PolyShift)-th pair, 0 = none.  Measured at B8 S4096 32x128: none 1073 TFLOP/s, every 4th pair 954,
// every 2nd pair 817 (profiles/r01_attention_v5.md) -- the packed-FMA sequence costs more issue slots than the
// MUFU time it frees, so it is off.
constexpr int kA6PolyShift = 2;   // used only when compiled with -DB200_ATTN5_POLY=1
#ifndef B200_ATTN5_POLY
#define B200_ATTN5_POLY 0
#endif

template <int D>
struct Attn6Cfg {
  static constexpr int kQBytes = kA6BM * D * 2;
  static constexpr int kKVBytes = kA6BN * D * 2;
  static constexpr int kXchgBytes = 2 * 2 * kA6BM * 4;   // [tile parity][group][row] partial maxima
  static constexpr int kBudget = 227 * 1024 - 1024 - 512 - kXchgBytes - kQBytes;
  static constexpr int kStages = kBudget / kKVBytes > 8 ? 8 : kBudget / kKVBytes;
  static constexpr int kSmemBytes = kQBytes + kStages * kKVBytes + kXchgBytes + 1024 + 512;
  static constexpr int kTmemCols = 512;
  static constexpr int kSCol = 0;       // S[b] at 128 b
  static constexpr int kPCol = 256;     // P (single buffer; bf16 pairs: 64 columns = 128 keys)
  static constexpr int kQCol = 320;     // Q: D/2 columns (bf16 pairs)
  static constexpr int kOCol = 384;     // O: D columns
};

struct Attn6Params {
  __nv_bfloat16* O;
  int Sq, Sk;
  int64_t o_stride_b, o_stride_s, o_stride_h;
  float scale_log2;
  int causal;
  int num_q_tiles;
  long long* prof;   // optional bring-up aid: per-phase cycle totals of block (0,0,0): softmax warp 4 -> [0..7], MMA warp -> [8..15]
};

__device__ __forceinline__ float max3k(float a, float b, float c) {
  float r;
  asm("max.f32 %0, %1, %2, %3;" : "=f"(r) : "f"(a), "f"(b), "f"(c));
  return r;
}
__device__ __forceinline__ void softmax_group_barrier6() {     // the 256 softmax threads (warps 4..11)
  asm volatile("bar.sync 1, 256;" ::: "memory");
}

template <int D>
__global__ void __launch_bounds__(kA6Threads, 1)
attention_fwd6_kernel(const __grid_constant__ CUtensorMap tmQ, const __grid_constant__ CUtensorMap tmK,
                      const __grid_constant__ CUtensorMap tmV, const Attn6Params p) {
  using Cfg = Attn6Cfg<D>;
  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));
  uint8_t* sQ = smem;
  uint8_t* sKV = sQ + Cfg::kQBytes;
  float* xchg = reinterpret_cast<float*>(sKV + Cfg::kStages * Cfg::kKVBytes);   // [2][2][128]
  uint64_t* bars = reinterpret_cast<uint64_t*>(reinterpret_cast<uint8_t*>(xchg) + Cfg::kXchgBytes);
  uint64_t* q_full = bars;                              // 1
  uint64_t* kv_full = bars + 1;                         // kStages
  uint64_t* kv_empty = kv_full + Cfg::kStages;          // kStages
  uint64_t* s_full = kv_empty + Cfg::kStages;           // 2 (one MMA commit)
  uint64_t* s_free = s_full + 2;                        // 2 (8 warp arrivals)
  uint64_t* p_full = s_free + 2;                        // 2 (8 warp arrivals)
  uint64_t* pv_done = p_full + 2;                       // 2 (one MMA commit)
  uint64_t* q_in_tmem = pv_done + 2;                    // 1 (8 warp arrivals)
  uint32_t* tmem_ptr_smem = reinterpret_cast<uint32_t*>(q_in_tmem + 1);

  const int warp = uniform_warp_idx();
  const int lane = threadIdx.x & 31;
  const int q_tile = p.causal ? (p.num_q_tiles - 1 - blockIdx.x) : blockIdx.x;   // heaviest first
  const int head = blockIdx.y;
  const int batch = blockIdx.z;
  const int f0 = q_tile * kA6BM;
  const int off = p.Sk - p.Sq;
  int kv_len = p.Sk;
  if (p.causal) kv_len = min(p.Sk, f0 + kA6BM + off);
  const int nt = (kv_len + kA6BN - 1) / kA6BN;

  if (warp == 0 && lane == 0) {
    tma_prefetch_desc(&tmQ);
    tma_prefetch_desc(&tmK);
    tma_prefetch_desc(&tmV);
  }
  if (warp == 1 && lane == 0) {
    mbar_init(q_full, 1);
    mbar_init(q_in_tmem, 8);
    for (int s = 0; s < Cfg::kStages; ++s) {
      mbar_init(&kv_full[s], 1);
      mbar_init(&kv_empty[s], 1);
    }
    for (int b = 0; b < 2; ++b) {
      mbar_init(&s_full[b], 1);
      mbar_init(&s_free[b], 8);
      mbar_init(&p_full[b], 8);
      mbar_init(&pv_done[b], 1);
    }
    fence_mbar_init();
  }
  if (warp == 2) tmem_alloc<Cfg::kTmemCols>(tmem_ptr_smem);
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = *tmem_ptr_smem;

  if (warp == 0) {
    // ===================== TMA producer =====================
    if (elect_one()) {
      mbar_arrive_expect_tx(q_full, Cfg::kQBytes);
#pragma unroll
      for (int c = 0; c < D / 64; ++c) tma_load_4d(sQ + c * (kA6BM * 128), &tmQ, q_full, c * 64, head, f0, batch);
      int item = 0;
      auto load = [&](const CUtensorMap* tm, int j) {
        const int st = item % Cfg::kStages;
        const uint32_t ph = (item / Cfg::kStages) & 1;
        ++item;
        mbar_wait(&kv_empty[st], ph ^ 1);
        mbar_arrive_expect_tx(&kv_full[st], Cfg::kKVBytes);
#pragma unroll
        for (int c = 0; c < D / 64; ++c)
          tma_load_4d(sKV + st * Cfg::kKVBytes + c * (kA6BN * 128), tm, &kv_full[st], c * 64, head, j * kA6BN, batch);
      };
      // K one tile ahead of V: K_0, K_1, V_0, K_2, V_1, ...
      if (nt > 0) load(&tmK, 0);
      for (int j = 0; j < nt; ++j) {
        if (j + 1 < nt) load(&tmK, j + 1);
        load(&tmV, j);
      }
    }
    __syncwarp();
  } else if (warp == 1) {
    // ===================== MMA issuer =====================
    constexpr uint32_t idesc_qk = make_idesc_bf16(kA6BM, kA6BN, false, false);
    constexpr uint32_t idesc_pv = make_idesc_bf16(kA6BM, D, false, true);
    const uint64_t k_desc0 = make_smem_desc_sw128(smem_u32(sKV), 16, 1024);
    const uint64_t v_desc0 = make_smem_desc_sw128(smem_u32(sKV), kA6BN * 128, 1024);
    int next_qk = 0, next_pv = 0;
    mbar_wait(q_in_tmem, 0);
    tc_fence_after();
    uint32_t idle = 0;
    const bool mprof = B200_ATTN_PROFILE && p.prof != nullptr && blockIdx.x == 0 && blockIdx.y == 0 && blockIdx.z == 0;
    long long m_t0 = mprof ? clock64() : 0, m_issue = 0, m_polls = 0;
    while (next_pv < nt) {
      bool progressed = false;
      // ---- O += P(j) . V_j : first, it is what the softmax two tiles later and the epilogue wait for
      if (next_pv < next_qk) {
        const int j = next_pv, b = j & 1;
        const int i = (j + 1 < nt) ? 2 * j + 2 : 2 * j + 1, st = i % Cfg::kStages;   // ring position of V_j
        const uint32_t ph = (i / Cfg::kStages) & 1;
        if (mbar_test_wait(&p_full[b], (j >> 1) & 1) && mbar_test_wait(&kv_full[st], ph)) {
          const long long ti = mprof ? clock64() : 0;
          tc_fence_after();
          if (elect_one()) {
            const uint64_t vd = desc_advance(v_desc0, st * Cfg::kKVBytes);
            const uint32_t p_tmem = tmem_base + Cfg::kPCol;              // 16 keys = 8 columns per MMA
            const uint32_t o_tmem = tmem_base + Cfg::kOCol;
            if (j == 0) mma_bf16_ts_c<false>(o_tmem, p_tmem, vd, idesc_pv);
            else mma_bf16_ts_c<true>(o_tmem, p_tmem, vd, idesc_pv);
#pragma unroll
            for (int kk = 1; kk < kA6BN / 16; ++kk)
              mma_bf16_ts_c<true>(o_tmem, p_tmem + kk * 8, desc_advance(vd, kk * (16 * 128)), idesc_pv);
            mma_commit(&pv_done[b]);
            mma_commit(&kv_empty[st]);
          }
          __syncwarp();
          if (mprof) m_issue += clock64() - ti;
          ++next_pv;
          progressed = true;
        }
      }
      // ---- S[b] = Q . K_j^T, up to two tiles ahead of the softmax
      if (next_qk < nt) {
        const int j = next_qk, b = j & 1;
        const int i = j == 0 ? 0 : 2 * j - 1, st = i % Cfg::kStages;                 // ring position of K_j
        const uint32_t ph = (i / Cfg::kStages) & 1;
        const bool s_ok = (j < 2) || mbar_test_wait(&s_free[b], ((j - 2) >> 1) & 1);
        if (s_ok && mbar_test_wait(&kv_full[st], ph)) {
          const long long ti = mprof ? clock64() : 0;
          tc_fence_after();
          if (elect_one()) {
            const uint64_t kd = desc_advance(k_desc0, st * Cfg::kKVBytes);
            const uint32_t s_tmem = tmem_base + Cfg::kSCol + b * 128;
            const uint32_t q_tmem = tmem_base + Cfg::kQCol;              // 16 head-dim elements = 8 columns per MMA
            mma_bf16_ts_c<false>(s_tmem, q_tmem, kd, idesc_qk);
#pragma unroll
            for (int kk = 1; kk < D / 16; ++kk)
              mma_bf16_ts_c<true>(s_tmem, q_tmem + kk * 8, desc_advance(kd, (kk >> 2) * (kA6BN * 128) + (kk & 3) * 32),
                                  idesc_qk);
            mma_commit(&s_full[b]);
            mma_commit(&kv_empty[st]);
          }
          __syncwarp();
          if (mprof) m_issue += clock64() - ti;
          ++next_qk;
          progressed = true;
        }
      }
      if (mprof) ++m_polls;
#if B200_WATCHDOG
      if (progressed) {
        idle = 0;
      } else if (++idle > (1u << 26)) {
        if (lane == 0)
          printf("b200: attention6 MMA scheduler stuck block=(%d,%d,%d) qk=%d pv=%d nt=%d\n", (int)blockIdx.x,
                 (int)blockIdx.y, (int)blockIdx.z, next_qk, next_pv, nt);
        __trap();
      }
#endif
    }
    if (mprof && lane == 0) {
      p.prof[8] = clock64() - m_t0;
      p.prof[9] = m_issue;
      p.prof[10] = m_polls;
    }
  } else if (warp >= 4) {
    // ===================== softmax / correction / epilogue =====================
    const int grp = (warp - 4) >> 2;             // 0: keys [0,64) and O columns [0,D/2); 1: the other halves
    const int quad = warp & 3;                   // TMEM lane quadrant
    const int row = quad * 32 + lane;
    const int f = f0 + row;
    const uint32_t lane_base = tmem_base + (static_cast<uint32_t>(quad * 32) << 16);
    const uint32_t o_tmem = lane_base + Cfg::kOCol + grp * (D / 2);
    const uint64_t scale2 = pack_f32x2(p.scale_log2, p.scale_log2);
    {
      // Q tile: smem (128B-swizzled, K-major) -> this thread's D/2 head-dim elements of its row -> TMEM
      mbar_wait(q_full, 0);
      constexpr int kUnits = D / 16;             // 16-byte units (8 bf16) per thread
      uint32_t qr[kUnits * 4];
      const int e0 = grp * (D / 2);              // first head-dim element of this thread
#pragma unroll
      for (int u = 0; u < kUnits; ++u) {
        const int e = e0 + u * 8;
        const uint8_t* src = sQ + (e >> 6) * (kA6BM * 128) + row * 128 + ((((e & 63) >> 3) ^ (row & 7)) << 4);
        const uint4 v = *reinterpret_cast<const uint4*>(src);
        qr[u * 4 + 0] = v.x;
        qr[u * 4 + 1] = v.y;
        qr[u * 4 + 2] = v.z;
        qr[u * 4 + 3] = v.w;
      }
      const uint32_t q_dst = lane_base + Cfg::kQCol + grp * (D / 4);
      if constexpr (D == 128) {
        tmem_st_32x32(q_dst, *reinterpret_cast<const uint32_t(*)[32]>(&qr[0]));
      } else {
        tmem_st_32x16(q_dst, *reinterpret_cast<const uint32_t(*)[16]>(&qr[0]));
      }
      tmem_wait_st();
      tc_fence_before();
      __syncwarp();
      if (lane == 0) mbar_arrive(q_in_tmem);
    }
    float m_used = -INFINITY;
    float l_part = 0.f;                          // sum over this thread's half of every row
    const bool prof = B200_ATTN_PROFILE && p.prof != nullptr && blockIdx.x == 0 && blockIdx.y == 0 && blockIdx.z == 0 && warp == 4 && lane == 0;
    long long tacc[8] = {0, 0, 0, 0, 0, 0, 0, 0};
    long long tp = prof ? clock64() : 0;
#if B200_ATTN_PROFILE
#define PROF_MARK6(k) do { if (prof) { long long _t = clock64(); tacc[k] += _t - tp; tp = _t; } } while (0)
#else
#define PROF_MARK6(k) do { } while (0)
#endif
    for (int j = 0; j < nt; ++j) {
      const int b = j & 1;
      const uint32_t use_ph = (j >> 1) & 1;
      const int t0 = j * kA6BN + grp * 64;       // first key of this thread's half tile
      int limit = p.Sk - t0;
      if (p.causal) limit = min(limit, f + off + 1 - t0);
      const bool need_mask = limit < 64;
      mbar_wait(&s_full[b], use_ph);
      tc_fence_after();
      PROF_MARK6(0);
      uint32_t s[64];
      {
        const uint32_t s_tmem = lane_base + Cfg::kSCol + b * 128 + grp * 64;
        uint32_t(&s0)[32] = *reinterpret_cast<uint32_t(*)[32]>(&s[0]);
        uint32_t(&s1)[32] = *reinterpret_cast<uint32_t(*)[32]>(&s[32]);
        tmem_ld_32x32(s_tmem, s0);
        tmem_ld_32x32(s_tmem + 32, s1);
        tmem_wait_ld();
      }
      tc_fence_before();
      __syncwarp();
      if (lane == 0) mbar_arrive(&s_free[b]);    // S[b] may be overwritten by Q.K^T of tile j+2
      PROF_MARK6(1);
      if (__any_sync(0xffffffffu, need_mask)) {
        asm volatile("" ::: "memory");           // keep this a branch: only diagonal / tail tiles pay for masking
#pragma unroll
        for (int i = 0; i < 64; ++i)
          if (i >= limit) s[i] = 0xff800000u;    // -inf
      }
      float mx4[4] = {-INFINITY, -INFINITY, -INFINITY, -INFINITY};
#pragma unroll
      for (int i = 0; i < 64; i += 8) {
#pragma unroll
        for (int c = 0; c < 4; ++c)
          mx4[c] = max3k(mx4[c], __uint_as_float(s[i + 2 * c]), __uint_as_float(s[i + 2 * c + 1]));
      }
      const float mx_half = fmaxf(fmaxf(mx4[0], mx4[1]), fmaxf(mx4[2], mx4[3]));
      // the row max needs the other half of the row: exchange through smem (slot by tile parity)
      PROF_MARK6(2);
      xchg[(b * 2 + grp) * kA6BM + row] = mx_half;
      softmax_group_barrier6();
      PROF_MARK6(3);
      const float mx = fmaxf(mx_half, xchg[(b * 2 + (grp ^ 1)) * kA6BM + row]);
      const float m_tile = mx * p.scale_log2;
      float factor = 1.0f;
      bool rescale = false;
      if (j == 0) {
        m_used = m_tile;                         // key 0 is visible to every stored row, so this is finite
      } else if (m_tile > m_used + kA6Rescale) {
        factor = exp2_approx(m_used - m_tile);
        m_used = m_tile;
        rescale = true;
      }
      if (__any_sync(0xffffffffu, rescale)) {
        // O may only be touched once every P.V issued so far (tile j-1 is the last) has completed
        mbar_wait(&pv_done[(j - 1) & 1], ((j - 1) >> 1) & 1);
        tc_fence_after();
        l_part *= factor;
#pragma unroll 1
        for (int c = 0; c < D / 64; ++c) {       // this group's D/2 columns of O
          uint32_t r[32];
          tmem_ld_32x32(o_tmem + c * 32, r);
          tmem_wait_ld();
#pragma unroll
          for (int i = 0; i < 32; ++i) r[i] = __float_as_uint(__uint_as_float(r[i]) * factor);
          tmem_st_32x32(o_tmem + c * 32, r);
        }
        tmem_wait_st();
      } else if (j >= 1) {
        mbar_wait(&pv_done[(j - 1) & 1], ((j - 1) >> 1) & 1);   // the single P buffer has been read by P.V of tile j-1
      }
      PROF_MARK6(4);
      // p = exp2(s * scale_log2 - m_used) -> bf16 pairs in place -> TMEM P[b], this group's 32 columns
      const uint64_t negm2 = pack_f32x2(-m_used, -m_used);
      uint64_t sum2[2] = {0ull, 0ull};
#if B200_ATTN5_POLY
      // MUFU.EX2 (16 per clock per SM) is what bounds this loop, so every other pair takes the FMA pipe
      // instead: 2^x = 2^n * 2^f with n = round(x) (magic-number add), f in [-0.5, 0.5], 2^f by a degree-3
      // polynomial (max rel. error 1.0e-4, far inside bf16's 2^-9), n added into the exponent bits.  Masked
      // tiles (x = -inf) keep the MUFU path for every element.
      const bool poly_ok = kA6PolyShift > 0 && !__any_sync(0xffffffffu, need_mask);
      const uint64_t magic2 = pack_f32x2(12582912.0f, 12582912.0f);          // 1.5 * 2^23
      const uint64_t nmagic2 = pack_f32x2(-12582912.0f, -12582912.0f);
      const uint64_t neg1 = pack_f32x2(-1.0f, -1.0f);
      const uint64_t c3 = pack_f32x2(0.05592203512787819f, 0.05592203512787819f);
      const uint64_t c2 = pack_f32x2(0.24264007806777954f, 0.24264007806777954f);
      const uint64_t c1 = pack_f32x2(0.6931210160255432f, 0.6931210160255432f);
      const uint64_t c0 = pack_f32x2(0.9999244809150696f, 0.9999244809150696f);
#endif
#pragma unroll
      for (int i = 0; i < 32; ++i) {
        const uint64_t x2 = fma_f32x2(pack_f32x2(__uint_as_float(s[2 * i]), __uint_as_float(s[2 * i + 1])), scale2, negm2);
        float p0, p1;
#if B200_ATTN5_POLY
        if (kA6PolyShift > 0 && (i & ((1 << kA6PolyShift) - 1)) == ((1 << kA6PolyShift) - 1) && poly_ok) {
          float x0, x1;
          unpack_f32x2(x2, x0, x1);
          const uint64_t xc = pack_f32x2(fmaxf(x0, -125.0f), fmaxf(x1, -125.0f));
          const uint64_t t2 = add_f32x2(xc, magic2);                         // n in the low mantissa bits
          const uint64_t f2 = fma_f32x2(add_f32x2(t2, nmagic2), neg1, xc);   // f = x - n
          uint64_t q2 = fma_f32x2(c3, f2, c2);
          q2 = fma_f32x2(q2, f2, c1);
          q2 = fma_f32x2(q2, f2, c0);
          float q0, q1, t0f, t1f;
          unpack_f32x2(q2, q0, q1);
          unpack_f32x2(t2, t0f, t1f);
          p0 = __int_as_float(__float_as_int(q0) + (__float_as_int(t0f) << 23));
          p1 = __int_as_float(__float_as_int(q1) + (__float_as_int(t1f) << 23));
        } else
#endif
        {
          float x0, x1;
          unpack_f32x2(x2, x0, x1);
          p0 = exp2_approx(x0);
          p1 = exp2_approx(x1);
        }
        sum2[i & 1] = add_f32x2(sum2[i & 1], pack_f32x2(p0, p1));
        s[i] = pack_bf16x2(p0, p1);              // keys (2i, 2i+1) of this half: low half = even key
      }
      {
        float a0, a1, b0, b1;
        unpack_f32x2(sum2[0], a0, a1);
        unpack_f32x2(sum2[1], b0, b1);
        l_part += (a0 + a1) + (b0 + b1);
      }
      PROF_MARK6(5);
      {
        const uint32_t(&pp)[32] = *reinterpret_cast<const uint32_t(*)[32]>(&s[0]);
        tmem_st_32x32(lane_base + Cfg::kPCol + grp * 32, pp);
        tmem_wait_st();
      }
      tc_fence_before();
      __syncwarp();
      if (lane == 0) mbar_arrive(&p_full[b]);
      PROF_MARK6(6);
    }
    if (prof) {
      for (int k = 0; k < 7; ++k) p.prof[k] = tacc[k];
      p.prof[7] = nt;
    }
    // ---- epilogue: O / l -> bf16 -> global [b, f, head, grp*D/2 ...]
    if (nt > 0) {
      // total row sum = both halves (slots of parity nt&1 are free: their last use was tile nt-2 or earlier)
      const int sl = nt & 1;
      xchg[(sl * 2 + grp) * kA6BM + row] = l_part;
      softmax_group_barrier6();
      const float l_sum = l_part + xchg[(sl * 2 + (grp ^ 1)) * kA6BM + row];
      mbar_wait(&pv_done[(nt - 1) & 1], ((nt - 1) >> 1) & 1);
      tc_fence_after();
      const float inv_l = 1.0f / l_sum;
      __nv_bfloat16* dst = p.O + batch * p.o_stride_b + static_cast<int64_t>(f) * p.o_stride_s + head * p.o_stride_h +
                           grp * (D / 2);
#pragma unroll 1
      for (int c = 0; c < D / 64; ++c) {
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
static int launch_attention6(const void* Q, const void* K, const void* V, const Attn6Params& p, int B, int Sq, int Sk,
                             int H, const int64_t* qs, const int64_t* ks, const int64_t* vs, cudaStream_t stream) {
  using Cfg = Attn6Cfg<D>;
  static_assert(Cfg::kStages >= 3, "KV ring needs at least three stages");
  CUtensorMap tmQ, tmK, tmV;
  int rc = make_bsnh_map(&tmQ, Q, B, Sq, H, D, qs[0], qs[1], qs[2], kA6BM);
  if (rc != B200_OK) return rc;
  rc = make_bsnh_map(&tmK, K, B, Sk, H, D, ks[0], ks[1], ks[2], kA6BN);
  if (rc != B200_OK) return rc;
  rc = make_bsnh_map(&tmV, V, B, Sk, H, D, vs[0], vs[1], vs[2], kA6BN);
  if (rc != B200_OK) return rc;
  auto kern = attention_fwd6_kernel<D>;
  B200_CUDA_CHECK(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, Cfg::kSmemBytes));
  dim3 grid(p.num_q_tiles, H, B);
  kern<<<grid, kA6Threads, Cfg::kSmemBytes, stream>>>(tmQ, tmK, tmV, p);
  B200_CUDA_CHECK(cudaGetLastError());
  launch_counter().fetch_add(1);
  return B200_OK;
}

extern long long* g_attn_prof;

int attention_fwd6_dispatch(const void* Q, const void* K, const void* V, void* O, int B, int Sq, int Sk, int H, int D,
                            const int64_t* strides, float scale, int causal, cudaStream_t stream) {
  Attn6Params p = {};
  p.O = reinterpret_cast<__nv_bfloat16*>(O);
  p.Sq = Sq;
  p.Sk = Sk;
  p.o_stride_b = strides[9];
  p.o_stride_s = strides[10];
  p.o_stride_h = strides[11];
  p.scale_log2 = scale * 1.4426950408889634f;
  p.causal = causal ? 1 : 0;
  p.num_q_tiles = (Sq + kA6BM - 1) / kA6BM;
  p.prof = g_attn_prof;
  if (D == 64) return launch_attention6<64>(Q, K, V, p, B, Sq, Sk, H, strides, strides + 3, strides + 6, stream);
  return launch_attention6<128>(Q, K, V, p, B, Sq, Sk, H, strides, strides + 3, strides + 6, stream);
}

}  // namespace b200
