// Attention forward, PERSISTENT: one CTA per SM walks a static list of (batch, head, query tile) work items, so that
// the ramp of one item (Q / K loads, first Q.K^T, first softmax) overlaps the tail of the previous one (last P.V,
// epilogue).  Inside an item it is attention7.cu: one 128-row query tile, S and P double-buffered in tensor memory,
// the two softmax warp groups on alternate KV tiles with a per-row hand-off of the running reference max.  Same math
// and layout as attention.cu (reference order case6_attention.py:121-135: f32 scores, f32 softmax, bf16 P,
// f32-accumulated P.V, bf16 out).
//
// Why (profiles/r01_attention_v7.md): with one CTA per (batch, head, query tile) and one CTA resident per SM (the
// kernel owns all 512 TMEM columns and ~224 KB of shared memory) nothing hides a CTA's start-up and drain: the
// clock64 profile of attention7.cu shows ~1600 cycles per KV tile inside the main loop but ~1970 per tile over the
// whole kernel at 32 tiles per CTA.
//
// Work list: non-causal: item = (batch, head, query tile), query tile fastest, so the CTAs running at the same time
// share K/V panels in L2.  Causal: item = (batch, head, pair of query tiles {n-1-i, i}): every item costs n+1 KV tiles,
// so the static round-robin over CTAs stays balanced.
//
// Buffers and barriers carry over from one item to the next; their phase is derived from T, the running count of KV
// tiles this CTA has processed, and from the running count of items:
//   S[T & 1], P[T & 1] and softmax group (T & 1) belong to tile T;
//   q_full / q_free / o_free flip once per item.
//
//   warp 0        TMA producer: Q of an item (after q_free of the previous one), then K_0,K_1,V_0,K_2,V_1,...
//   warp 1        MMA issuer: two cursors (Q.K^T up to two tiles ahead, P.V trailing), non-blocking polls
//   warp 2        TMEM allocator
//   warps 4..7    softmax group 0 (even T), warps 8..11 softmax group 1 (odd T); both groups share the epilogue
#include <cuda.h>
#include <math.h>

#include "host_common.h"
#include "sm100_ptx.cuh"

#ifndef B200_ATTN_LATE_PWAIT
#define B200_ATTN_LATE_PWAIT 0   // 1: wait for "P[b] read by P.V of tile T-2" right before the P store instead of before the exp
                                 // pass.  Measured 8-12 % SLOWER (interleaved A/B, tools/diag/attn_ab.py: 974 vs 1097 TFLOP/s at
                                 // B8 S4096 32x128, 572 vs 636 at head dim 64): the early wait paces the two softmax groups
#endif

namespace b200 {

constexpr int kA8BM = 128;      // query rows per item
constexpr int kA8BN = 128;      // keys per KV tile
constexpr int kA8Threads = 384;
constexpr float kA8Rescale = 8.0f;

template <int D>
struct Attn8Cfg {
  static constexpr int kQBytes = kA8BM * D * 2;
  static constexpr int kKVBytes = kA8BN * D * 2;
  // reference max [tile parity][row] + row sums [item parity][group][row]
  static constexpr int kXchgBytes = (2 + 4) * kA8BM * 4;
  static constexpr int kBudget = 227 * 1024 - 1024 - 512 - kXchgBytes - kQBytes;
  static constexpr int kStages = kBudget / kKVBytes > 8 ? 8 : kBudget / kKVBytes;
  static constexpr int kSmemBytes = kQBytes + kStages * kKVBytes + kXchgBytes + 1024 + 512;
  static constexpr int kTmemCols = 512;
  static constexpr int kSCol = 0;       // S[b] at 128 b
  static constexpr int kPCol = 256;     // P[b] at 256 + 64 b (bf16 pairs: 64 columns = 128 keys)
  static constexpr int kOCol = 384;     // O: D columns
};

struct Attn8Params {
  __nv_bfloat16* O;
  int Sq, Sk;
  int64_t o_stride_b, o_stride_s, o_stride_h;
  float scale_log2;
  int causal;
  int num_q_tiles;
  int H;
  int items_per_head;   // query tiles (non-causal) or pairs of query tiles (causal) per (batch, head)
  int total_items;      // items_per_head * H * B
  float* lse;           // optional [B, H, Sq] f32: log2 of the softmax denominator per row, scale included (for the backward)
};

// Walks this CTA's work list; every role runs the same walk, so the item / tile counters agree without communication.
struct Attn8Walk {
  const Attn8Params& p;
  int w;         // current work item
  int half;      // causal: 0 = heavy query tile of the pair, 1 = light one
  int q_tile, head, batch, nt;
  int s;         // ordinal of the current (sub-)item in this CTA
  int T0;        // KV tiles processed before it
  __device__ __forceinline__ Attn8Walk(const Attn8Params& pp) : p(pp), w(blockIdx.x), half(0), q_tile(0), head(0), batch(0), nt(0), s(-1), T0(0) {}
  __device__ __forceinline__ bool next() {
    for (;;) {
      if (w >= p.total_items) return false;
      const int i = w % p.items_per_head;
      const int hb = w / p.items_per_head;
      int qt;
      bool last_half = true;
      if (p.causal) {
        qt = half == 0 ? p.num_q_tiles - 1 - i : i;
        const bool single = (p.num_q_tiles - 1 - i) == i;     // odd tile count: the middle tile has no partner
        last_half = (half == 1) || single;
      } else {
        qt = i;
      }
      T0 += nt;                                               // tiles of the item just finished (0 the first time)
      ++s;
      q_tile = qt;
      head = hb % p.H;
      batch = hb / p.H;
      int kv_len = p.Sk;
      if (p.causal) kv_len = min(p.Sk, qt * kA8BM + kA8BM + (p.Sk - p.Sq));
      nt = (kv_len + kA8BN - 1) / kA8BN;
      if (last_half) { w += gridDim.x; half = 0; } else { half = 1; }
      return true;
    }
  }
};

__device__ __forceinline__ float max3m(float a, float b, float c) {
  float r;
  asm("max.f32 %0, %1, %2, %3;" : "=f"(r) : "f"(a), "f"(b), "f"(c));
  return r;
}
__device__ __forceinline__ void softmax_pair_barrier8() {     // the 256 softmax threads (warps 4..11)
  asm volatile("bar.sync 1, 256;" ::: "memory");
}

template <int D>
__global__ void __launch_bounds__(kA8Threads, 1)
attention_fwd8_kernel(const __grid_constant__ CUtensorMap tmQ, const __grid_constant__ CUtensorMap tmK,
                      const __grid_constant__ CUtensorMap tmV, const Attn8Params p) {
  using Cfg = Attn8Cfg<D>;
  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));
  uint8_t* sQ = smem;
  uint8_t* sKV = sQ + Cfg::kQBytes;
  float* mref = reinterpret_cast<float*>(sKV + Cfg::kStages * Cfg::kKVBytes);   // [tile parity][row]
  float* lsum_x = mref + 2 * kA8BM;                                             // [item parity][group][row]
  uint64_t* bars = reinterpret_cast<uint64_t*>(reinterpret_cast<uint8_t*>(mref) + Cfg::kXchgBytes);
  uint64_t* q_full = bars;                              // 1 (TMA bytes)
  uint64_t* q_free = bars + 1;                          // 1 (MMA commit after the last Q.K^T of an item)
  uint64_t* o_free = bars + 2;                          // 1 (8 warp arrivals: O of an item is in registers)
  uint64_t* kv_full = bars + 3;                         // kStages
  uint64_t* kv_empty = kv_full + Cfg::kStages;          // kStages
  uint64_t* s_full = kv_empty + Cfg::kStages;           // 2 (one MMA commit)
  uint64_t* s_free = s_full + 2;                        // 2 (4 warp arrivals: one group each)
  uint64_t* p_full = s_free + 2;                        // 2 (4 warp arrivals)
  uint64_t* pv_done = p_full + 2;                       // 2 (one MMA commit)
  uint64_t* m_ready = pv_done + 2;                      // [tile parity][quadrant]: reference max of a tile published
  uint32_t* tmem_ptr_smem = reinterpret_cast<uint32_t*>(m_ready + 8);

  const int warp = uniform_warp_idx();
  const int lane = threadIdx.x & 31;
  const int off = p.Sk - p.Sq;

  if (warp == 0 && lane == 0) {
    tma_prefetch_desc(&tmQ);
    tma_prefetch_desc(&tmK);
    tma_prefetch_desc(&tmV);
  }
  if (warp == 1 && lane == 0) {
    mbar_init(q_full, 1);
    mbar_init(q_free, 1);
    mbar_init(o_free, 8);
    for (int s = 0; s < Cfg::kStages; ++s) {
      mbar_init(&kv_full[s], 1);
      mbar_init(&kv_empty[s], 1);
    }
    for (int b = 0; b < 2; ++b) {
      mbar_init(&s_full[b], 1);
      mbar_init(&s_free[b], 4);
      mbar_init(&p_full[b], 4);
      mbar_init(&pv_done[b], 1);
    }
    for (int i = 0; i < 8; ++i) mbar_init(&m_ready[i], 1);
    fence_mbar_init();
  }
  if (warp == 2) tmem_alloc<Cfg::kTmemCols>(tmem_ptr_smem);
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = *tmem_ptr_smem;

  // register re-split: the control warp group needs few, each softmax thread keeps a 128-wide score row
  if (warp < 4) {
  asm volatile("setmaxnreg.dec.sync.aligned.u32 64;");
  if (warp == 0) {
    // ===================== TMA producer =====================
    if (elect_one()) {
      Attn8Walk wk(p);
      int item = 0;                              // K/V ring position, runs over all items
      while (wk.next()) {
        const int f0 = wk.q_tile * kA8BM;
        if (wk.s > 0) mbar_wait(q_free, (wk.s - 1) & 1);     // every Q.K^T of the previous item has read sQ
        mbar_arrive_expect_tx(q_full, Cfg::kQBytes);
#pragma unroll
        for (int c = 0; c < D / 64; ++c) tma_load_4d(sQ + c * (kA8BM * 128), &tmQ, q_full, c * 64, wk.head, f0, wk.batch);
        auto load = [&](const CUtensorMap* tm, int j) {
          const int st = item % Cfg::kStages;
          const uint32_t ph = (item / Cfg::kStages) & 1;
          ++item;
          mbar_wait(&kv_empty[st], ph ^ 1);
          mbar_arrive_expect_tx(&kv_full[st], Cfg::kKVBytes);
#pragma unroll
          for (int c = 0; c < D / 64; ++c)
            tma_load_4d(sKV + st * Cfg::kKVBytes + c * (kA8BN * 128), tm, &kv_full[st], c * 64, wk.head, j * kA8BN, wk.batch);
        };
        // K one tile ahead of V: K_0, K_1, V_0, K_2, V_1, ...  (2 nt loads per item)
        load(&tmK, 0);
        for (int j = 0; j < wk.nt; ++j) {
          if (j + 1 < wk.nt) load(&tmK, j + 1);
          load(&tmV, j);
        }
      }
    }
    __syncwarp();
  } else if (warp == 1) {
    // ===================== MMA issuer =====================
    constexpr uint32_t idesc_qk = make_idesc_bf16(kA8BM, kA8BN, false, false);
    constexpr uint32_t idesc_pv = make_idesc_bf16(kA8BM, D, false, true);
    const uint64_t q_desc0 = make_smem_desc_sw128(smem_u32(sQ), 16, 1024);
    const uint64_t k_desc0 = make_smem_desc_sw128(smem_u32(sKV), 16, 1024);
    const uint64_t v_desc0 = make_smem_desc_sw128(smem_u32(sKV), kA8BN * 128, 1024);
    // two cursors over the same walk: Q.K^T runs ahead (into the next item too), P.V trails
    Attn8Walk qk(p), pv(p);
    bool qk_ok = qk.next();
    bool pv_ok = pv.next();
    int qk_j = 0, pv_j = 0;
    uint32_t idle = 0;
    while (pv_ok) {
      bool progressed = false;
      // ---- O (+)= P(T) . V : first, it is what the softmax two tiles later and the epilogue wait for
      const int Tpv = pv.T0 + pv_j;
      const int Tqk = qk_ok ? qk.T0 + qk_j : 0x7fffffff;
      if (Tpv < Tqk) {
        const int b = Tpv & 1;
        const int i = 2 * pv.T0 + ((pv_j + 1 < pv.nt) ? 2 * pv_j + 2 : 2 * pv_j + 1), st = i % Cfg::kStages;   // V_j
        const uint32_t ph = (i / Cfg::kStages) & 1;
        // the first P.V of an item overwrites O: the epilogue of the previous item must have read it
        const bool o_ok = pv_j > 0 || pv.s == 0 || mbar_test_wait(o_free, (pv.s - 1) & 1);
        if (o_ok && mbar_test_wait(&p_full[b], (Tpv >> 1) & 1) && mbar_test_wait(&kv_full[st], ph)) {
          tc_fence_after();
          if (elect_one()) {
            const uint64_t vd = desc_advance(v_desc0, st * Cfg::kKVBytes);
            const uint32_t p_tmem = tmem_base + Cfg::kPCol + b * 64;     // 16 keys = 8 columns per MMA
            const uint32_t o_tmem = tmem_base + Cfg::kOCol;
            if (pv_j == 0) mma_bf16_ts_c<false>(o_tmem, p_tmem, vd, idesc_pv);
            else mma_bf16_ts_c<true>(o_tmem, p_tmem, vd, idesc_pv);
#pragma unroll
            for (int kk = 1; kk < kA8BN / 16; ++kk)
              mma_bf16_ts_c<true>(o_tmem, p_tmem + kk * 8, desc_advance(vd, kk * (16 * 128)), idesc_pv);
            mma_commit(&pv_done[b]);
            mma_commit(&kv_empty[st]);
          }
          __syncwarp();
          if (++pv_j == pv.nt) {
            pv_ok = pv.next();
            pv_j = 0;
          }
          progressed = true;
        }
      }
      // ---- S[T & 1] = Q . K_j^T, up to two tiles ahead of the softmax
      if (qk_ok) {
        const int T = qk.T0 + qk_j, b = T & 1;
        const int i = 2 * qk.T0 + (qk_j == 0 ? 0 : 2 * qk_j - 1), st = i % Cfg::kStages;                      // K_j
        const uint32_t ph = (i / Cfg::kStages) & 1;
        const bool q_ok = qk_j > 0 || mbar_test_wait(q_full, qk.s & 1);
        const bool s_ok = (T < 2) || mbar_test_wait(&s_free[b], ((T - 2) >> 1) & 1);
        if (q_ok && s_ok && mbar_test_wait(&kv_full[st], ph)) {
          tc_fence_after();
          const bool last = qk_j + 1 == qk.nt;
          if (elect_one()) {
            const uint64_t kd = desc_advance(k_desc0, st * Cfg::kKVBytes);
            const uint32_t s_tmem = tmem_base + Cfg::kSCol + b * 128;
            mma_bf16_ss_c<false>(s_tmem, q_desc0, kd, idesc_qk);
#pragma unroll
            for (int kk = 1; kk < D / 16; ++kk) {
              const uint32_t o = (kk >> 2) * (kA8BM * 128) + (kk & 3) * 32;
              mma_bf16_ss_c<true>(s_tmem, desc_advance(q_desc0, o), desc_advance(kd, o), idesc_qk);
            }
            mma_commit(&s_full[b]);
            mma_commit(&kv_empty[st]);
            if (last) mma_commit(q_free);
          }
          __syncwarp();
          if (++qk_j == qk.nt) {
            qk_ok = qk.next();
            qk_j = 0;
          }
          progressed = true;
        }
      }
#if B200_WATCHDOG
      if (progressed) {
        idle = 0;
      } else if (++idle > (1u << 26)) {
        if (lane == 0)
          printf("b200: attention8 MMA scheduler stuck block=%d qk=(%d,%d) pv=(%d,%d)\n", (int)blockIdx.x, qk.s, qk_j, pv.s, pv_j);
        __trap();
      }
#endif
    }
  }
  } else {
    asm volatile("setmaxnreg.inc.sync.aligned.u32 216;");
    // ===================== softmax / correction / epilogue =====================
    const int grp = (warp - 4) >> 2;             // tiles T with (T & 1) == grp; O columns [grp*D/2, +D/2) in the epilogue
    const int quad = warp & 3;                   // TMEM lane quadrant
    const int row = quad * 32 + lane;
    const int b = grp;                           // S[b], P[b]: this group's buffers
    const uint32_t lane_base = tmem_base + (static_cast<uint32_t>(quad * 32) << 16);
    const uint32_t s_tmem = lane_base + Cfg::kSCol + b * 128;
    const uint32_t p_tmem = lane_base + Cfg::kPCol + b * 64;
    const uint32_t o_all = lane_base + Cfg::kOCol;
    const uint64_t scale2 = pack_f32x2(p.scale_log2, p.scale_log2);
    Attn8Walk wk(p);
    while (wk.next()) {
      const int f = wk.q_tile * kA8BM + row;
      const int nt = wk.nt;
      float m_local = -INFINITY;                 // the reference max l_part is relative to
      float l_part = 0.f;                        // row sum over this group's tiles of the item
      for (int j = (grp - wk.T0) & 1; j < nt; j += 2) {
        const int T = wk.T0 + j;                 // (T & 1) == grp
        const uint32_t use_ph = (T >> 1) & 1;
        const int t0 = j * kA8BN;
        int limit = p.Sk - t0;
        if (p.causal) limit = min(limit, f + off + 1 - t0);
        const bool need_mask = limit < kA8BN;
        mbar_wait(&s_full[b], use_ph);
        tc_fence_after();
        uint32_t s[128];
        {
          uint32_t(&s0)[32] = *reinterpret_cast<uint32_t(*)[32]>(&s[0]);
          uint32_t(&s1)[32] = *reinterpret_cast<uint32_t(*)[32]>(&s[32]);
          uint32_t(&s2)[32] = *reinterpret_cast<uint32_t(*)[32]>(&s[64]);
          uint32_t(&s3)[32] = *reinterpret_cast<uint32_t(*)[32]>(&s[96]);
          tmem_ld_32x32(s_tmem, s0);
          tmem_ld_32x32(s_tmem + 32, s1);
          tmem_ld_32x32(s_tmem + 64, s2);
          tmem_ld_32x32(s_tmem + 96, s3);
          tmem_wait_ld();
        }
        tc_fence_before();
        __syncwarp();
        if (lane == 0) mbar_arrive(&s_free[b]);  // S[b] may be overwritten by Q.K^T of tile T+2
        if (__any_sync(0xffffffffu, need_mask)) {
          asm volatile("" ::: "memory");         // keep this a branch: only diagonal / tail tiles pay for masking
#pragma unroll
          for (int i = 0; i < 128; ++i)
            if (i >= limit) s[i] = 0xff800000u;  // -inf
        }
        float mx4[4] = {-INFINITY, -INFINITY, -INFINITY, -INFINITY};
#pragma unroll
        for (int i = 0; i < 128; i += 8) {
#pragma unroll
          for (int c = 0; c < 4; ++c)
            mx4[c] = max3m(mx4[c], __uint_as_float(s[i + 2 * c]), __uint_as_float(s[i + 2 * c + 1]));
        }
        const float m_tile = fmaxf(fmaxf(mx4[0], mx4[1]), fmaxf(mx4[2], mx4[3])) * p.scale_log2;
        // The reference max is ONE sequence over the tiles of an item: tile j starts from what tile j-1 (the other
        // group's thread of the same row) settled on.  That hand-off is the only coupling between the two groups.
        float m_prev = -INFINITY;
        if (j > 0) {
          mbar_wait(&m_ready[((T - 1) & 1) * 4 + quad], ((T - 1) >> 1) & 1);
          m_prev = mref[((T - 1) & 1) * kA8BM + row];
        }
        float m_used = m_prev;
        float factor = 1.0f;
        bool rescale = false;
        if (j == 0) {
          m_used = m_tile;                       // key 0 is visible to every stored row, so this is finite
        } else if (m_tile > m_prev + kA8Rescale) {
          factor = exp2_approx(m_prev - m_tile);
          m_used = m_tile;
          rescale = true;
        }
        mref[(T & 1) * kA8BM + row] = m_used;
        __syncwarp();
        if (lane == 0) mbar_arrive(&m_ready[(T & 1) * 4 + quad]);
        if (m_used != m_local) {                 // bring this group's partial row sum to the current reference
          l_part *= exp2_approx(m_local - m_used);
          m_local = m_used;
        }
        const bool any_rescale = __any_sync(0xffffffffu, rescale);
        if (any_rescale) {
          // O may only be touched once every P.V issued so far (tile T-1 is the last) has completed; P.V(T) cannot
          // start before this warp has delivered P(T) below, and a rescale by tile T+1 waits for P.V(T) in turn.
          mbar_wait(&pv_done[(T - 1) & 1], ((T - 1) >> 1) & 1);
          tc_fence_after();
#pragma unroll 1
          for (int c = 0; c < D / 32; ++c) {     // all D columns of this warp's 32 rows
            uint32_t r[32];
            tmem_ld_32x32(o_all + c * 32, r);
            tmem_wait_ld();
#pragma unroll
            for (int i = 0; i < 32; ++i) r[i] = __float_as_uint(__uint_as_float(r[i]) * factor);
            tmem_st_32x32(o_all + c * 32, r);
          }
          tmem_wait_st();
        }
#if !B200_ATTN_LATE_PWAIT
        else if (T >= 2) mbar_wait(&pv_done[b], ((T - 2) >> 1) & 1);   // P[b] has been read by P.V of tile T-2
#endif
        // (P[b] is written only after the exp pass: its "read by P.V of tile T-2" wait sits there, under the MUFU work)
        // p = exp2(s * scale_log2 - m_used) -> bf16 pairs in place -> TMEM P[b]
        const uint64_t negm2 = pack_f32x2(-m_used, -m_used);
        uint64_t sum2[2] = {0ull, 0ull};
#pragma unroll
        for (int i = 0; i < 64; ++i) {
          const uint64_t x2 = fma_f32x2(pack_f32x2(__uint_as_float(s[2 * i]), __uint_as_float(s[2 * i + 1])), scale2, negm2);
          float x0, x1;
          unpack_f32x2(x2, x0, x1);
          const float p0 = exp2_approx(x0), p1 = exp2_approx(x1);
          sum2[i & 1] = add_f32x2(sum2[i & 1], pack_f32x2(p0, p1));
          s[i] = pack_bf16x2(p0, p1);            // keys (2i, 2i+1): low half = even key
        }
        {
          float a0, a1, b0, b1;
          unpack_f32x2(sum2[0], a0, a1);
          unpack_f32x2(sum2[1], b0, b1);
          l_part += (a0 + a1) + (b0 + b1);
        }
        if (B200_ATTN_LATE_PWAIT && !any_rescale && T >= 2) {   // (the rescale path has already waited for P.V of tile T-1)
          mbar_wait(&pv_done[b], ((T - 2) >> 1) & 1);   // P[b] has been read by P.V of tile T-2
          tc_fence_after();
        }
        {
          const uint32_t(&pp0)[32] = *reinterpret_cast<const uint32_t(*)[32]>(&s[0]);
          const uint32_t(&pp1)[32] = *reinterpret_cast<const uint32_t(*)[32]>(&s[32]);
          tmem_st_32x32(p_tmem, pp0);
          tmem_st_32x32(p_tmem + 32, pp1);
          tmem_wait_st();
        }
        tc_fence_before();
        __syncwarp();
        if (lane == 0) mbar_arrive(&p_full[b]);
      }
      // ---- epilogue of the item: O / l -> bf16 -> global [batch, f, head, grp*D/2 ...]
      {
        const int Tl = wk.T0 + nt - 1;           // last tile of the item
        if ((Tl & 1) != grp) mbar_wait(&m_ready[(Tl & 1) * 4 + quad], (Tl >> 1) & 1);
        const float m_final = mref[(Tl & 1) * kA8BM + row];
        if (m_final != m_local) l_part *= exp2_approx(m_local - m_final);
        float* lx = lsum_x + (wk.s & 1) * 2 * kA8BM;
        lx[grp * kA8BM + row] = l_part;
        softmax_pair_barrier8();
        const float l_sum = l_part + lx[(grp ^ 1) * kA8BM + row];
        mbar_wait(&pv_done[Tl & 1], (Tl >> 1) & 1);
        tc_fence_after();
        const float inv_l = 1.0f / l_sum;
        if (p.lse != nullptr && grp == 0 && f < p.Sq)
          p.lse[(static_cast<int64_t>(wk.batch) * p.H + wk.head) * p.Sq + f] = m_final + log2f(l_sum);
        uint32_t r[D / 2];
#pragma unroll
        for (int c = 0; c < D / 64; ++c) {
          uint32_t(&rc)[32] = *reinterpret_cast<uint32_t(*)[32]>(&r[c * 32]);
          tmem_ld_32x32(o_all + grp * (D / 2) + c * 32, rc);
        }
        tmem_wait_ld();
        tc_fence_before();
        __syncwarp();
        if (lane == 0) mbar_arrive(o_free);      // O is in registers: the first P.V of the next item may overwrite it
        if (f < p.Sq) {
          __nv_bfloat16* dst = p.O + wk.batch * p.o_stride_b + static_cast<int64_t>(f) * p.o_stride_s +
                               wk.head * p.o_stride_h + grp * (D / 2);
#pragma unroll
          for (int i = 0; i < D / 2; i += 8) {
            uint4 o;
            o.x = pack_bf16x2(__uint_as_float(r[i]) * inv_l, __uint_as_float(r[i + 1]) * inv_l);
            o.y = pack_bf16x2(__uint_as_float(r[i + 2]) * inv_l, __uint_as_float(r[i + 3]) * inv_l);
            o.z = pack_bf16x2(__uint_as_float(r[i + 4]) * inv_l, __uint_as_float(r[i + 5]) * inv_l);
            o.w = pack_bf16x2(__uint_as_float(r[i + 6]) * inv_l, __uint_as_float(r[i + 7]) * inv_l);
            *reinterpret_cast<uint4*>(dst + i) = o;
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
static int launch_attention8(const void* Q, const void* K, const void* V, const Attn8Params& p, int B, int Sq, int Sk,
                             int H, const int64_t* qs, const int64_t* ks, const int64_t* vs, cudaStream_t stream) {
  using Cfg = Attn8Cfg<D>;
  static_assert(Cfg::kStages >= 3, "KV ring needs at least three stages");
  CUtensorMap tmQ, tmK, tmV;
  int rc = make_bsnh_map(&tmQ, Q, B, Sq, H, D, qs[0], qs[1], qs[2], kA8BM);
  if (rc != B200_OK) return rc;
  rc = make_bsnh_map(&tmK, K, B, Sk, H, D, ks[0], ks[1], ks[2], kA8BN);
  if (rc != B200_OK) return rc;
  rc = make_bsnh_map(&tmV, V, B, Sk, H, D, vs[0], vs[1], vs[2], kA8BN);
  if (rc != B200_OK) return rc;
  auto kern = attention_fwd8_kernel<D>;
  B200_CUDA_CHECK(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, Cfg::kSmemBytes));
  int dev = 0, sms = 0;
  B200_CUDA_CHECK(cudaGetDevice(&dev));
  B200_CUDA_CHECK(cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev));
  const int grid = p.total_items < sms ? p.total_items : sms;
  kern<<<grid, kA8Threads, Cfg::kSmemBytes, stream>>>(tmQ, tmK, tmV, p);
  B200_CUDA_CHECK(cudaGetLastError());
  launch_counter().fetch_add(1);
  return B200_OK;
}

int attention_fwd8_dispatch(const void* Q, const void* K, const void* V, void* O, int B, int Sq, int Sk, int H, int D,
                            const int64_t* strides, float scale, int causal, float* lse, cudaStream_t stream) {
  Attn8Params p = {};
  p.O = reinterpret_cast<__nv_bfloat16*>(O);
  p.Sq = Sq;
  p.Sk = Sk;
  p.o_stride_b = strides[9];
  p.o_stride_s = strides[10];
  p.o_stride_h = strides[11];
  p.scale_log2 = scale * 1.4426950408889634f;
  p.causal = causal ? 1 : 0;
  p.num_q_tiles = (Sq + kA8BM - 1) / kA8BM;
  p.H = H;
  p.items_per_head = p.causal ? (p.num_q_tiles + 1) / 2 : p.num_q_tiles;
  const long long total = static_cast<long long>(p.items_per_head) * H * B;
  if (total > 0x7fffffffLL) return fail(B200_ERR_INVALID_ARGUMENT, "b200_attention_fwd: too many work items");
  p.total_items = static_cast<int>(total);
  p.lse = lse;
  if (D == 64) return launch_attention8<64>(Q, K, V, p, B, Sq, Sk, H, strides, strides + 3, strides + 6, stream);
  return launch_attention8<128>(Q, K, V, p, B, Sq, Sk, H, strides, strides + 3, strides + 6, stream);
}

}  // namespace b200
