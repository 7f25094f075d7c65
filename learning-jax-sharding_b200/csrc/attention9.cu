// Attention forward on CTA PAIRS (cluster of 2, tcgen05 cta_group::2), persistent, head dim 128.  A pair owns a
// 256-row query block: CTA r holds query tile 2i + r, its own S / P / O in tensor memory and its own softmax warps, exactly
// as in attention8.cu; what the pair shares is K and V: every KV tile is loaded ONCE per pair, half into each CTA's
// shared memory (Q.K^T: CTA r holds keys [64 r, 64 r + 64) of the tile, the N split of the M = 256 MMA; P.V: CTA r holds
// V[:, 64 r .. 64 r + 64), the N split over the head dim).  Same math and layout as attention.cu (reference order
// case6_attention.py:121-135: f32 scores, f32 softmax, bf16 P, f32-accumulated P.V, bf16 out).
//
// Why (profiles/r01_attention_v7.md): with one query tile per CTA every CTA streams the whole K and V of its head
// through TMA, 64 KB per 128 keys; at ~1400-1600 cycles per KV tile that is ~6-7 KB per clock over the chip, which is
// the measured L2 -> SM throughput limit (the per-tile time of attention7/8 stays ~1950 cycles although their main loops
// run at ~1600 in a lightly loaded first wave).  A pair halves the bytes per query row, and the M = 256 MMA also reads
// half the B operand bytes per CTA.
//
// Work list (static round-robin over pairs): non-causal: item = (batch, head, query block); causal: item = (batch, head,
// two query blocks {n-1-i, i}) so that every item costs about the same number of KV tiles.  For a causal block the KV
// range is the later tile's; the earlier tile sees its last KV tile fully masked.
//
// Barriers: full / "P ready" / "S free" / "O free" barriers live in the leader CTA (rank 0), whose warp 1 issues every
// MMA; the other CTA's TMA bytes and warp arrivals are sent there.  Completion barriers (S full, P.V done, ring slot
// empty, Q free) exist in both CTAs and are signalled by multicast tcgen05.commit.
//
//   warp 0        TMA producer (both CTAs): own Q tile, own halves of K_j / V_j
//   warp 1        MMA issuer (leader only): two cursors (Q.K^T up to two tiles ahead, P.V trailing), non-blocking polls
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

#ifndef B200_ATTN_PROFILE
#define B200_ATTN_PROFILE 0   // 1: per-phase clock64 totals of one warp into Params::prof (tools/gpu_check.py attn_phases)
#endif

#ifndef B200_ATTN9_QK_FIRST
#define B200_ATTN9_QK_FIRST 1   // MMA issue order when both are ready: Q.K^T (what the softmax waits for next) before P.V
#endif

#ifndef B200_ATTN9_OVERLAP
#define B200_ATTN9_OVERLAP 0    // (not compatible with the late P-buffer wait: an r01 experiment, kept for the record)
#endif
#if B200_ATTN9_OVERLAP
#error "B200_ATTN9_OVERLAP stores half of P before the P-buffer wait: rebuild r01's attention9.cu to repeat that A/B"
#endif
// (1 = second half of the S load under the max of the first, first half of the P store under the exp of the second; A/B in
//  profiles/r01_attention_v9_ab.log: no gain)
#ifndef B200_ATTN9_POLY
#define B200_ATTN9_POLY 0       // n > 0: every (1 << n)-th pair of exp2 goes to a degree-3 polynomial on the FMA pipe
#endif

constexpr int kA9BM = 128;      // query rows per CTA (256 per pair)
constexpr int kA9BN = 128;      // keys per KV tile
constexpr int kA9Threads = 384;
constexpr float kA9Rescale = 8.0f;

template <int D>
struct Attn9Cfg {
  static constexpr int kQBytes = kA9BM * D * 2;
  static constexpr int kKVBytes = kA9BN * D * 2 / 2;     // this CTA's half of a K or V tile
  // reference max [tile parity][row] + row sums [item parity][group][row]
  static constexpr int kXchgBytes = (2 + 4) * kA9BM * 4;
  static constexpr int kBudget = 227 * 1024 - 1024 - 512 - kXchgBytes - kQBytes;
  static constexpr int kStages = kBudget / kKVBytes > 8 ? 8 : kBudget / kKVBytes;
  static constexpr int kSmemBytes = kQBytes + kStages * kKVBytes + kXchgBytes + 1024 + 512;
  static constexpr int kTmemCols = 512;
  static constexpr int kSCol = 0;       // S[b] at 128 b
  static constexpr int kPCol = 256;     // P[b] at 256 + 64 b (bf16 pairs: 64 columns = 128 keys)
  static constexpr int kOCol = 384;     // O: D columns
};

struct Attn9Params {
  __nv_bfloat16* O;
  int Sq, Sk;
  int64_t o_stride_b, o_stride_s, o_stride_h;
  float scale_log2;
  int causal;
  int num_q_tiles;
  int H;
  int num_q_blocks;     // 256-row query blocks per (batch, head)
  int items_per_head;   // query blocks (non-causal) or pairs of query blocks (causal) per (batch, head)
  int total_items;      // items_per_head * H * B
  float* lse;           // optional [B, H, Sq] f32: log2 of the softmax denominator per row, scale included (for the backward)
  long long* prof;      // bring-up aid (B200_ATTN_PROFILE builds): softmax warp 4 of block 0 -> [0..7], MMA warp -> [8..10]
};

// Walks this CTA's work list; every role runs the same walk, so the item / tile counters agree without communication.
struct Attn9Walk {
  const Attn9Params& p;
  int w;         // current work item
  int half;      // causal: 0 = heavy query tile of the pair, 1 = light one
  int q_blk, head, batch, nt;
  int s;         // ordinal of the current (sub-)item in this CTA
  int T0;        // KV tiles processed before it
  __device__ __forceinline__ Attn9Walk(const Attn9Params& pp) : p(pp), w(blockIdx.x >> 1), half(0), q_blk(0), head(0), batch(0), nt(0), s(-1), T0(0) {}
  __device__ __forceinline__ bool next() {
    for (;;) {
      if (w >= p.total_items) return false;
      const int i = w % p.items_per_head;
      const int hb = w / p.items_per_head;
      int qt;
      bool last_half = true;
      if (p.causal) {
        qt = half == 0 ? p.num_q_blocks - 1 - i : i;
        const bool single = (p.num_q_blocks - 1 - i) == i;    // odd block count: the middle block has no partner
        last_half = (half == 1) || single;
      } else {
        qt = i;
      }
      T0 += nt;                                               // tiles of the item just finished (0 the first time)
      ++s;
      q_blk = qt;
      head = hb % p.H;
      batch = hb / p.H;
      int kv_len = p.Sk;
      if (p.causal) kv_len = min(p.Sk, qt * (2 * kA9BM) + 2 * kA9BM + (p.Sk - p.Sq));   // the later tile's range
      nt = (kv_len + kA9BN - 1) / kA9BN;
      if (last_half) { w += gridDim.x >> 1; half = 0; } else { half = 1; }
      return true;
    }
  }
};

__device__ __forceinline__ float max3n(float a, float b, float c) {
  float r;
  asm("max.f32 %0, %1, %2, %3;" : "=f"(r) : "f"(a), "f"(b), "f"(c));
  return r;
}
__device__ __forceinline__ void softmax_pair_barrier9() {     // the 256 softmax threads (warps 4..11)
  asm volatile("bar.sync 1, 256;" ::: "memory");
}

template <int D>
__global__ void __launch_bounds__(kA9Threads, 1)
attention_fwd9_kernel(const __grid_constant__ CUtensorMap tmQ, const __grid_constant__ CUtensorMap tmK,
                      const __grid_constant__ CUtensorMap tmV, const Attn9Params p) {
  using Cfg = Attn9Cfg<D>;
  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));
  uint8_t* sQ = smem;
  uint8_t* sKV = sQ + Cfg::kQBytes;
  float* mref = reinterpret_cast<float*>(sKV + Cfg::kStages * Cfg::kKVBytes);   // [tile parity][row]
  float* lsum_x = mref + 2 * kA9BM;                                             // [item parity][group][row]
  uint64_t* bars = reinterpret_cast<uint64_t*>(reinterpret_cast<uint8_t*>(mref) + Cfg::kXchgBytes);
  // L = used in the leader CTA only (the other CTA's bytes / arrivals are sent there); B = one per CTA, multicast commit
  uint64_t* q_full = bars;                              // L 1 (TMA bytes of both Q tiles)
  uint64_t* q_free = bars + 1;                          // B 1 (MMA commit after the last Q.K^T of an item)
  uint64_t* o_free = bars + 2;                          // L 1 (16 warp arrivals: O of an item is in registers)
  uint64_t* kv_full = bars + 3;                         // L kStages (TMA bytes of both halves)
  uint64_t* kv_empty = kv_full + Cfg::kStages;          // B kStages
  uint64_t* s_full = kv_empty + Cfg::kStages;           // B 2 (one MMA commit)
  uint64_t* s_free = s_full + 2;                        // L 2 (8 warp arrivals: one group of each CTA)
  uint64_t* p_full = s_free + 2;                        // L 2 (8 warp arrivals)
  uint64_t* pv_done = p_full + 2;                       // B 2 (one MMA commit)
  uint64_t* m_ready = pv_done + 2;                      // local [tile parity][quadrant]: reference max of a tile published
  uint32_t* tmem_ptr_smem = reinterpret_cast<uint32_t*>(m_ready + 8);

  const int warp = uniform_warp_idx();
  const int lane = threadIdx.x & 31;
  const int off = p.Sk - p.Sq;
  uint32_t cta_rank;
  asm volatile("mov.u32 %0, %%cluster_ctarank;" : "=r"(cta_rank));
  const bool leader = cta_rank == 0;

  if (warp == 0 && lane == 0) {
    tma_prefetch_desc(&tmQ);
    tma_prefetch_desc(&tmK);
    tma_prefetch_desc(&tmV);
  }
  if (warp == 1 && lane == 0) {
    mbar_init(q_full, 1);
    mbar_init(q_free, 1);
    mbar_init(o_free, 16);
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
    for (int i = 0; i < 8; ++i) mbar_init(&m_ready[i], 1);
    fence_mbar_init();
  }
  if (warp == 2) tmem_alloc_cg2<Cfg::kTmemCols>(tmem_ptr_smem);
  tc_fence_before();
  __syncthreads();
  // both CTAs' barriers are initialised before anything is sent across
  asm volatile("barrier.cluster.arrive.release.aligned;\n\tbarrier.cluster.wait.acquire.aligned;" ::: "memory");
  tc_fence_after();
  const uint32_t tmem_base = *tmem_ptr_smem;

  // register re-split: the control warp group needs few, each softmax thread keeps a 128-wide score row
  if (warp < 4) {
  asm volatile("setmaxnreg.dec.sync.aligned.u32 64;");
  if (warp == 0) {
    // ===================== TMA producer =====================
    if (elect_one()) {
      Attn9Walk wk(p);
      int item = 0;                              // K/V ring position, runs over all items
      while (wk.next()) {
        const int f0 = (wk.q_blk * 2 + static_cast<int>(cta_rank)) * kA9BM;
        if (wk.s > 0) mbar_wait(q_free, (wk.s - 1) & 1);     // every Q.K^T of the previous item has read sQ (both CTAs')
        if (leader) mbar_arrive_expect_tx(q_full, 2 * Cfg::kQBytes);
#pragma unroll
        for (int c = 0; c < D / 64; ++c) tma_load_4d_cg2(sQ + c * (kA9BM * 128), &tmQ, q_full, c * 64, wk.head, f0, wk.batch);
        auto load = [&](bool is_v, int j) {
          const int st = item % Cfg::kStages;
          const uint32_t ph = (item / Cfg::kStages) & 1;
          ++item;
          mbar_wait(&kv_empty[st], ph ^ 1);
          if (leader) mbar_arrive_expect_tx(&kv_full[st], 2 * Cfg::kKVBytes);
          uint8_t* dst = sKV + st * Cfg::kKVBytes;
          if (!is_v) {
            // K: this CTA's 64 keys of the tile, all of D (two 64-wide chunks, K-major)
#pragma unroll
            for (int c = 0; c < D / 64; ++c)
              tma_load_4d_cg2(dst + c * (64 * 128), &tmK, &kv_full[st], c * 64, wk.head,
                              j * kA9BN + static_cast<int>(cta_rank) * 64, wk.batch);
          } else {
            // V: all 128 keys, this CTA's 64 head-dim columns (one MN-major chunk)
            tma_load_4d_cg2(dst, &tmV, &kv_full[st], static_cast<int>(cta_rank) * 64, wk.head, j * kA9BN, wk.batch);
          }
        };
        // K one tile ahead of V: K_0, K_1, V_0, K_2, V_1, ...  (2 nt loads per item)
        load(false, 0);
        for (int j = 0; j < wk.nt; ++j) {
          if (j + 1 < wk.nt) load(false, j + 1);
          load(true, j);
        }
      }
    }
    __syncwarp();
  } else if (warp == 1 && leader) {
    // ===================== MMA issuer (leader CTA only) =====================
    constexpr uint32_t idesc_qk = make_idesc_bf16(2 * kA9BM, kA9BN, false, false);   // M = 256 over the pair
    constexpr uint32_t idesc_pv = make_idesc_bf16(2 * kA9BM, D, false, true);
    const uint64_t q_desc0 = make_smem_desc_sw128(smem_u32(sQ), 16, 1024);
    const uint64_t k_desc0 = make_smem_desc_sw128(smem_u32(sKV), 16, 1024);
    const uint64_t v_desc0 = make_smem_desc_sw128(smem_u32(sKV), kA9BN * 128, 1024);   // one 64-column chunk per CTA
    // two cursors over the same walk: Q.K^T runs ahead (into the next item too), P.V trails
    Attn9Walk qk(p), pv(p);
    bool qk_ok = qk.next();
    bool pv_ok = pv.next();
    int qk_j = 0, pv_j = 0;
    uint32_t idle = 0;
    const bool mprof = B200_ATTN_PROFILE && p.prof != nullptr && blockIdx.x == 0;
    long long m_t0 = mprof ? clock64() : 0, m_issue = 0, m_polls = 0;
    while (pv_ok) {
      bool progressed = false;
      if (mprof) ++m_polls;
#if B200_ATTN9_QK_FIRST
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
          const long long ti = mprof ? clock64() : 0;
          if (elect_one()) {
            const uint64_t kd = desc_advance(k_desc0, st * Cfg::kKVBytes);
            const uint32_t s_tmem = tmem_base + Cfg::kSCol + b * 128;
            mma_bf16_ss_cg2_c<false>(s_tmem, q_desc0, kd, idesc_qk);
#pragma unroll
            for (int kk = 1; kk < D / 16; ++kk) {
              // 64-wide chunk kk/4 of the head dim: 128 Q rows x 128 B per chunk, 64 key rows x 128 B per chunk
              const uint32_t oq = (kk >> 2) * (kA9BM * 128) + (kk & 3) * 32;
              const uint32_t ok = (kk >> 2) * (64 * 128) + (kk & 3) * 32;
              mma_bf16_ss_cg2_c<true>(s_tmem, desc_advance(q_desc0, oq), desc_advance(kd, ok), idesc_qk);
            }
            mma_commit_cg2(&s_full[b], 0x3);
            mma_commit_cg2(&kv_empty[st], 0x3);
            if (last) mma_commit_cg2(q_free, 0x3);
          }
          __syncwarp();
          if (mprof) m_issue += clock64() - ti;
          if (++qk_j == qk.nt) {
            qk_ok = qk.next();
            qk_j = 0;
          }
          progressed = true;
        }
      }
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
          const long long ti = mprof ? clock64() : 0;
          if (elect_one()) {
            const uint64_t vd = desc_advance(v_desc0, st * Cfg::kKVBytes);
            const uint32_t p_tmem = tmem_base + Cfg::kPCol + b * 64;     // 16 keys = 8 columns per MMA
            const uint32_t o_tmem = tmem_base + Cfg::kOCol;
            if (pv_j == 0) mma_bf16_ts_cg2_c<false>(o_tmem, p_tmem, vd, idesc_pv);
            else mma_bf16_ts_cg2_c<true>(o_tmem, p_tmem, vd, idesc_pv);
#pragma unroll
            for (int kk = 1; kk < kA9BN / 16; ++kk)
              mma_bf16_ts_cg2_c<true>(o_tmem, p_tmem + kk * 8, desc_advance(vd, kk * (16 * 128)), idesc_pv);
            mma_commit_cg2(&pv_done[b], 0x3);
            mma_commit_cg2(&kv_empty[st], 0x3);
          }
          __syncwarp();
          if (mprof) m_issue += clock64() - ti;
          if (++pv_j == pv.nt) {
            pv_ok = pv.next();
            pv_j = 0;
          }
          progressed = true;
        }
      }
#else
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
          const long long ti = mprof ? clock64() : 0;
          if (elect_one()) {
            const uint64_t vd = desc_advance(v_desc0, st * Cfg::kKVBytes);
            const uint32_t p_tmem = tmem_base + Cfg::kPCol + b * 64;     // 16 keys = 8 columns per MMA
            const uint32_t o_tmem = tmem_base + Cfg::kOCol;
            if (pv_j == 0) mma_bf16_ts_cg2_c<false>(o_tmem, p_tmem, vd, idesc_pv);
            else mma_bf16_ts_cg2_c<true>(o_tmem, p_tmem, vd, idesc_pv);
#pragma unroll
            for (int kk = 1; kk < kA9BN / 16; ++kk)
              mma_bf16_ts_cg2_c<true>(o_tmem, p_tmem + kk * 8, desc_advance(vd, kk * (16 * 128)), idesc_pv);
            mma_commit_cg2(&pv_done[b], 0x3);
            mma_commit_cg2(&kv_empty[st], 0x3);
          }
          __syncwarp();
          if (mprof) m_issue += clock64() - ti;
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
          const long long ti = mprof ? clock64() : 0;
          if (elect_one()) {
            const uint64_t kd = desc_advance(k_desc0, st * Cfg::kKVBytes);
            const uint32_t s_tmem = tmem_base + Cfg::kSCol + b * 128;
            mma_bf16_ss_cg2_c<false>(s_tmem, q_desc0, kd, idesc_qk);
#pragma unroll
            for (int kk = 1; kk < D / 16; ++kk) {
              // 64-wide chunk kk/4 of the head dim: 128 Q rows x 128 B per chunk, 64 key rows x 128 B per chunk
              const uint32_t oq = (kk >> 2) * (kA9BM * 128) + (kk & 3) * 32;
              const uint32_t ok = (kk >> 2) * (64 * 128) + (kk & 3) * 32;
              mma_bf16_ss_cg2_c<true>(s_tmem, desc_advance(q_desc0, oq), desc_advance(kd, ok), idesc_qk);
            }
            mma_commit_cg2(&s_full[b], 0x3);
            mma_commit_cg2(&kv_empty[st], 0x3);
            if (last) mma_commit_cg2(q_free, 0x3);
          }
          __syncwarp();
          if (mprof) m_issue += clock64() - ti;
          if (++qk_j == qk.nt) {
            qk_ok = qk.next();
            qk_j = 0;
          }
          progressed = true;
        }
      }
#endif
#if B200_WATCHDOG
      if (progressed) {
        idle = 0;
      } else if (++idle > (1u << 26)) {
        if (lane == 0)
          printf("b200: attention9 MMA scheduler stuck block=%d qk=(%d,%d) pv=(%d,%d)\n", (int)blockIdx.x, qk.s, qk_j, pv.s, pv_j);
        __trap();
      }
#endif
    }
    if (mprof && lane == 0) {
      p.prof[8] = clock64() - m_t0;
      p.prof[9] = m_issue;
      p.prof[10] = m_polls;
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
    const bool prof = B200_ATTN_PROFILE && p.prof != nullptr && blockIdx.x == 0 && warp == 4 && lane == 0;
    long long tacc[8] = {0, 0, 0, 0, 0, 0, 0, 0};
    long long tp = prof ? clock64() : 0;
#if B200_ATTN_PROFILE
#define PROF_MARK9(k) do { if (prof) { long long _t = clock64(); tacc[k] += _t - tp; tp = _t; } } while (0)
#else
#define PROF_MARK9(k) do { } while (0)
#endif
    Attn9Walk wk(p);
    while (wk.next()) {
      const int f = (wk.q_blk * 2 + static_cast<int>(cta_rank)) * kA9BM + row;
      const int nt = wk.nt;
      float m_local = -INFINITY;                 // the reference max l_part is relative to
      float l_part = 0.f;                        // row sum over this group's tiles of the item
      for (int j = (grp - wk.T0) & 1; j < nt; j += 2) {
        const int T = wk.T0 + j;                 // (T & 1) == grp
        const uint32_t use_ph = (T >> 1) & 1;
        const int t0 = j * kA9BN;
        int limit = p.Sk - t0;
        if (p.causal) limit = min(limit, f + off + 1 - t0);
        const bool need_mask = limit < kA9BN;
        mbar_wait(&s_full[b], use_ph);
        tc_fence_after();
        PROF_MARK9(0);
        uint32_t s[128];
        {
          uint32_t(&s0)[32] = *reinterpret_cast<uint32_t(*)[32]>(&s[0]);
          uint32_t(&s1)[32] = *reinterpret_cast<uint32_t(*)[32]>(&s[32]);
          uint32_t(&s2)[32] = *reinterpret_cast<uint32_t(*)[32]>(&s[64]);
          uint32_t(&s3)[32] = *reinterpret_cast<uint32_t(*)[32]>(&s[96]);
          tmem_ld_32x32(s_tmem, s0);
          tmem_ld_32x32(s_tmem + 32, s1);
#if B200_ATTN9_OVERLAP
          tmem_wait_ld();
#endif
          tmem_ld_32x32(s_tmem + 64, s2);
          tmem_ld_32x32(s_tmem + 96, s3);
#if !B200_ATTN9_OVERLAP
          tmem_wait_ld();
#endif
        }
        float mx4[4] = {-INFINITY, -INFINITY, -INFINITY, -INFINITY};
        const bool any_mask = __any_sync(0xffffffffu, need_mask);
#if B200_ATTN9_OVERLAP
        if (!any_mask) {                         // max of keys [0, 64) while keys [64, 128) are still arriving
#pragma unroll
          for (int i = 0; i < 64; i += 8) {
#pragma unroll
            for (int c = 0; c < 4; ++c)
              mx4[c] = max3n(mx4[c], __uint_as_float(s[i + 2 * c]), __uint_as_float(s[i + 2 * c + 1]));
          }
        }
        tmem_wait_ld();
#endif
        tc_fence_before();
        __syncwarp();
        if (lane == 0) mbar_arrive_remote(&s_free[b], 0);   // S[b] may be overwritten by Q.K^T of tile T+2
        PROF_MARK9(1);
        if (any_mask) {
          asm volatile("" ::: "memory");         // keep this a branch: only diagonal / tail tiles pay for masking
#pragma unroll
          for (int i = 0; i < 128; ++i)
            if (i >= limit) s[i] = 0xff800000u;  // -inf
        }
#pragma unroll
        for (int i = (B200_ATTN9_OVERLAP ? 64 : 0); i < 128; i += 8) {
#pragma unroll
          for (int c = 0; c < 4; ++c)
            mx4[c] = max3n(mx4[c], __uint_as_float(s[i + 2 * c]), __uint_as_float(s[i + 2 * c + 1]));
        }
#if B200_ATTN9_OVERLAP
        if (any_mask) {                          // the masked path did not take the early half
#pragma unroll
          for (int i = 0; i < 64; i += 8) {
#pragma unroll
            for (int c = 0; c < 4; ++c)
              mx4[c] = max3n(mx4[c], __uint_as_float(s[i + 2 * c]), __uint_as_float(s[i + 2 * c + 1]));
          }
        }
#endif
        const float m_tile = fmaxf(fmaxf(mx4[0], mx4[1]), fmaxf(mx4[2], mx4[3])) * p.scale_log2;
        // The reference max is ONE sequence over the tiles of an item: tile j starts from what tile j-1 (the other
        // group's thread of the same row) settled on.  That hand-off is the only coupling between the two groups.
        PROF_MARK9(2);
        float m_prev = -INFINITY;
        if (j > 0) {
          mbar_wait(&m_ready[((T - 1) & 1) * 4 + quad], ((T - 1) >> 1) & 1);
          m_prev = mref[((T - 1) & 1) * kA9BM + row];
        }
        PROF_MARK9(3);
        float m_used = m_prev;
        float factor = 1.0f;
        bool rescale = false;
        if (j == 0) {
          m_used = m_tile;                       // key 0 is visible to every stored row, so this is finite
        } else if (m_tile > m_prev + kA9Rescale) {
          factor = exp2_approx(m_prev - m_tile);
          m_used = m_tile;
          rescale = true;
        }
        mref[(T & 1) * kA9BM + row] = m_used;
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
        PROF_MARK9(4);
        // p = exp2(s * scale_log2 - m_used) -> bf16 pairs in place -> TMEM P[b]
        const uint64_t negm2 = pack_f32x2(-m_used, -m_used);
        uint64_t sum2[2] = {0ull, 0ull};
#if B200_ATTN9_POLY
        // 2^x = 2^n * 2^f, n = round(x) by the magic-number add, f in [-0.5, 0.5], 2^f by a degree-3 polynomial (max
        // relative error 1.0e-4, inside bf16's 2^-9), n added into the exponent bits; masked tiles keep MUFU throughout
        const bool poly_ok = !__any_sync(0xffffffffu, need_mask);
        const uint64_t magic2 = pack_f32x2(12582912.0f, 12582912.0f);          // 1.5 * 2^23
        const uint64_t nmagic2 = pack_f32x2(-12582912.0f, -12582912.0f);
        const uint64_t neg1 = pack_f32x2(-1.0f, -1.0f);
        const uint64_t c3 = pack_f32x2(0.05592203512787819f, 0.05592203512787819f);
        const uint64_t c2 = pack_f32x2(0.24264007806777954f, 0.24264007806777954f);
        const uint64_t c1 = pack_f32x2(0.6931210160255432f, 0.6931210160255432f);
        const uint64_t c0 = pack_f32x2(0.9999244809150696f, 0.9999244809150696f);
#endif
#pragma unroll
        for (int i = 0; i < 64; ++i) {
          const uint64_t x2 = fma_f32x2(pack_f32x2(__uint_as_float(s[2 * i]), __uint_as_float(s[2 * i + 1])), scale2, negm2);
          float x0, x1;
          unpack_f32x2(x2, x0, x1);
          float p0, p1;
#if B200_ATTN9_POLY
          if ((i & ((1 << B200_ATTN9_POLY) - 1)) == ((1 << B200_ATTN9_POLY) - 1) && poly_ok) {
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
            p0 = exp2_approx(x0);
            p1 = exp2_approx(x1);
          }
          sum2[i & 1] = add_f32x2(sum2[i & 1], pack_f32x2(p0, p1));
          s[i] = pack_bf16x2(p0, p1);            // keys (2i, 2i+1): low half = even key
#if B200_ATTN9_OVERLAP
          if (i == 31) tmem_st_32x32(p_tmem, *reinterpret_cast<const uint32_t(*)[32]>(&s[0]));   // keys [0, 64) go out early
#endif
        }
        {
          float a0, a1, b0, b1;
          unpack_f32x2(sum2[0], a0, a1);
          unpack_f32x2(sum2[1], b0, b1);
          l_part += (a0 + a1) + (b0 + b1);
        }
        PROF_MARK9(5);
        if (B200_ATTN_LATE_PWAIT && !any_rescale && T >= 2) {   // (the rescale path has already waited for P.V of tile T-1)
          mbar_wait(&pv_done[b], ((T - 2) >> 1) & 1);   // P[b] has been read by P.V of tile T-2
          tc_fence_after();
        }
        {
          const uint32_t(&pp0)[32] = *reinterpret_cast<const uint32_t(*)[32]>(&s[0]);
          const uint32_t(&pp1)[32] = *reinterpret_cast<const uint32_t(*)[32]>(&s[32]);
#if !B200_ATTN9_OVERLAP
          tmem_st_32x32(p_tmem, pp0);
#endif
          tmem_st_32x32(p_tmem + 32, pp1);
          tmem_wait_st();
        }
        tc_fence_before();
        __syncwarp();
        if (lane == 0) mbar_arrive_remote(&p_full[b], 0);
        PROF_MARK9(6);
        if (prof) ++tacc[7];
      }
      // ---- epilogue of the item: O / l -> bf16 -> global [batch, f, head, grp*D/2 ...]
      const long long te = prof ? clock64() : 0;
      {
        const int Tl = wk.T0 + nt - 1;           // last tile of the item
        if ((Tl & 1) != grp) mbar_wait(&m_ready[(Tl & 1) * 4 + quad], (Tl >> 1) & 1);
        const float m_final = mref[(Tl & 1) * kA9BM + row];
        if (m_final != m_local) l_part *= exp2_approx(m_local - m_final);
        float* lx = lsum_x + (wk.s & 1) * 2 * kA9BM;
        lx[grp * kA9BM + row] = l_part;
        softmax_pair_barrier9();
        const float l_sum = l_part + lx[(grp ^ 1) * kA9BM + row];
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
        if (lane == 0) mbar_arrive_remote(o_free, 0);   // O is in registers: the first P.V of the next item may overwrite it
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
      if (prof) { const long long t = clock64(); p.prof[11] += t - te; tp = t; }
    }
    if (prof) {
      for (int k = 0; k < 8; ++k) p.prof[k] = tacc[k];
    }
  }

  __syncwarp();
  tc_fence_before();
  __syncthreads();
  // no CTA of the pair may exit (or free its TMEM) while the other can still send to it or the leader's MMAs run
  asm volatile("barrier.cluster.arrive.release.aligned;\n\tbarrier.cluster.wait.acquire.aligned;" ::: "memory");
  if (warp == 2) {
    tc_fence_after();
    tmem_dealloc_cg2<Cfg::kTmemCols>(tmem_base);
  }
}

int make_bsnh_map(CUtensorMap* tm, const void* base, int B, int S, int H, int D, int64_t sb, int64_t ss, int64_t sh,
                  int rows);

template <int D>
static int launch_attention9(const void* Q, const void* K, const void* V, const Attn9Params& p, int B, int Sq, int Sk,
                             int H, const int64_t* qs, const int64_t* ks, const int64_t* vs, cudaStream_t stream) {
  using Cfg = Attn9Cfg<D>;
  static_assert(Cfg::kStages >= 3, "KV ring needs at least three stages");
  CUtensorMap tmQ, tmK, tmV;
  int rc = make_bsnh_map(&tmQ, Q, B, Sq, H, D, qs[0], qs[1], qs[2], kA9BM);
  if (rc != B200_OK) return rc;
  rc = make_bsnh_map(&tmK, K, B, Sk, H, D, ks[0], ks[1], ks[2], kA9BN / 2);   // a CTA's 64 keys of a tile
  if (rc != B200_OK) return rc;
  rc = make_bsnh_map(&tmV, V, B, Sk, H, D, vs[0], vs[1], vs[2], kA9BN);
  if (rc != B200_OK) return rc;
  auto kern = attention_fwd9_kernel<D>;
  B200_CUDA_CHECK(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, Cfg::kSmemBytes));
  int dev = 0, sms = 0;
  B200_CUDA_CHECK(cudaGetDevice(&dev));
  B200_CUDA_CHECK(cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev));
  const int pairs = p.total_items < sms / 2 ? p.total_items : sms / 2;
  cudaLaunchConfig_t cfg = {};
  cfg.gridDim = dim3(pairs * 2);
  cfg.blockDim = dim3(kA9Threads);
  cfg.dynamicSmemBytes = Cfg::kSmemBytes;
  cfg.stream = stream;
  cudaLaunchAttribute attrs[1];
  attrs[0].id = cudaLaunchAttributeClusterDimension;
  attrs[0].val.clusterDim.x = 2;
  attrs[0].val.clusterDim.y = 1;
  attrs[0].val.clusterDim.z = 1;
  cfg.attrs = attrs;
  cfg.numAttrs = 1;
  B200_CUDA_CHECK(cudaLaunchKernelEx(&cfg, kern, tmQ, tmK, tmV, p));
  launch_counter().fetch_add(1);
  return B200_OK;
}

int attention_fwd9_dispatch(const void* Q, const void* K, const void* V, void* O, int B, int Sq, int Sk, int H, int D,
                            const int64_t* strides, float scale, int causal, float* lse, cudaStream_t stream) {
  Attn9Params p = {};
  p.O = reinterpret_cast<__nv_bfloat16*>(O);
  p.Sq = Sq;
  p.Sk = Sk;
  p.o_stride_b = strides[9];
  p.o_stride_s = strides[10];
  p.o_stride_h = strides[11];
  p.scale_log2 = scale * 1.4426950408889634f;
  p.causal = causal ? 1 : 0;
  p.num_q_tiles = (Sq + kA9BM - 1) / kA9BM;
  p.H = H;
  p.num_q_blocks = (p.num_q_tiles + 1) / 2;
  p.items_per_head = p.causal ? (p.num_q_blocks + 1) / 2 : p.num_q_blocks;
  const long long total = static_cast<long long>(p.items_per_head) * H * B;
  if (total > 0x7fffffffLL) return fail(B200_ERR_INVALID_ARGUMENT, "b200_attention_fwd: too many work items");
  p.total_items = static_cast<int>(total);
  p.lse = lse;
  p.prof = profile_buffer();
  if (D != 128) return fail(B200_ERR_INVALID_ARGUMENT, "attention variant 9 (CTA pairs) needs head dim 128");
  return launch_attention9<128>(Q, K, V, p, B, Sq, Sk, H, strides, strides + 3, strides + 6, stream);
}

}  // namespace b200
