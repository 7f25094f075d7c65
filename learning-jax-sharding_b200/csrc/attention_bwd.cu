// Attention backward (dQ, dK, dV of  O = softmax(scale Q K^T [+ causal mask]) V), first correct tcgen05 path: the compute
// that jax.grad adds to the reference's train_step for the attention module (case6_attention.py:206-220; SURVEY.md section
// 8(f) row 1).  Same layout as the forward (b s n h, head dim innermost), bf16 in / out, f32 accumulation.
//
// One CTA per (batch, head, 128-key KV tile) walks the query tiles that see it:
//   S  = Q_i K_j^T            (recomputed; MMA 1, smem x smem)           P  = exp2(log2e scale S - lse2_i)     (no max pass:
//   dP = dO_i V_j^T           (MMA 2)                                    dS = scale P (dP - delta_i)            lse from the fwd)
//   dV_j += P^T dO_i          (MMA 3: A = P from smem read MN-major, i.e. transposed for free)
//   dK_j += dS^T Q_i          (MMA 4, same trick)
//   dQ_i += dS K_j            (MMA 5: per-step result, staged through the P / dS buffers and added to an f32 dQ
//                              accumulator in global memory by TMA reduce-add: per-thread vector reds cost 36 % of the time)
// with delta_i = rowsum(dO_i * O_i) from a small pre-pass.  P and dS go through shared memory as bf16, stored row-major
// [query][key] in 128-byte-swizzled 64-key chunks: that one image is K-major for dS (dQ = dS K) and MN-major for its
// transpose (dV = P^T dO, dK = dS^T Q).  Tensor memory: S | dP | dV | dK (4 x 128 columns); dQ reuses S's columns once the
// compute warps have read S.  Shared memory (D = 128): K_j, V_j, Q_i, 2 x dO_i, P, dS = 7 x 32 KB.  dO is double-buffered so
// that dP of the next step (which needs neither S's columns nor Q) runs under the dQ read-out and the Q load of that step;
// S, dP, P and dS each have their own barrier, so dV += P^T dO runs under the dS pass (DESIGN.md section 4.6 has the measured
// rates of this and of the earlier versions).
//
//   warp 0   TMA producer: K_j, V_j once; dO_i one step ahead, Q_i per step
//   warp 1   MMA issuer
//   warp 2   TMEM allocator
//   warps 4..11 two threads per query row (keys [0,64) and [64,128) of the tile; D/2 columns each of dQ, dK, dV):
//               P, dS, the dQ reduction, and at the end dK / dV -> global
#include <cuda.h>
#include <math.h>

#include "host_common.h"
#include "sm100_ptx.cuh"

namespace b200 {

#ifndef B200_ATTN_PROFILE
#define B200_ATTN_PROFILE 0   // 1: per-phase clock64 totals of block (0,0,0) into AttnBwdParams::prof (tools/gpu_check.py attn_bwd_phases)
#endif

constexpr int kBwBM = 128;      // query rows per step
constexpr int kBwBN = 128;      // keys per CTA
constexpr int kBwThreads = 384;

template <int D>
struct AttnBwdCfg {
  static constexpr int kTileBytes = 128 * D * 2;          // one of K_j, V_j, Q_i, dO_i
  static constexpr int kPBytes = kBwBM * kBwBN * 2;       // P or dS as bf16
  static constexpr int kSmemBytes = 5 * kTileBytes + 2 * kPBytes + 1024 + 256;
  static constexpr int kSCol = 0, kDPCol = 128, kDVCol = 256, kDKCol = 384;
};

struct AttnBwdParams {
  const float* lse2;      // [B, H, Sq]
  const float* delta;     // [B, H, Sq]
  float* dq_acc;          // [B, Sq, H, D] f32, zeroed by the caller of the kernel (written through tmDQ)
  __nv_bfloat16 *dK, *dV; // [B, Sk, H, D]
  int Sq, Sk, H;
  int64_t dk_stride_b, dk_stride_s, dk_stride_h, dv_stride_b, dv_stride_s, dv_stride_h;
  float scale, scale_log2;
  int causal;
  int num_q_tiles;
  long long* prof;        // bring-up aid: compute warp 4 -> [0..7], MMA warp -> [8..15]
};

__device__ __forceinline__ void bwd_group_barrier() {      // the 256 compute threads (warps 4..11)
  asm volatile("bar.sync 2, 256;" ::: "memory");
}

template <int D>
__global__ void __launch_bounds__(kBwThreads, 1)
attention_bwd_kernel(const __grid_constant__ CUtensorMap tmQ, const __grid_constant__ CUtensorMap tmK,
                     const __grid_constant__ CUtensorMap tmV, const __grid_constant__ CUtensorMap tmDO,
                     const __grid_constant__ CUtensorMap tmDQ, const AttnBwdParams p) {
  using Cfg = AttnBwdCfg<D>;
  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));
  uint8_t* sK = smem;
  uint8_t* sV = sK + Cfg::kTileBytes;
  uint8_t* sQ = sV + Cfg::kTileBytes;
  uint8_t* sDO = sQ + Cfg::kTileBytes;                    // two buffers
  uint8_t* sP = sDO + 2 * Cfg::kTileBytes;
  uint8_t* sDS = sP + Cfg::kPBytes;
  uint64_t* bars = reinterpret_cast<uint64_t*>(sDS + Cfg::kPBytes);
  uint64_t* kv_full = bars;          // TMA bytes of K_j and V_j
  uint64_t* q_full = bars + 1;       // TMA bytes of Q_i
  uint64_t* step_done = bars + 2;    // MMA commit after MMAs 3-5 of a step: sQ / sDO[.] / sP / sDS are free, dQ is in TMEM
  uint64_t* s_full = bars + 3;       // MMA commit: S is in TMEM
  uint64_t* p_full = bars + 4;       // 8 warp arrivals: P is in shared memory (S has been read)
  uint64_t* dq_free = bars + 5;      // 8 warp arrivals: dQ has been read out of S's columns
  uint64_t* do_full = bars + 6;      // [2] TMA bytes of dO_i
  uint64_t* dp_full = bars + 8;      // MMA commit: dP is in TMEM
  uint64_t* ds_full = bars + 9;      // 8 warp arrivals: dS is in shared memory (dP has been read)
  uint32_t* tmem_ptr_smem = reinterpret_cast<uint32_t*>(bars + 10);

  const int warp = uniform_warp_idx();
  const int lane = threadIdx.x & 31;
  const int j = blockIdx.x, head = blockIdx.y, batch = blockIdx.z;
  const int t0 = j * kBwBN;
  const int off = p.Sk - p.Sq;
  // first query tile that sees any key of this KV tile (causal: rows f with f + off >= t0)
  int i_start = 0;
  if (p.causal && t0 - off > 0) i_start = (t0 - off) / kBwBM;
  const int nsteps = p.num_q_tiles - i_start;

  if (warp == 0 && lane == 0) {
    tma_prefetch_desc(&tmQ);
    tma_prefetch_desc(&tmK);
    tma_prefetch_desc(&tmV);
    tma_prefetch_desc(&tmDO);
  }
  if (warp == 1 && lane == 0) {
    mbar_init(kv_full, 1);
    mbar_init(q_full, 1);
    mbar_init(step_done, 1);
    mbar_init(s_full, 1);
    mbar_init(p_full, 8);
    mbar_init(dp_full, 1);
    mbar_init(ds_full, 8);
    mbar_init(dq_free, 8);
    mbar_init(&do_full[0], 1);
    mbar_init(&do_full[1], 1);
    fence_mbar_init();
  }
  if (warp == 2) tmem_alloc<512>(tmem_ptr_smem);
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = *tmem_ptr_smem;

  if (warp == 0) {
    // ===================== TMA producer =====================
    if (elect_one()) {
      mbar_arrive_expect_tx(kv_full, 2 * Cfg::kTileBytes);
#pragma unroll
      for (int c = 0; c < D / 64; ++c) {
        tma_load_4d(sK + c * (128 * 128), &tmK, kv_full, c * 64, head, t0, batch);
        tma_load_4d(sV + c * (128 * 128), &tmV, kv_full, c * 64, head, t0, batch);
      }
      auto load_do = [&](int st) {
        uint8_t* dst = sDO + (st & 1) * Cfg::kTileBytes;
        mbar_arrive_expect_tx(&do_full[st & 1], Cfg::kTileBytes);
#pragma unroll
        for (int c = 0; c < D / 64; ++c)
          tma_load_4d(dst + c * (128 * 128), &tmDO, &do_full[st & 1], c * 64, head, (i_start + st) * kBwBM, batch);
      };
      auto load_q = [&](int st) {
        mbar_arrive_expect_tx(q_full, Cfg::kTileBytes);
#pragma unroll
        for (int c = 0; c < D / 64; ++c)
          tma_load_4d(sQ + c * (128 * 128), &tmQ, q_full, c * 64, head, (i_start + st) * kBwBM, batch);
      };
      load_do(0);
      load_q(0);
      for (int s = 0; s + 1 < nsteps; ++s) {
        load_do(s + 1);                       // its buffer was last read in step s-1, whose step_done was waited for below
        mbar_wait(step_done, s & 1);          // MMAs 3-5 of step s are done with sQ
        load_q(s + 1);
      }
    }
    __syncwarp();
  } else if (warp == 1) {
    // ===================== MMA issuer =====================
    constexpr uint32_t idesc_s = make_idesc_bf16(128, 128, false, false);     // S, dP: both operands K-major
    constexpr uint32_t idesc_t = make_idesc_bf16(128, D, true, true);         // dV, dK: A = (P | dS)^T, B = dO | Q
    constexpr uint32_t idesc_q = make_idesc_bf16(128, D, false, true);        // dQ: A = dS K-major, B = K MN-major
    const uint64_t q_k = make_smem_desc_sw128(smem_u32(sQ), 16, 1024);        // K-major views ([row][d])
    const uint64_t do_k0 = make_smem_desc_sw128(smem_u32(sDO), 16, 1024);
    const uint64_t k_k = make_smem_desc_sw128(smem_u32(sK), 16, 1024);
    const uint64_t v_k = make_smem_desc_sw128(smem_u32(sV), 16, 1024);
    const uint64_t q_mn = make_smem_desc_sw128(smem_u32(sQ), 128 * 128, 1024);   // MN-major views ([k = row][n = d])
    const uint64_t do_mn0 = make_smem_desc_sw128(smem_u32(sDO), 128 * 128, 1024);
    const uint64_t k_mn = make_smem_desc_sw128(smem_u32(sK), 128 * 128, 1024);
    const uint64_t p_mn = make_smem_desc_sw128(smem_u32(sP), 128 * 128, 1024);   // P^T: [k = query][m = key]
    const uint64_t ds_mn = make_smem_desc_sw128(smem_u32(sDS), 128 * 128, 1024);
    const uint64_t ds_k = make_smem_desc_sw128(smem_u32(sDS), 16, 1024);         // dS: [m = query][k = key]
    const uint32_t t_s = tmem_base + Cfg::kSCol, t_dp = tmem_base + Cfg::kDPCol;
    const uint32_t t_dv = tmem_base + Cfg::kDVCol, t_dk = tmem_base + Cfg::kDKCol;
    mbar_wait(kv_full, 0);
    // The tensor pipe never waits for a whole softmax pass: per step S -> [P pass] -> dV while [dS pass] -> dK, dQ -> dP of
    // the NEXT step (it needs neither Q nor S's columns), so that dP is ready when the dS pass of that step wants it.
    auto issue_dp = [&](int st) {
      const uint64_t do_k = desc_advance(do_k0, (st & 1) * Cfg::kTileBytes);
      mbar_wait(&do_full[st & 1], (st >> 1) & 1);
      tc_fence_after();
      if (elect_one()) {
#pragma unroll
        for (int kk = 0; kk < D / 16; ++kk) {           // contraction over the head dim, 64-wide chunks 16 KB apart
          const uint32_t o = (kk >> 2) * (128 * 128) + (kk & 3) * 32;
          mma_bf16_ss(t_dp, desc_advance(do_k, o), desc_advance(v_k, o), idesc_s, kk > 0);
        }
        mma_commit(dp_full);
      }
      __syncwarp();
    };
    const bool mprof = B200_ATTN_PROFILE && p.prof != nullptr && blockIdx.x == 0 && blockIdx.y == 0 && blockIdx.z == 0 && lane == 0;
    long long macc[8] = {0, 0, 0, 0, 0, 0, 0, 0};
    long long mtp = mprof ? clock64() : 0;
#if B200_ATTN_PROFILE
#define MMARK(k) do { if (mprof) { long long _t = clock64(); macc[k] += _t - mtp; mtp = _t; } } while (0)
#else
#define MMARK(k) do { } while (0)
#endif
    issue_dp(0);
    MMARK(6);
    for (int s = 0; s < nsteps; ++s) {
      const uint64_t do_mn = desc_advance(do_mn0, (s & 1) * Cfg::kTileBytes);
      mbar_wait(q_full, s & 1);
      MMARK(0);
      if (s > 0) mbar_wait(dq_free, (s - 1) & 1);       // S's columns hold dQ of the previous step until it has been read out
      tc_fence_after();
      MMARK(1);
      if (elect_one()) {
#pragma unroll
        for (int kk = 0; kk < D / 16; ++kk) {
          const uint32_t o = (kk >> 2) * (128 * 128) + (kk & 3) * 32;
          mma_bf16_ss(t_s, desc_advance(q_k, o), desc_advance(k_k, o), idesc_s, kk > 0);
        }
        mma_commit(s_full);
      }
      __syncwarp();
      MMARK(2);
      mbar_wait(p_full, s & 1);
      tc_fence_after();
      MMARK(3);
      if (elect_one()) {
#pragma unroll
        for (int kk = 0; kk < kBwBM / 16; ++kk) {       // contraction over the 128 query rows: 16 rows = 2 KB per step
          const uint32_t o = kk * (16 * 128);
          mma_bf16_ss(t_dv, desc_advance(p_mn, o), desc_advance(do_mn, o), idesc_t, (s > 0 || kk > 0) ? 1u : 0u);
        }
      }
      __syncwarp();
      MMARK(2);
      mbar_wait(ds_full, s & 1);
      tc_fence_after();
      MMARK(4);
      if (elect_one()) {
#pragma unroll
        for (int kk = 0; kk < kBwBM / 16; ++kk) {
          const uint32_t o = kk * (16 * 128);
          mma_bf16_ss(t_dk, desc_advance(ds_mn, o), desc_advance(q_mn, o), idesc_t, (s > 0 || kk > 0) ? 1u : 0u);
        }
#pragma unroll
        for (int kk = 0; kk < kBwBN / 16; ++kk) {       // contraction over the 128 keys
          const uint32_t oa = (kk >> 2) * (128 * 128) + (kk & 3) * 32;
          mma_bf16_ss(t_s, desc_advance(ds_k, oa), desc_advance(k_mn, kk * (16 * 128)), idesc_q, kk > 0);
        }
        mma_commit(step_done);
      }
      __syncwarp();
      MMARK(2);
      if (s + 1 < nsteps) issue_dp(s + 1);
      MMARK(5);
    }
    if (mprof) {
      for (int k = 0; k < 7; ++k) p.prof[8 + k] = macc[k];
      p.prof[15] = nsteps;
    }
  } else if (warp >= 4) {
    // ===================== P, dS, dQ reduction, dK / dV epilogue =====================
    const int quad = warp & 3;
    const int grp = (warp - 4) >> 2;                   // keys [64 grp, 64 grp + 64) of the tile; columns [grp D/2, +D/2) of dQ, dK, dV
    const int r = quad * 32 + lane;                    // query row of the step's tile (and key row in the epilogue)
    const uint32_t lane_base = tmem_base + (static_cast<uint32_t>(quad * 32) << 16);
    const int64_t bh = static_cast<int64_t>(batch) * p.H + head;
    const bool cprof = B200_ATTN_PROFILE && p.prof != nullptr && blockIdx.x == 0 && blockIdx.y == 0 && blockIdx.z == 0 && warp == 4 && lane == 0;
    long long cacc[8] = {0, 0, 0, 0, 0, 0, 0, 0};
    long long ctp = cprof ? clock64() : 0;
#if B200_ATTN_PROFILE
#define CMARK(k) do { if (cprof) { long long _t = clock64(); cacc[k] += _t - ctp; ctp = _t; } } while (0)
#else
#define CMARK(k) do { } while (0)
#endif
    for (int s = 0; s < nsteps; ++s) {
      const int f = (i_start + s) * kBwBM + r;
      const bool row_valid = f < p.Sq;
      const float lse = row_valid ? p.lse2[bh * p.Sq + f] : 0.f;
      const float dl = row_valid ? p.delta[bh * p.Sq + f] : 0.f;
      int limit = p.Sk - t0;                           // keys [0, limit) of this tile are visible to this row
      if (p.causal) limit = min(limit, f + off + 1 - t0);
      if (!row_valid) limit = 0;
      const bool all_visible = __all_sync(0xffffffffu, limit >= kBwBN);   // warp-uniform: interior tiles skip the per-key select
      const float neg_lse = -lse, dls = dl * p.scale;
      mbar_wait(s_full, s & 1);
      tc_fence_after();
      CMARK(0);
      if (s > 0) {
        // sP / sDS doubled as the dQ staging tile of the previous step, handed to TMA as D/32 bulk groups in column order:
        // sP is free once all but the last D/64 groups have been read (the reduce runs at ~34 B/clk per SM, so waiting for
        // the whole tile here cost ~1900 cycles per step); sDS is waited for before the dS pass
        if (warp == 4 && lane == 0) tma_store_wait_read<D / 64>();
        bwd_group_barrier();
      }
      CMARK(1);
      // ---- P pass (this thread: its row, 64 keys), kept in registers for the dS pass
      float pf[64];
#pragma unroll
      for (int cc = 0; cc < 2; ++cc) {
        const int c = 2 * grp + cc;
        uint32_t sv[32];
        tmem_ld_32x32(lane_base + Cfg::kSCol + c * 32, sv);
        tmem_wait_ld();
        uint32_t pp[16];
#pragma unroll
        for (int k = 0; k < 32; k += 2) {
#pragma unroll
          for (int e = 0; e < 2; ++e) {
            const int key = c * 32 + k + e;
            const float pe = exp2_approx(fmaf(__uint_as_float(sv[k + e]), p.scale_log2, neg_lse));
            pf[cc * 32 + k + e] = (all_visible || key < limit) ? pe : 0.f;
          }
          pp[k >> 1] = pack_bf16x2(pf[cc * 32 + k], pf[cc * 32 + k + 1]);
        }
        // row r, keys [32 c, 32 c + 32): four 16-byte units of the 128-byte-swizzled 64-key chunk
        const uint32_t row_off = (c >> 1) * (kBwBM * 128) + r * 128;
#pragma unroll
        for (int u = 0; u < 4; ++u) {
          const uint32_t unit = ((c & 1) * 4 + u) ^ (r & 7);
          *reinterpret_cast<uint4*>(sP + row_off + (unit << 4)) = make_uint4(pp[4 * u], pp[4 * u + 1], pp[4 * u + 2], pp[4 * u + 3]);
        }
      }
      fence_async_smem();                              // generic-proxy stores -> visible to the MMAs' async-proxy reads
      tc_fence_before();
      __syncwarp();
      if (lane == 0) mbar_arrive(p_full);              // dV += P^T dO may start
      CMARK(2);
      // ---- dS pass
      mbar_wait(dp_full, s & 1);
      tc_fence_after();
      if (s > 0) {
        if (warp == 4 && lane == 0) tma_store_wait_read<0>();
        bwd_group_barrier();
      }
      CMARK(3);
#pragma unroll
      for (int cc = 0; cc < 2; ++cc) {
        const int c = 2 * grp + cc;
        uint32_t dv[32];
        tmem_ld_32x32(lane_base + Cfg::kDPCol + c * 32, dv);
        tmem_wait_ld();
        uint32_t dd[16];
#pragma unroll
        for (int k = 0; k < 32; k += 2) {
          const float d0 = pf[cc * 32 + k] * fmaf(__uint_as_float(dv[k]), p.scale, -dls);       // p * scale * (dP - delta)
          const float d1 = pf[cc * 32 + k + 1] * fmaf(__uint_as_float(dv[k + 1]), p.scale, -dls);
          dd[k >> 1] = pack_bf16x2(d0, d1);
        }
        const uint32_t row_off = (c >> 1) * (kBwBM * 128) + r * 128;
#pragma unroll
        for (int u = 0; u < 4; ++u) {
          const uint32_t unit = ((c & 1) * 4 + u) ^ (r & 7);
          *reinterpret_cast<uint4*>(sDS + row_off + (unit << 4)) = make_uint4(dd[4 * u], dd[4 * u + 1], dd[4 * u + 2], dd[4 * u + 3]);
        }
      }
      fence_async_smem();
      tc_fence_before();
      __syncwarp();
      if (lane == 0) mbar_arrive(ds_full);             // dK += dS^T Q and dQ = dS K may start
      CMARK(4);
      // ---- dQ_i of this step: TMEM (S's columns) -> f32 accumulator in global memory
      mbar_wait(step_done, s & 1);
      tc_fence_after();
      CMARK(5);
      // this thread's D/2 columns of its row -> the staging tile (f32, [32-column chunk][row][128 B], 16-byte units XOR-swizzled
      // like the tensor map), then ONE thread hands the tile to TMA: dq_acc[b, f0.., head, :] += tile, rows past Sq clipped
#pragma unroll 1
      for (int c = grp * (D / 64); c < (grp + 1) * (D / 64); ++c) {
        uint32_t q[32];
        tmem_ld_32x32(lane_base + Cfg::kSCol + c * 32, q);
        tmem_wait_ld();
        uint8_t* dst = sP + c * (kBwBM * 128) + r * 128;
#pragma unroll
        for (int u = 0; u < 8; ++u)
          *reinterpret_cast<uint4*>(dst + ((u ^ (r & 7)) << 4)) = make_uint4(q[4 * u], q[4 * u + 1], q[4 * u + 2], q[4 * u + 3]);
      }
      // (arriving on dq_free before these stores -- all columns loaded first -- was measured and is slower: 730 vs 774 TFLOP/s)
      tc_fence_before();
      __syncwarp();
      if (lane == 0) mbar_arrive(dq_free);             // S's columns may take the next step's scores
      fence_async_smem();
      bwd_group_barrier();
      if (warp == 4 && lane == 0) {
#pragma unroll
        for (int c = 0; c < D / 32; ++c) {
          tma_reduce_add_4d(&tmDQ, sP + c * (kBwBM * 128), c * 32, head, (i_start + s) * kBwBM, batch);
          tma_store_commit();                          // one bulk group per 32-column chunk, in buffer order
        }
      }
      CMARK(6);
    }
    if (cprof) {
      for (int k = 0; k < 7; ++k) p.prof[k] = cacc[k];
    }
    if (warp == 4 && lane == 0) tma_store_wait_all<0>();  // the last tile has left shared memory (and landed) before the CTA exits
    // ---- dK_j, dV_j: complete after the last step's MMAs (step_done of the last step has been waited for above)
    const int t = t0 + r;
    if (nsteps > 0) {
      __nv_bfloat16* dk_row = p.dK + batch * p.dk_stride_b + static_cast<int64_t>(t) * p.dk_stride_s + head * p.dk_stride_h;
      __nv_bfloat16* dv_row = p.dV + batch * p.dv_stride_b + static_cast<int64_t>(t) * p.dv_stride_s + head * p.dv_stride_h;
#pragma unroll 1
      for (int c = grp * (D / 64); c < (grp + 1) * (D / 64); ++c) {
        uint32_t a[32], b[32];
        tmem_ld_32x32(lane_base + Cfg::kDKCol + c * 32, a);
        tmem_ld_32x32(lane_base + Cfg::kDVCol + c * 32, b);
        tmem_wait_ld();
        if (t < p.Sk) {
#pragma unroll
          for (int k = 0; k < 32; k += 8) {
            uint4 o;
            o.x = pack_bf16x2(__uint_as_float(a[k]), __uint_as_float(a[k + 1]));
            o.y = pack_bf16x2(__uint_as_float(a[k + 2]), __uint_as_float(a[k + 3]));
            o.z = pack_bf16x2(__uint_as_float(a[k + 4]), __uint_as_float(a[k + 5]));
            o.w = pack_bf16x2(__uint_as_float(a[k + 6]), __uint_as_float(a[k + 7]));
            *reinterpret_cast<uint4*>(dk_row + c * 32 + k) = o;
            o.x = pack_bf16x2(__uint_as_float(b[k]), __uint_as_float(b[k + 1]));
            o.y = pack_bf16x2(__uint_as_float(b[k + 2]), __uint_as_float(b[k + 3]));
            o.z = pack_bf16x2(__uint_as_float(b[k + 4]), __uint_as_float(b[k + 5]));
            o.w = pack_bf16x2(__uint_as_float(b[k + 6]), __uint_as_float(b[k + 7]));
            *reinterpret_cast<uint4*>(dv_row + c * 32 + k) = o;
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
    tmem_dealloc<512>(tmem_base);
  }
}

// delta[b, h, s] = sum_d dO[b, s, h, d] * O[b, s, h, d]: one warp per (b, s, h) row, 16-byte loads
template <int D>
__global__ void __launch_bounds__(256) attention_delta_kernel(const __nv_bfloat16* __restrict__ O, const __nv_bfloat16* __restrict__ dO,
                                                             float* __restrict__ delta, int B, int S, int H, int64_t o_sb,
                                                             int64_t o_ss, int64_t o_sh, int64_t d_sb, int64_t d_ss,
                                                             int64_t d_sh) {
  const int lane = threadIdx.x & 31;
  const int64_t rows = static_cast<int64_t>(B) * S * H;
  const int64_t wstride = static_cast<int64_t>(gridDim.x) * (blockDim.x >> 5);
  for (int64_t row = static_cast<int64_t>(blockIdx.x) * (blockDim.x >> 5) + (threadIdx.x >> 5); row < rows; row += wstride) {
    const int h = static_cast<int>(row % H);
    const int s = static_cast<int>((row / H) % S);
    const int b = static_cast<int>(row / (static_cast<int64_t>(H) * S));
    const __nv_bfloat16* o = O + b * o_sb + s * o_ss + h * o_sh;
    const __nv_bfloat16* g = dO + b * d_sb + s * d_ss + h * d_sh;
    float acc = 0.f;
    if (lane * 8 < D) {
      const uint4 ov = *reinterpret_cast<const uint4*>(o + lane * 8);
      const uint4 gv = *reinterpret_cast<const uint4*>(g + lane * 8);
      const __nv_bfloat162* oh = reinterpret_cast<const __nv_bfloat162*>(&ov);
      const __nv_bfloat162* gh = reinterpret_cast<const __nv_bfloat162*>(&gv);
#pragma unroll
      for (int e = 0; e < 4; ++e) {
        const float2 a = __bfloat1622float2(oh[e]), c = __bfloat1622float2(gh[e]);
        acc += a.x * c.x + a.y * c.y;
      }
    }
#pragma unroll
    for (int m = 16; m > 0; m >>= 1) acc += __shfl_xor_sync(0xffffffffu, acc, m);
    if (lane == 0) delta[(static_cast<int64_t>(b) * H + h) * S + s] = acc;
  }
}

int make_bsnh_map(CUtensorMap* tm, const void* base, int B, int S, int H, int D, int64_t sb, int64_t ss, int64_t sh,
                  int rows);

template <int D>
static int launch_attention_bwd(const void* Q, const void* K, const void* V, const void* O, const void* dO, AttnBwdParams p,
                                float* delta, int B, int Sq, int Sk, int H, const int64_t* st, cudaStream_t stream) {
  using Cfg = AttnBwdCfg<D>;
  static_assert(Cfg::kSmemBytes <= 227 * 1024, "attention backward tiles do not fit shared memory");
  CUtensorMap tmQ, tmK, tmV, tmDO;
  int rc = make_bsnh_map(&tmQ, Q, B, Sq, H, D, st[0], st[1], st[2], 128);
  if (rc == B200_OK) rc = make_bsnh_map(&tmK, K, B, Sk, H, D, st[3], st[4], st[5], 128);
  if (rc == B200_OK) rc = make_bsnh_map(&tmV, V, B, Sk, H, D, st[6], st[7], st[8], 128);
  if (rc == B200_OK) rc = make_bsnh_map(&tmDO, dO, B, Sq, H, D, st[12], st[13], st[14], 128);
  if (rc != B200_OK) return rc;
  CUtensorMap tmDQ;
  {
    uint64_t dims[4] = {static_cast<uint64_t>(D), static_cast<uint64_t>(H), static_cast<uint64_t>(Sq), static_cast<uint64_t>(B)};
    uint64_t strides[4] = {0, static_cast<uint64_t>(D) * 4, static_cast<uint64_t>(H) * D * 4,
                           static_cast<uint64_t>(Sq) * H * D * 4};
    uint32_t box[4] = {32, 1, 128, 1};
    rc = make_tensor_map(&tmDQ, p.dq_acc, 4, 4, dims, strides, box, 3);
    if (rc != B200_OK) return rc;
  }
  const int64_t rows = static_cast<int64_t>(B) * Sq * H;
  const int64_t blocks = (rows + 7) / 8;
  attention_delta_kernel<D><<<static_cast<int>(blocks < 148 * 8 ? blocks : 148 * 8), 256, 0, stream>>>(
      reinterpret_cast<const __nv_bfloat16*>(O), reinterpret_cast<const __nv_bfloat16*>(dO), delta, B, Sq, H, st[9], st[10],
      st[11], st[12], st[13], st[14]);
  B200_CUDA_CHECK(cudaGetLastError());
  launch_counter().fetch_add(1);
  p.delta = delta;
  auto kern = attention_bwd_kernel<D>;
  B200_CUDA_CHECK(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, Cfg::kSmemBytes));
  dim3 grid((Sk + kBwBN - 1) / kBwBN, H, B);
  kern<<<grid, kBwThreads, Cfg::kSmemBytes, stream>>>(tmQ, tmK, tmV, tmDO, tmDQ, p);
  B200_CUDA_CHECK(cudaGetLastError());
  launch_counter().fetch_add(1);
  return B200_OK;
}

}  // namespace b200

using namespace b200;

extern "C" int b200_attention_bwd(const b200_attention_bwd_args* a, void* stream) {
  B200_REQUIRE(a, "b200_attention_bwd: null args");
  B200_REQUIRE(a->Q && a->K && a->V && a->O && a->dO && a->lse2 && a->dQ && a->dK && a->dV && a->workspace,
               "b200_attention_bwd: null pointer");
  const int B = a->B, Sq = a->Sq, Sk = a->Sk, H = a->H, D = a->D;
  B200_REQUIRE(B > 0 && Sq > 0 && Sk > 0 && H > 0, "b200_attention_bwd: B, Sq, Sk, H must be positive");
  B200_REQUIRE(D == 64 || D == 128, "b200_attention_bwd: head dim must be 64 or 128 (got %d)", D);
  B200_REQUIRE(B <= 65535 && H <= 65535, "b200_attention_bwd: B and H must be <= 65535");
  B200_REQUIRE(!a->causal || Sk >= Sq, "b200_attention_bwd: causal attention needs Sk >= Sq");
  for (int i = 0; i < 24; ++i)
    B200_REQUIRE(a->strides[i] > 0 && a->strides[i] % 8 == 0, "b200_attention_bwd: strides must be positive multiples of 8");
  // dQ is produced by a flat f32 -> bf16 pass over the accumulator: it has to be the contiguous [B, Sq, H, D] tensor
  B200_REQUIRE(a->strides[17] == D && a->strides[16] == static_cast<int64_t>(H) * D &&
                   a->strides[15] == static_cast<int64_t>(Sq) * H * D,
               "b200_attention_bwd: dQ must be contiguous [B, Sq, H, D]");
  const uintptr_t all = reinterpret_cast<uintptr_t>(a->Q) | reinterpret_cast<uintptr_t>(a->K) | reinterpret_cast<uintptr_t>(a->V) |
                        reinterpret_cast<uintptr_t>(a->O) | reinterpret_cast<uintptr_t>(a->dO) | reinterpret_cast<uintptr_t>(a->dQ) |
                        reinterpret_cast<uintptr_t>(a->dK) | reinterpret_cast<uintptr_t>(a->dV) |
                        reinterpret_cast<uintptr_t>(a->workspace);
  B200_REQUIRE(all % 16 == 0, "b200_attention_bwd: operands must be 16-byte aligned");
  cudaStream_t st = reinterpret_cast<cudaStream_t>(stream);
  const int64_t nq = static_cast<int64_t>(B) * Sq * H * D;
  float* dq_acc = a->workspace;                       // [B, Sq, H, D] f32
  float* delta = a->workspace + nq;                   // [B, H, Sq] f32
  B200_CUDA_CHECK(cudaMemsetAsync(dq_acc, 0, static_cast<size_t>(nq) * sizeof(float), st));
  AttnBwdParams p = {};
  p.lse2 = a->lse2;
  p.dq_acc = dq_acc;
  p.dK = reinterpret_cast<__nv_bfloat16*>(a->dK);
  p.dV = reinterpret_cast<__nv_bfloat16*>(a->dV);
  p.Sq = Sq;
  p.Sk = Sk;
  p.H = H;
  p.dk_stride_b = a->strides[18];
  p.dk_stride_s = a->strides[19];
  p.dk_stride_h = a->strides[20];
  p.dv_stride_b = a->strides[21];
  p.dv_stride_s = a->strides[22];
  p.dv_stride_h = a->strides[23];
  p.scale = a->scale;
  p.scale_log2 = a->scale * 1.4426950408889634f;
  p.causal = a->causal ? 1 : 0;
  p.num_q_tiles = (Sq + kBwBM - 1) / kBwBM;
  p.prof = profile_buffer();
  int rc = D == 64 ? launch_attention_bwd<64>(a->Q, a->K, a->V, a->O, a->dO, p, delta, B, Sq, Sk, H, a->strides, st)
                   : launch_attention_bwd<128>(a->Q, a->K, a->V, a->O, a->dO, p, delta, B, Sq, Sk, H, a->strides, st);
  if (rc != B200_OK) return rc;
  return b200_cast_f32_to_bf16(dq_acc, a->dQ, nq, stream);
}
