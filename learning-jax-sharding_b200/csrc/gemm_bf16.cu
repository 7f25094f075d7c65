// bf16 GEMM with fused epilogue for sm_100a: C = epi(A . B + bias).
//
// Persistent, warp-specialised kernel; one CTA (or one CTA pair, cta_group::2) per SM:
//   warp 0      TMA producer   global -> 128B-swizzled smem ring (mbarrier full/empty)
//   warp 1      MMA issuer     one thread issues tcgen05.mma, accumulators live in TMEM
//   warp 2      TMEM allocator (512 columns = two accumulator stages of up to 256 columns)
//   warps 4..7  epilogue       tcgen05.ld TMEM -> registers -> bias / activation -> global
// The two accumulator stages let the epilogue of tile i overlap the main loop of tile i+1.
//
// This is the per-device compute behind flax nn.Dense(dtype=bf16) (reference
// case6_attention.py:61-90,96-99,139; case5_attention_dense.py:59-67) and both halves
// of the FFN (case6_attention.py:36-40).  B is taken in the layout the reference keeps
// Dense kernels in, (in_features, out_features) row-major = "[K,N]", which tcgen05 reads
// as an MN-major operand; the [N,K] layout (K-major) is also supported.
#include <cuda.h>

#include "host_common.h"
#include "sm100_ptx.cuh"

#ifndef B200_PROGRESS_FENCE
#define B200_PROGRESS_FENCE 1   // fence between a warp's tile stores and its progress count: 1 = fence.acq_rel.sys, 2 = fence.sc.sys
#endif                          // (0 = none: timing experiments only, the flags are then not ordered after the data)

namespace b200 {

constexpr int kBM = 128;        // rows of C per CTA (TMEM lanes)
constexpr int kBK = 64;         // K elements per pipeline stage = one 128-byte swizzle row
constexpr int kUmmaK = 16;      // K per tcgen05.mma for 16-bit inputs
constexpr int kGemmThreads = 256;

template <int BN, int G>
struct GemmCfg {
  static constexpr int kBNcta = BN / G;                         // B rows/cols loaded by one CTA
  static constexpr int kABytes = kBM * kBK * 2;                 // 16 KB
  static constexpr int kBBytes = kBNcta * kBK * 2;
  static constexpr int kStageBytes = kABytes + kBBytes;
  static constexpr int kStoreStageBytes = 4 * 32 * 128;         // per epilogue warp: 32 rows x 128 B
  static constexpr int kSmemBudget = 227 * 1024 - 1024 /*align slack*/ - 256 /*barriers*/ - kStoreStageBytes;
  static constexpr int kStages = kSmemBudget / kStageBytes > 8 ? 8 : kSmemBudget / kStageBytes;
  static constexpr int kAccStages = 2;
  static constexpr int kTmemCols = 512;
  static constexpr int kSmemBytes = kStages * kStageBytes + kStoreStageBytes + 1024 /*align slack*/ + 256 /*barriers*/;
};

struct GemmParams {
  void* C;
  const __nv_bfloat16* bias;
  int M, N, K;
  int64_t ldc;
  int epilogue;
  int out_f32;
  int tiles_m, tiles_n;
  // optional row-blocked view of A: logical row m lives at physical row
  // (m / a_rows_per_block) * a_block_stride + m % a_rows_per_block   (0 = plain row-major)
  int a_rows_per_block, a_block_stride;
  // two-level form of that view (a_blocks_inner > 0): block q of a_rows_per_block rows starts at physical row
  // (q % a_blocks_inner) * a_block_stride + (q / a_blocks_inner) * a_outer_stride -- "sequence block j of every batch
  // item, owner by owner": the row order in which a [batch, seq]-shaped activation is reduce-scattered along seq
  int a_blocks_inner, a_outer_stride;
  int group_m;   // tile rasterisation: group_m row-blocks sweep all column-blocks together (L2 reuse)
  // optional row-block scatter of the OUTPUT (fused reduce-scatter push): logical row m belongs to row block
  // (m / m_rows_per_block + m_block_rot) % m_num_blocks; A is read from that physical block and C is written
  // to c_block_ptrs[block] (a peer GPU's receive slot, or local memory) at row m % m_rows_per_block.
  void* const* c_block_ptrs;
  int m_rows_per_block, m_block_rot, m_num_blocks;
  // 1: consecutive tiles cycle through the owner blocks (tile i -> block (i % blocks + rot) % blocks, position i / blocks
  // inside it) instead of finishing one block after the other: every wave then pushes to ALL peers at once and the NVLink
  // traffic of a scatter to peer memory is spread over the whole kernel (measured: 677 GB/s per direction is what peer
  // stores reach here whatever their shape, tools/diag/nvlink_store.cu; block after block, the 7/8 of the tiles that
  // are remote need more than that while they run and the link idles during the own block)
  int m_interleave;
  // Progress signals of a row-block scatter (all null / 0 = off).  Tiles are produced raster group by raster group
  // (group_m row-blocks x all column-blocks); once every tile of group g has been stored -- and made visible system-wide --
  // the kernel writes `done_epoch` into flag [done_sender * done_groups + g] of EVERY owner's flag array, so that an owner
  // can start reducing the rows of group g while this kernel is still producing the later groups.
  unsigned int* done_counters;          // local, [done_groups], monotonic over launches (never reset)
  unsigned int* const* done_flag_ptrs;  // device array [m_num_blocks] of the owners' flag arrays (peer memory)
  int done_sender, done_groups;
  unsigned int done_epoch;
  int num_stages;   // smem ring depth actually used (<= Cfg::kStages); tuning experiments only
  int a_mn_major;   // 1: A is stored [K, M] row-major (the transposed operand of a weight-gradient GEMM dW = X^T . dY):
                    // the A tile is loaded as two 64-row-wide MN-major chunks and the MMA reads it with the MN-major flag
  int k_serpentine; // 1: every other tile of a CTA walks K downwards, so a wave starts on the K range the previous wave
                    // has just left in L2 (the panels of one wave exceed L2 when K is large)
};

__device__ __forceinline__ float gelu_tanh_f(float x) {
  // 0.5 x (1 + tanh( sqrt(2/pi) (x + 0.044715 x^3) ))  -- jax.nn.gelu(approximate=True)
  const float k0 = 0.7978845608028654f, k1 = 0.044715f;
  float u = k0 * (x + k1 * x * x * x);
  return 0.5f * x * (1.0f + tanh_approx(u));
}

__device__ __forceinline__ float apply_epilogue(float acc, float bias, int epi, bool out_f32) {
  if (out_f32 && epi == B200_EPI_NONE) return acc;   // raw f32 partial sums (row-parallel reductions)
  float v = bf16_round(acc);                          // Dense: result is cast to bf16 first ...
  if (epi == B200_EPI_BIAS || epi == B200_EPI_BIAS_RELU || epi == B200_EPI_BIAS_GELU_TANH)
    v = bf16_round(v + bias);                         // ... then the bf16 bias is added in bf16
  if (epi == B200_EPI_BIAS_RELU || epi == B200_EPI_RELU) v = fmaxf(v, 0.0f);
  if (epi == B200_EPI_BIAS_GELU_TANH || epi == B200_EPI_GELU_TANH) v = gelu_tanh_f(v);
  return v;
}

// first logical row of the 128-row tile (m_blk, cta_rank) after the row-block scatter's permutation
template <int G>
__device__ __forceinline__ int scatter_row0(const GemmParams& p, int m_blk, uint32_t cta_rank) {
  int m0 = (m_blk * G + cta_rank) * kBM;
  if (p.m_rows_per_block > 0) {
    if (p.m_interleave) {
      const int blk = (m_blk % p.m_num_blocks + p.m_block_rot) % p.m_num_blocks;
      m0 = blk * p.m_rows_per_block + (m_blk / p.m_num_blocks) * (kBM * G) + cta_rank * kBM;
    } else {
      m0 = ((m0 / p.m_rows_per_block + p.m_block_rot) % p.m_num_blocks) * p.m_rows_per_block + m0 % p.m_rows_per_block;
    }
  }
  return m0;
}

__device__ __forceinline__ void tile_coords(int tile, int tiles_m, int tiles_n, int group_m, int& m_blk,
                                            int& n_blk) {
  const int per_group = group_m * tiles_n;
  const int group = tile / per_group;
  const int first_m = group * group_m;
  const int gsize = min(tiles_m - first_m, group_m);
  const int in_group = tile - group * per_group;
  m_blk = first_m + in_group % gsize;
  n_blk = in_group / gsize;
}

template <int BN, int G, bool kBKmajor>
__global__ void __launch_bounds__(kGemmThreads, 1)
gemm_bf16_kernel(const __grid_constant__ CUtensorMap tmA, const __grid_constant__ CUtensorMap tmB,
                 const GemmParams p) {
  using Cfg = GemmCfg<BN, G>;
  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));
  uint8_t* store_stage = smem + Cfg::kStages * Cfg::kStageBytes;
  uint64_t* bars = reinterpret_cast<uint64_t*>(store_stage + Cfg::kStoreStageBytes);
  uint64_t* full_bar = bars;
  uint64_t* empty_bar = bars + Cfg::kStages;
  uint64_t* tmem_full_bar = bars + 2 * Cfg::kStages;
  uint64_t* tmem_empty_bar = tmem_full_bar + Cfg::kAccStages;
  uint64_t* prog_full = tmem_empty_bar + Cfg::kAccStages;     // progress signals: the 4 epilogue warps -> warp 3
  uint64_t* prog_empty = prog_full + 1;                       //                   warp 3 -> the epilogue warps
  uint32_t* tmem_ptr_smem = reinterpret_cast<uint32_t*>(prog_empty + 1);

  const int warp = uniform_warp_idx();
  const int lane = threadIdx.x & 31;
  uint32_t cta_rank = 0;
  if (G == 2) asm volatile("mov.u32 %0, %%cluster_ctarank;" : "=r"(cta_rank));
  const bool leader = (cta_rank == 0);

  const int num_tiles = p.tiles_m * p.tiles_n;
  const int num_workers = gridDim.x / G;       // CTAs (G=1) or CTA pairs (G=2)
  const int worker = blockIdx.x / G;
  const int num_kb = (p.K + kBK - 1) / kBK;

  if (warp == 0 && lane == 0) {
    tma_prefetch_desc(&tmA);
    tma_prefetch_desc(&tmB);
  }
  if (warp == 1 && lane == 0) {
    for (int s = 0; s < Cfg::kStages; ++s) {
      mbar_init(&full_bar[s], 1);
      mbar_init(&empty_bar[s], 1);
    }
    for (int a = 0; a < Cfg::kAccStages; ++a) {
      mbar_init(&tmem_full_bar[a], 1);
      mbar_init(&tmem_empty_bar[a], 4 * G);    // one arrival per epilogue warp of every CTA of the group
    }
    mbar_init(prog_full, 4);
    mbar_init(prog_empty, 1);
    fence_mbar_init();
  }
  if (warp == 2) {
    if (G == 1) tmem_alloc<Cfg::kTmemCols>(tmem_ptr_smem);
    else tmem_alloc_cg2<Cfg::kTmemCols>(tmem_ptr_smem);
  }
  tc_fence_before();
  if (G == 2) {
    asm volatile("barrier.cluster.arrive.release.aligned;\n\tbarrier.cluster.wait.acquire.aligned;" ::: "memory");
  } else {
    __syncthreads();
  }
  tc_fence_after();
  const uint32_t tmem_base = *tmem_ptr_smem;
  // Programmatic dependent launch: a kernel enqueued behind this one with the programmatic-serialization attribute (the
  // progressive reduction, b200_peer_progressive_sum) may start once EVERY CTA of this grid has got here -- i.e. once the
  // whole persistent grid is resident, so that the dependent's CTAs can only fill in beside it, never take an SM a CTA of
  // this grid still needs.  Without such a dependent the instruction does nothing.
  asm volatile("griddepcontrol.launch_dependents;" ::: "memory");

  if (warp == 0) {
    // ===================== TMA producer (whole warp in the loop, one elected lane issues) =====================
    int stage = 0;
    uint32_t phase = 0;
    // (L2 eviction hints were tried and removed: evict_first on B made the B tiles shared by neighbouring CTAs
    //  miss, evict_last on A did not keep a 67 MB panel group resident -- profiles/r01_l2_hint_experiment.md.)
    bool k_down = false;
    for (int tile = worker; tile < num_tiles; tile += num_workers) {
      int m_blk, n_blk;
      tile_coords(tile, p.tiles_m, p.tiles_n, p.group_m, m_blk, n_blk);
      int m0 = scatter_row0<G>(p, m_blk, cta_rank);      // row of the product in "owner block, row inside it" order
      if (p.a_rows_per_block > 0) {
        const int q = m0 / p.a_rows_per_block, r = m0 % p.a_rows_per_block;
        m0 = p.a_blocks_inner > 0
                 ? (q % p.a_blocks_inner) * p.a_block_stride + (q / p.a_blocks_inner) * p.a_outer_stride + r
                 : q * p.a_block_stride + r;
      }
      const int n0 = n_blk * BN + cta_rank * Cfg::kBNcta;
      for (int kb = 0; kb < num_kb; ++kb) {
        mbar_wait(&empty_bar[stage], phase ^ 1);
        uint8_t* sa = smem + stage * Cfg::kStageBytes;
        uint8_t* sb = sa + Cfg::kABytes;
        const int k0 = (k_down ? num_kb - 1 - kb : kb) * kBK;   // the MMA side only sees ring slots: any K order works
        if (elect_one()) {
        if (G == 1) {
          mbar_arrive_expect_tx(&full_bar[stage], Cfg::kStageBytes);
          if (p.a_mn_major) {
#pragma unroll
            for (int c = 0; c < kBM / 64; ++c) tma_load_2d(sa + c * (kBK * 128), &tmA, &full_bar[stage], m0 + c * 64, k0);
          } else {
            tma_load_2d(sa, &tmA, &full_bar[stage], k0, m0);
          }
          if (kBKmajor) {
            tma_load_2d(sb, &tmB, &full_bar[stage], k0, n0);
          } else {
#pragma unroll
            for (int c = 0; c < Cfg::kBNcta / 64; ++c)
              tma_load_2d(sb + c * (kBK * 128), &tmB, &full_bar[stage], n0 + c * 64, k0);
          }
        } else {
          // both CTAs' bytes complete on the leader's barrier
          if (leader) mbar_arrive_expect_tx(&full_bar[stage], Cfg::kStageBytes * 2);
          if (p.a_mn_major) {
#pragma unroll
            for (int c = 0; c < kBM / 64; ++c)
              tma_load_2d_cg2(sa + c * (kBK * 128), &tmA, &full_bar[stage], m0 + c * 64, k0);
          } else {
            tma_load_2d_cg2(sa, &tmA, &full_bar[stage], k0, m0);
          }
          if (kBKmajor) {
            tma_load_2d_cg2(sb, &tmB, &full_bar[stage], k0, n0);
          } else {
#pragma unroll
            for (int c = 0; c < Cfg::kBNcta / 64; ++c)
              tma_load_2d_cg2(sb + c * (kBK * 128), &tmB, &full_bar[stage], n0 + c * 64, k0);
          }
        }
        }
        __syncwarp();
        if (++stage == p.num_stages) { stage = 0; phase ^= 1; }
      }
      if (p.k_serpentine) k_down = !k_down;
    }
  } else if (warp == 1 && leader) {
    // ===================== MMA issuer (leader CTA only) =====================
    const uint32_t idesc = p.a_mn_major ? make_idesc_bf16(kBM * G, BN, true, !kBKmajor)
                                        : make_idesc_bf16(kBM * G, BN, false, !kBKmajor);
    // Descriptors of stage 0, built once; per MMA only a constant is added (the issuing thread has 128 cycles
    // per MMA at most -- rebuilding two 64-bit descriptors each time cost ~20 % of the tensor pipe).
    //   A: K-major, 32 bytes of K per MMA inside the 128-byte swizzle row.
    //   B K-major: same.  B MN-major ([K,N] row-major): 16 K-rows of 128 B per MMA, 64-column chunks kBK*128 B apart.
    //   A MN-major ([K,M] row-major): like B MN-major, 64-row chunks kBK*128 B apart.
    const uint64_t a_desc0 = p.a_mn_major ? make_smem_desc_sw128(smem_u32(smem), kBK * 128, 1024)
                                          : make_smem_desc_sw128(smem_u32(smem), 16, 1024);
    const uint64_t b_desc0 = kBKmajor ? make_smem_desc_sw128(smem_u32(smem) + Cfg::kABytes, 16, 1024)
                                      : make_smem_desc_sw128(smem_u32(smem) + Cfg::kABytes, kBK * 128, 1024);
    const uint32_t kAStep = p.a_mn_major ? kUmmaK * 128 : kUmmaK * 2;
    constexpr uint32_t kBStep = kBKmajor ? kUmmaK * 2 : kUmmaK * 128;
    int stage = 0;
    uint32_t phase = 0;
    int acc = 0;
    uint32_t acc_phase = 0;
    for (int tile = worker; tile < num_tiles; tile += num_workers) {
      mbar_wait(&tmem_empty_bar[acc], acc_phase ^ 1);
      tc_fence_after();
      const uint32_t tmem_d = tmem_base + acc * BN;
      for (int kb = 0; kb < num_kb; ++kb) {
        mbar_wait(&full_bar[stage], phase);
        tc_fence_after();
        const uint64_t da = desc_advance(a_desc0, stage * Cfg::kStageBytes);
        const uint64_t db = desc_advance(b_desc0, stage * Cfg::kStageBytes);
        if (elect_one()) {
        if (G == 1) {
          if (kb == 0) mma_bf16_ss_c<false>(tmem_d, da, db, idesc);
          else mma_bf16_ss_c<true>(tmem_d, da, db, idesc);
#pragma unroll
          for (int k = 1; k < kBK / kUmmaK; ++k)
            mma_bf16_ss_c<true>(tmem_d, desc_advance(da, k * kAStep), desc_advance(db, k * kBStep), idesc);
          mma_commit(&empty_bar[stage]);          // frees this smem stage once the MMAs above have read it
          if (kb == num_kb - 1) mma_commit(&tmem_full_bar[acc]);
        } else {
          if (kb == 0) mma_bf16_ss_cg2_c<false>(tmem_d, da, db, idesc);
          else mma_bf16_ss_cg2_c<true>(tmem_d, da, db, idesc);
#pragma unroll
          for (int k = 1; k < kBK / kUmmaK; ++k)
            mma_bf16_ss_cg2_c<true>(tmem_d, desc_advance(da, k * kAStep), desc_advance(db, k * kBStep), idesc);
          mma_commit_cg2(&empty_bar[stage], 0x3);  // ... in both CTAs of the pair
          if (kb == num_kb - 1) mma_commit_cg2(&tmem_full_bar[acc], 0x3);
        }
        }
        __syncwarp();
        if (++stage == p.num_stages) { stage = 0; phase ^= 1; }
      }
      if (++acc == Cfg::kAccStages) { acc = 0; acc_phase ^= 1; }
    }
  } else if (warp == 3 && p.done_counters != nullptr) {
    // ===================== progress signals (see the epilogue) =====================
    const int per_group = p.group_m * p.tiles_n;
    int group = -1, pending = 0, flushes = 0;
    auto publish = [&]() {
      mbar_wait(prog_full, flushes & 1);
      if (lane == 0) {
#if B200_PROGRESS_FENCE == 2
        __threadfence_system();                                        // fence.sc.sys
#elif B200_PROGRESS_FENCE == 1
        asm volatile("fence.acq_rel.sys;" ::: "memory");
#endif
        const int rows_in_group = min(p.tiles_m - group * p.group_m, p.group_m);
        const unsigned int target = static_cast<unsigned int>(rows_in_group * p.tiles_n * G);   // (tile, CTA) pairs per launch
        unsigned int old;
        asm volatile("atom.acq_rel.gpu.global.add.u32 %0, [%1], %2;"
                     : "=r"(old) : "l"(p.done_counters + group), "r"(static_cast<unsigned int>(pending)) : "memory");
        if ((old + pending) % target == 0) {
          for (int o = 0; o < p.m_num_blocks; ++o) {
            unsigned int* flag = p.done_flag_ptrs[o] + p.done_sender * p.done_groups + group;
            asm volatile("st.release.sys.global.u32 [%0], %1;" ::"l"(flag), "r"(p.done_epoch) : "memory");
          }
        }
        mbar_arrive(prog_empty);
      }
      __syncwarp();
      ++flushes;
      pending = 0;
    };
    for (int tile = worker; tile < num_tiles; tile += num_workers) {
      const int g_tile = tile / per_group;
      if (g_tile != group && pending > 0) publish();
      group = g_tile;
      ++pending;
    }
    if (pending > 0) publish();
  } else if (warp >= 4) {
    // ===================== epilogue =====================
    const int q = warp & 3;                    // TMEM lane quadrant this warp may read
    int acc = 0;
    uint32_t acc_phase = 0;
    const bool out_f32 = p.out_f32 != 0;
    const bool has_bias = p.bias != nullptr;
    // Progress signals: when this CTA has stored its last tile of a raster group, the four epilogue warps hand the group
    // over to warp 3 (an mbarrier; they only wait for warp 3 to have finished the PREVIOUS hand-over, a group ago) and go on
    // with the next tile.  Warp 3 does the slow part: ONE system-scope fence per CTA and group -- which, by cumulativity
    // through the mbarrier, orders the tile stores of all four warps before what follows -- the count and, if it completes
    // the group, the flags.  (The fence takes as long as this SM's outstanding peer stores need to be acknowledged; inside an
    // epilogue warp it held back the next tile's accumulator: 0.557 instead of 0.462 ms per step at 8 GPUs.)
    int done_group = -1, done_pending = 0, done_flushes = 0;
    auto progress_flush = [&]() {
      if (done_flushes > 0) mbar_wait(prog_empty, (done_flushes - 1) & 1);
      __syncwarp();
      if (lane == 0) mbar_arrive(prog_full);
      ++done_flushes;
      done_pending = 0;
    };
    for (int tile = worker; tile < num_tiles; tile += num_workers) {
      int m_blk, n_blk;
      tile_coords(tile, p.tiles_m, p.tiles_n, p.group_m, m_blk, n_blk);
      if (p.done_counters != nullptr) {
        const int g_tile = tile / (p.group_m * p.tiles_n);
        if (g_tile != done_group && done_pending > 0) progress_flush();
        done_group = g_tile;
      }
      const int row = (m_blk * G + cta_rank) * kBM + q * 32 + lane;
      const int n0 = n_blk * BN;
      mbar_wait(&tmem_full_bar[acc], acc_phase);
      tc_fence_after();
      const uint32_t taddr = tmem_base + (static_cast<uint32_t>(q * 32) << 16) + acc * BN;
      if (out_f32) {
#pragma unroll 1
        for (int c = 0; c < BN / 32; ++c) {
          uint32_t r[32];
          tmem_ld_32x32(taddr + c * 32, r);
          tmem_wait_ld();
          const int col0 = n0 + c * 32;
          if (row < p.M && col0 < p.N) {
            float* dst = reinterpret_cast<float*>(p.C) + static_cast<int64_t>(row) * p.ldc + col0;
#pragma unroll
            for (int j = 0; j < 32; j += 4) {
              if (col0 + j < p.N) {
                float4 v;
                float b0 = 0.f, b1 = 0.f, b2 = 0.f, b3 = 0.f;
                if (has_bias) {
                  b0 = __bfloat162float(p.bias[col0 + j]);
                  b1 = __bfloat162float(p.bias[col0 + j + 1]);
                  b2 = __bfloat162float(p.bias[col0 + j + 2]);
                  b3 = __bfloat162float(p.bias[col0 + j + 3]);
                }
                v.x = apply_epilogue(__uint_as_float(r[j]), b0, p.epilogue, true);
                v.y = apply_epilogue(__uint_as_float(r[j + 1]), b1, p.epilogue, true);
                v.z = apply_epilogue(__uint_as_float(r[j + 2]), b2, p.epilogue, true);
                v.w = apply_epilogue(__uint_as_float(r[j + 3]), b3, p.epilogue, true);
                *reinterpret_cast<float4*>(dst + j) = v;
              }
            }
          }
        }
      } else {
        // bf16: 64 columns at a time through this warp's 4 KB staging tile (128-byte rows, 16-byte units
        // XOR-swizzled by row), so every warp store instruction writes 4 rows x 128 contiguous bytes --
        // full sectors in local HBM and full-size packets when the destination is a peer GPU.
        const int row_base = (m_blk * G + cta_rank) * kBM + q * 32;       // logical row of lane 0
        __nv_bfloat16* cbase;
        if (p.m_rows_per_block > 0) {
          const int prow = scatter_row0<G>(p, m_blk, cta_rank) + q * 32;  // (owner block) * rows_per_block + row inside it
          const int blk = prow / p.m_rows_per_block;
          const int local = prow % p.m_rows_per_block;
          cbase = p.c_block_ptrs != nullptr
                      ? reinterpret_cast<__nv_bfloat16*>(p.c_block_ptrs[blk]) + static_cast<int64_t>(local) * p.ldc
                      : reinterpret_cast<__nv_bfloat16*>(p.C) +
                            (static_cast<int64_t>(blk) * p.m_rows_per_block + local) * p.ldc;
        } else {
          cbase = reinterpret_cast<__nv_bfloat16*>(p.C) + static_cast<int64_t>(row_base) * p.ldc;
        }
        uint8_t* stg = store_stage + q * (32 * 128);
#pragma unroll 1
        for (int c2 = 0; c2 < BN / 64; ++c2) {
          const int col0 = n0 + c2 * 64;
#pragma unroll
          for (int half = 0; half < 2; ++half) {
            uint32_t r[32];
            tmem_ld_32x32(taddr + c2 * 64 + half * 32, r);
            tmem_wait_ld();
#pragma unroll
            for (int j = 0; j < 32; j += 8) {
              const int col = col0 + half * 32 + j;
              float bv[8];
#pragma unroll
              for (int t = 0; t < 8; ++t) bv[t] = 0.f;
              if (has_bias && col < p.N) {      // N % 8 == 0: a group of 8 columns is all in or all out
                uint4 braw = *reinterpret_cast<const uint4*>(p.bias + col);
                const __nv_bfloat162* b2 = reinterpret_cast<const __nv_bfloat162*>(&braw);
#pragma unroll
                for (int t = 0; t < 4; ++t) {
                  float2 f = __bfloat1622float2(b2[t]);
                  bv[2 * t] = f.x;
                  bv[2 * t + 1] = f.y;
                }
              }
              uint4 o;
              o.x = pack_bf16x2(apply_epilogue(__uint_as_float(r[j + 0]), bv[0], p.epilogue, false),
                                apply_epilogue(__uint_as_float(r[j + 1]), bv[1], p.epilogue, false));
              o.y = pack_bf16x2(apply_epilogue(__uint_as_float(r[j + 2]), bv[2], p.epilogue, false),
                                apply_epilogue(__uint_as_float(r[j + 3]), bv[3], p.epilogue, false));
              o.z = pack_bf16x2(apply_epilogue(__uint_as_float(r[j + 4]), bv[4], p.epilogue, false),
                                apply_epilogue(__uint_as_float(r[j + 5]), bv[5], p.epilogue, false));
              o.w = pack_bf16x2(apply_epilogue(__uint_as_float(r[j + 6]), bv[6], p.epilogue, false),
                                apply_epilogue(__uint_as_float(r[j + 7]), bv[7], p.epilogue, false));
              const int unit = half * 4 + (j >> 3);
              *reinterpret_cast<uint4*>(stg + lane * 128 + ((unit ^ (lane & 7)) << 4)) = o;
            }
          }
          __syncwarp();
#pragma unroll
          for (int it = 0; it < 8; ++it) {
            const int rr = it * 4 + (lane >> 3);
            const int unit = lane & 7;
            const uint4 v = *reinterpret_cast<const uint4*>(stg + rr * 128 + ((unit ^ (rr & 7)) << 4));
            const int gcol = col0 + unit * 8;
            if (row_base + rr < p.M && gcol < p.N)
              *reinterpret_cast<uint4*>(cbase + static_cast<int64_t>(rr) * p.ldc + gcol) = v;
          }
          __syncwarp();
        }
      }
      // accumulator stage drained: hand it back to the MMA issuer (in the leader CTA)
      tc_fence_before();
      __syncwarp();
      if (lane == 0) {
        if (G == 1) mbar_arrive(&tmem_empty_bar[acc]);
        else mbar_arrive_cluster(&tmem_empty_bar[acc], 0);
      }
      if (++acc == Cfg::kAccStages) { acc = 0; acc_phase ^= 1; }
      if (p.done_counters != nullptr) ++done_pending;     // counted at the next raster-group boundary (progress_flush)
    }
    if (done_pending > 0) progress_flush();
  }

  // ===================== teardown =====================
  __syncwarp();
  tc_fence_before();
  if (G == 2) {
    asm volatile("barrier.cluster.arrive.release.aligned;\n\tbarrier.cluster.wait.acquire.aligned;" ::: "memory");
  } else {
    __syncthreads();
  }
  if (warp == 2) {
    tc_fence_after();
    if (G == 1) tmem_dealloc<Cfg::kTmemCols>(tmem_base);
    else tmem_dealloc_cg2<Cfg::kTmemCols>(tmem_base);
  }
}

// ------------------------------------------------------------------ host side
// No tuning state lives here: overrides of the heuristics arrive with the call (b200_gemm_args.tune).
template <int BN, int G, bool kBKmajor>
static int launch_gemm(const void* A, int64_t lda, const void* B, int64_t ldb, const GemmParams& p0,
                       const b200_gemm_tuning& tune, cudaStream_t stream) {
  using Cfg = GemmCfg<BN, G>;
  GemmParams p = p0;
  p.tiles_m = (p.M + kBM * G - 1) / (kBM * G);
  p.tiles_n = (p.N + BN - 1) / BN;
  // Ring depth: measured on B200 (profiles/r01_gemm_stage_sweep.log), CTA pairs run fastest with FOUR
  // 32 KB stages (1348 / 1417 TFLOP/s) and lose 5-15 % with five or six although the smem is there: a deeper
  // ring lets the CTAs of a wave drift apart in K, which costs L2 locality on the shared A/B panels.
  int stages = Cfg::kStages;
  if (G == 2 && stages > 4) stages = 4;
  if (tune.max_stages > 0 && tune.max_stages < stages) stages = tune.max_stages;
  p.num_stages = stages;
  p.k_serpentine = tune.k_serpentine ? 1 : 0;
  if (tune.group_m > 0) {
    p.group_m = tune.group_m;
  } else {
    // Raster group (row-blocks that sweep all column-blocks together).  Measured with ncu on the two FFN GEMMs
    // (profiles/r01_gemm_dram_traffic_vs_raster.log): at K = 4096 group 16 reads 0.95 GB from DRAM (8: 1.29,
    // 32: 2.36, 4: 2.40; 0.27 GB algorithmic) and is also the fastest; at K = 16384 the panels of one wave
    // alone are ~145 MB > L2 and every group size reads 3.2-4 GB, so the choice is flat.
    p.group_m = p.K <= 8192 ? 16 : 8;
  }
  CUtensorMap tmA, tmB;
  {
    uint64_t a_rows = static_cast<uint64_t>(p.M);
    if (p.a_rows_per_block > 0) {
      const uint64_t nblk = static_cast<uint64_t>((p.M + p.a_rows_per_block - 1) / p.a_rows_per_block);
      if (p.a_blocks_inner > 0)
        a_rows = static_cast<uint64_t>(p.a_blocks_inner - 1) * p.a_block_stride +
                 (nblk / p.a_blocks_inner - 1) * p.a_outer_stride + p.a_rows_per_block;
      else
        a_rows = (nblk - 1) * p.a_block_stride + p.a_rows_per_block;
    }
    uint64_t dims[2] = {static_cast<uint64_t>(p.K), a_rows};
    uint64_t strides[2] = {0, static_cast<uint64_t>(lda) * 2};
    uint32_t box[2] = {kBK, kBM};
    if (p.a_mn_major) {          // stored [K, M]: M is the contiguous dimension
      dims[0] = static_cast<uint64_t>(p.M);
      dims[1] = static_cast<uint64_t>(p.K);
      box[0] = 64;
      box[1] = kBK;
    }
    int rc = make_tensor_map(&tmA, A, 2, 2, dims, strides, box, 3);
    if (rc != B200_OK) return rc;
  }
  if (kBKmajor) {
    uint64_t dims[2] = {static_cast<uint64_t>(p.K), static_cast<uint64_t>(p.N)};
    uint64_t strides[2] = {0, static_cast<uint64_t>(ldb) * 2};
    uint32_t box[2] = {kBK, static_cast<uint32_t>(Cfg::kBNcta)};
    int rc = make_tensor_map(&tmB, B, 2, 2, dims, strides, box, 3);
    if (rc != B200_OK) return rc;
  } else {
    uint64_t dims[2] = {static_cast<uint64_t>(p.N), static_cast<uint64_t>(p.K)};
    uint64_t strides[2] = {0, static_cast<uint64_t>(ldb) * 2};
    uint32_t box[2] = {64, kBK};
    int rc = make_tensor_map(&tmB, B, 2, 2, dims, strides, box, 3);
    if (rc != B200_OK) return rc;
  }
  auto kern = gemm_bf16_kernel<BN, G, kBKmajor>;
  {
    // the opt-in to > 48 KB of dynamic shared memory is per device and per kernel instantiation: do it once each
    static std::atomic<bool> smem_opted_in[64];
    int dev = 0;
    B200_CUDA_CHECK(cudaGetDevice(&dev));
    if (dev < 0 || dev >= 64 || !smem_opted_in[dev].load(std::memory_order_acquire)) {
      B200_CUDA_CHECK(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, Cfg::kSmemBytes));
      if (dev >= 0 && dev < 64) smem_opted_in[dev].store(true, std::memory_order_release);
    }
  }
  const int sms = num_sms();
  const int tiles = p.tiles_m * p.tiles_n;
  int workers = sms / G;
  if (tune.max_ctas > 0 && tune.max_ctas / G < workers) workers = tune.max_ctas / G > 0 ? tune.max_ctas / G : 1;
  if (tiles < workers) workers = tiles;
  cudaLaunchConfig_t cfg = {};
  cfg.gridDim = dim3(workers * G);
  cfg.blockDim = dim3(kGemmThreads);
  cfg.dynamicSmemBytes = Cfg::kSmemBytes;
  cfg.stream = stream;
  cudaLaunchAttribute attrs[1];
  attrs[0].id = cudaLaunchAttributeClusterDimension;
  attrs[0].val.clusterDim.x = G;
  attrs[0].val.clusterDim.y = 1;
  attrs[0].val.clusterDim.z = 1;
  cfg.attrs = attrs;
  cfg.numAttrs = 1;
  B200_CUDA_CHECK(cudaLaunchKernelEx(&cfg, kern, tmA, tmB, p));
  launch_counter().fetch_add(1);
  return B200_OK;
}

template <int BN, int G>
static int launch_gemm_layout(bool kmajor, const void* A, int64_t lda, const void* B, int64_t ldb,
                              const GemmParams& p, const b200_gemm_tuning& tune, cudaStream_t stream) {
  return kmajor ? launch_gemm<BN, G, true>(A, lda, B, ldb, p, tune, stream)
                : launch_gemm<BN, G, false>(A, lda, B, ldb, p, tune, stream);
}

}  // namespace b200

using namespace b200;

extern "C" int b200_gemm_bf16(const void* A, int64_t lda, const void* B, int64_t ldb, int b_layout,
                              const void* bias, void* C, int64_t ldc, int out_dtype, int M, int N, int K,
                              int epilogue, void* stream) {
  b200_gemm_args a = {};
  a.A = A; a.lda = lda; a.B = B; a.ldb = ldb; a.b_layout = b_layout; a.bias = bias; a.C = C; a.ldc = ldc;
  a.out_dtype = out_dtype; a.M = M; a.N = N; a.K = K; a.epilogue = epilogue;
  return b200_gemm_bf16_ex(&a, stream);
}

extern "C" int b200_gemm_bf16_rowblocked(const void* A, int64_t lda, int a_rows_per_block, int a_block_stride,
                                         const void* B, int64_t ldb, int b_layout, const void* bias, void* C,
                                         int64_t ldc, int out_dtype, int M, int N, int K, int epilogue,
                                         void* stream) {
  b200_gemm_args a = {};
  a.A = A; a.lda = lda; a.a_rows_per_block = a_rows_per_block; a.a_block_stride = a_block_stride;
  a.B = B; a.ldb = ldb; a.b_layout = b_layout; a.bias = bias; a.C = C; a.ldc = ldc;
  a.out_dtype = out_dtype; a.M = M; a.N = N; a.K = K; a.epilogue = epilogue;
  return b200_gemm_bf16_ex(&a, stream);
}

extern "C" int b200_gemm_bf16_ex(const b200_gemm_args* args, void* stream) {
  B200_REQUIRE(args != nullptr, "b200_gemm_bf16: null argument block");
  const void* A = args->A;
  const void* B = args->B;
  const void* bias = args->bias;
  void* C = args->C;
  const int64_t lda = args->lda, ldb = args->ldb, ldc = args->ldc;
  const int b_layout = args->b_layout, out_dtype = args->out_dtype, epilogue = args->epilogue;
  const int M = args->M, N = args->N, K = args->K;
  const int a_rows_per_block = args->a_rows_per_block, a_block_stride = args->a_block_stride;
  const bool scatter = args->c_block_ptrs != nullptr;
  B200_REQUIRE(A && B && (C || scatter), "b200_gemm_bf16: null operand pointer");
  B200_REQUIRE(a_rows_per_block >= 0 && a_block_stride >= 0, "b200_gemm_bf16: negative row-block parameters");
  B200_REQUIRE(a_rows_per_block == 0 || (a_rows_per_block % 128 == 0 && a_block_stride >= a_rows_per_block),
               "b200_gemm_bf16: a_rows_per_block must be a multiple of 128 and <= a_block_stride");
  B200_REQUIRE(args->m_rows_per_block >= 0 && args->m_block_rot >= 0, "b200_gemm_bf16: negative m-block parameters");
  if (args->m_rows_per_block > 0) {
    B200_REQUIRE(args->m_rows_per_block % 256 == 0 && M % args->m_rows_per_block == 0,
                 "b200_gemm_bf16: m_rows_per_block must be a multiple of 256 that divides M");
    B200_REQUIRE(out_dtype == B200_BF16, "b200_gemm_bf16: row-block scatter writes bf16 only");
    B200_REQUIRE(a_rows_per_block == 0 || a_rows_per_block == args->m_rows_per_block ||
                     (args->a_blocks_inner > 0 && args->m_rows_per_block == a_rows_per_block * args->a_blocks_inner),
                 "b200_gemm_bf16: with the row-block scatter, a_rows_per_block must be 0, equal m_rows_per_block, or "
                 "(two-level view) m_rows_per_block / a_blocks_inner");
  } else {
    B200_REQUIRE(!scatter, "b200_gemm_bf16: c_block_ptrs needs m_rows_per_block");
  }
  B200_REQUIRE(args->a_blocks_inner >= 0 && args->a_outer_stride >= 0, "b200_gemm_bf16: negative two-level view parameters");
  if (args->a_blocks_inner > 0)
    B200_REQUIRE(a_rows_per_block > 0 && M % (a_rows_per_block * args->a_blocks_inner) == 0 &&
                     args->a_outer_stride >= a_rows_per_block,
                 "b200_gemm_bf16: the two-level A view needs a_rows_per_block > 0, M a multiple of a_rows_per_block * "
                 "a_blocks_inner and a_outer_stride >= a_rows_per_block");
  B200_REQUIRE(M > 0 && N > 0 && K > 0, "b200_gemm_bf16: M, N, K must be positive (got %d, %d, %d)", M, N, K);
  B200_REQUIRE(K % 8 == 0 && N % 8 == 0, "b200_gemm_bf16: K and N must be multiples of 8 (got K=%d N=%d)", K, N);
  B200_REQUIRE(b_layout == B200_B_KN || b_layout == B200_B_NK, "b200_gemm_bf16: bad b_layout %d", b_layout);
  B200_REQUIRE(out_dtype == B200_BF16 || out_dtype == B200_F32, "b200_gemm_bf16: bad out_dtype %d", out_dtype);
  B200_REQUIRE(epilogue >= B200_EPI_NONE && epilogue <= B200_EPI_GELU_TANH, "b200_gemm_bf16: bad epilogue %d",
               epilogue);
  const bool needs_bias =
      epilogue == B200_EPI_BIAS || epilogue == B200_EPI_BIAS_RELU || epilogue == B200_EPI_BIAS_GELU_TANH;
  B200_REQUIRE(!needs_bias || bias, "b200_gemm_bf16: epilogue %d needs a bias vector", epilogue);
  const int64_t min_ldb = b_layout == B200_B_KN ? N : K;
  const bool a_km = args->a_layout == B200_A_KM;
  B200_REQUIRE(args->a_layout == B200_A_MK || a_km, "b200_gemm_bf16: bad a_layout %d", args->a_layout);
  B200_REQUIRE(!a_km || (a_rows_per_block == 0 && args->m_rows_per_block == 0 && M % 8 == 0),
               "b200_gemm_bf16: a_layout KM needs M %% 8 == 0 and excludes the row-blocked / row-block-scatter options");
  B200_REQUIRE(lda >= (a_km ? M : K) && ldb >= min_ldb && ldc >= N, "b200_gemm_bf16: leading dimension smaller than row");
  B200_REQUIRE(lda % 8 == 0 && ldb % 8 == 0 && ldc % 8 == 0, "b200_gemm_bf16: lda/ldb/ldc must be multiples of 8");
  B200_REQUIRE((reinterpret_cast<uintptr_t>(A) | reinterpret_cast<uintptr_t>(B) | reinterpret_cast<uintptr_t>(C) |
                reinterpret_cast<uintptr_t>(bias)) % 16 == 0,
               "b200_gemm_bf16: operands must be 16-byte aligned");

  GemmParams p = {};
  p.C = C;
  p.bias = needs_bias ? reinterpret_cast<const __nv_bfloat16*>(bias) : nullptr;
  p.M = M;
  p.N = N;
  p.K = K;
  p.ldc = ldc;
  p.epilogue = epilogue;
  p.out_f32 = out_dtype == B200_F32;
  p.a_rows_per_block = a_rows_per_block;
  p.a_block_stride = a_block_stride;
  p.a_blocks_inner = args->a_blocks_inner;
  p.a_outer_stride = args->a_outer_stride;
  p.c_block_ptrs = args->c_block_ptrs;
  p.m_rows_per_block = args->m_rows_per_block;
  p.m_block_rot = args->m_block_rot;
  p.m_num_blocks = args->m_rows_per_block > 0 ? M / args->m_rows_per_block : 0;
  p.m_interleave = (args->m_block_interleave && args->m_rows_per_block > 0) ? 1 : 0;
  if (args->done_counters != nullptr) {
    B200_REQUIRE(args->m_rows_per_block > 0 && args->done_flag_ptrs != nullptr && args->done_groups > 0 &&
                     args->done_sender >= 0 && args->tune.group_m > 0 && args->tune.cta_group > 0 && args->tune.block_n > 0,
                 "b200_gemm_bf16: progress signals need the row-block scatter, the owners' flag arrays, and an explicit tile "
                 "shape and raster group in `tune` (the caller sizes its groups from them)");
    p.done_counters = args->done_counters;
    p.done_flag_ptrs = args->done_flag_ptrs;
    p.done_sender = args->done_sender;
    p.done_groups = args->done_groups;
    p.done_epoch = args->done_epoch;
  }
  p.a_mn_major = a_km ? 1 : 0;
  cudaStream_t st = reinterpret_cast<cudaStream_t>(stream);
  const bool kmajor = b_layout == B200_B_NK;

  // Tile choice: the largest tile that still gives every SM work; CTA pairs once the
  // problem has at least one 256 x BN tile per pair.
  const int sms = num_sms();
  int bn = 256;
  auto tiles_for = [&](int bm, int bnn) { return ((M + bm - 1) / bm) * ((N + bnn - 1) / bnn); };
  if (N <= 64 || tiles_for(128, 128) < sms / 2) bn = 64;
  else if (N <= 128 || tiles_for(128, 256) < sms) bn = 128;
  int g = (bn >= 128 && tiles_for(256, bn) >= sms / 2) ? 2 : 1;
  const b200_gemm_tuning& tune = args->tune;
  B200_REQUIRE(tune.cta_group >= 0 && tune.cta_group <= 2, "b200_gemm_bf16: tune.cta_group must be 0, 1 or 2");
  B200_REQUIRE(tune.block_n == 0 || tune.block_n == 64 || tune.block_n == 128 || tune.block_n == 256,
               "b200_gemm_bf16: tune.block_n must be 0, 64, 128 or 256");
  B200_REQUIRE(tune.group_m >= 0 && tune.group_m <= 1024 && tune.max_stages >= 0 && tune.max_stages <= 8 &&
                   tune.max_ctas >= 0 && tune.max_ctas <= 4096,
               "b200_gemm_bf16: tune.group_m / max_stages / max_ctas out of range");
  if (tune.block_n) bn = tune.block_n;
  if (tune.cta_group) g = tune.cta_group;
  if (g == 2 && bn == 64) g = 1;   // a CTA would load 32 B columns: not a whole swizzle chunk

  if (g == 2) {
    if (bn == 256) return launch_gemm_layout<256, 2>(kmajor, A, lda, B, ldb, p, tune, st);
    return launch_gemm_layout<128, 2>(kmajor, A, lda, B, ldb, p, tune, st);
  }
  if (bn == 256) return launch_gemm_layout<256, 1>(kmajor, A, lda, B, ldb, p, tune, st);
  if (bn == 128) return launch_gemm_layout<128, 1>(kmajor, A, lda, B, ldb, p, tune, st);
  return launch_gemm_layout<64, 1>(kmajor, A, lda, B, ldb, p, tune, st);
}
