// Micro-benchmark: cycles per tcgen05.mma (kind::f16, bf16 in, f32 accumulate, M = 128, K = 16) on sm_100a,
// as a function of N, operand source (A from smem or TMEM), B major-ness, accumulator rotation and of what
// the other warps of the CTA do meanwhile (TMEM loads, MUFU, shared-memory traffic).  It answers one question
// for csrc/attention*.cu: what does a 128x128x16 MMA cost when nothing else is in its way?
// Build: tools/build_mma_rate.sh ; run on the GPU box: tools/mma_rate  -> table on stdout.
#include <cstdio>
#include <cstdint>
#include <cuda_runtime.h>
#include <cuda_bf16.h>
#include "../learning-jax-sharding_b200/csrc/sm100_ptx.cuh"
using namespace b200;

struct RateCfg {
  int n;          // MMA N
  int a_tmem;     // 1: A operand from TMEM (TS form)
  int b_mn;       // 1: B operand MN-major (the P.V layout of V), 0: K-major
  int nacc;       // accumulators rotated over (1, 2)
  int mix;        // 1: alternate groups of 8 MMAs: SS K-major (acc 0) / TS MN-major (acc 1), like attention v5
  int bg;         // background in warps 4..11: 0 none, 1 tcgen05.ld, 2 ex2, 3 shared ld/st, 4 tcgen05.st
  int iters;      // groups of 4 MMAs (one 64-wide K block)
  int cg2;        // 1: CTA pair, cta_group::2 MMAs with M = 256 (n = N of the pair MMA; each CTA holds N/2 of B)
};

template <bool kPair>
__global__ void __launch_bounds__(384, 1) mma_rate_kernel(RateCfg cfg, unsigned long long* out) {
  extern __shared__ __align__(1024) unsigned char smem_raw[];
  unsigned char* smem = reinterpret_cast<unsigned char*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));
  unsigned char* sA = smem;                 // 128 x 64 bf16, K-major, 16 KB
  unsigned char* sB = smem + 16384;         // up to 256 x 64 bf16 = 32 KB (either major-ness)
  unsigned char* sX = smem + 49152;         // 48 KB scratch for background shared-memory traffic
  __shared__ uint64_t done_bar;
  __shared__ uint32_t tmem_base_s;
  __shared__ volatile int stop_flag;
  const int warp = uniform_warp_idx();
  if (threadIdx.x == 0) { mbar_init(&done_bar, 1); fence_mbar_init(); stop_flag = 0; }
  for (int i = threadIdx.x; i < 49152 / 4; i += blockDim.x) reinterpret_cast<uint32_t*>(smem)[i] = 0x3c003c00u + i;
  fence_async_smem();
  if (warp == 2) { if (kPair) tmem_alloc_cg2<512>(&tmem_base_s); else tmem_alloc<512>(&tmem_base_s); }
  tc_fence_before();
  __syncthreads();
  uint32_t cta_rank = 0;
  if (kPair) {
    asm volatile("mov.u32 %0, %%cluster_ctarank;" : "=r"(cta_rank));
    asm volatile("barrier.cluster.arrive.release.aligned;\n\tbarrier.cluster.wait.acquire.aligned;" ::: "memory");
  }
  tc_fence_after();
  const uint32_t tmem = tmem_base_s;

  if (warp == 1 && kPair) {
    // pair: the leader issues M = 256 MMAs, both CTAs wait for the multicast commit
    const uint32_t idesc_k = make_idesc_bf16(256, cfg.n, false, false);
    const uint32_t idesc_mn = make_idesc_bf16(256, cfg.n, false, true);
    const uint64_t a_desc = make_smem_desc_sw128(smem_u32(sA), 16, 1024);
    const uint64_t bk_desc = make_smem_desc_sw128(smem_u32(sB), 16, 1024);
    const uint64_t bmn_desc = make_smem_desc_sw128(smem_u32(sB), 16384, 1024);
    const uint32_t a_tmem = tmem + 448;
    __syncwarp();
    const long long t0 = clock64();
    if (cta_rank == 0) {
      for (int it = 0; it < cfg.iters; ++it) {
        const bool pv = cfg.mix ? ((it >> 1) & 1) : (cfg.a_tmem != 0);
        if (elect_one()) {
#pragma unroll
          for (int k = 0; k < 4; ++k) {
            if (!pv) mma_bf16_ss_cg2(tmem, desc_advance(a_desc, k * 32), cfg.b_mn && !cfg.mix ? desc_advance(bmn_desc, k * 2048) : desc_advance(bk_desc, k * 32),
                                     cfg.b_mn && !cfg.mix ? idesc_mn : idesc_k, 1);
            else mma_bf16_ts_cg2_c<true>(tmem + 256, a_tmem + k * 8, (cfg.b_mn || cfg.mix) ? desc_advance(bmn_desc, k * 2048) : desc_advance(bk_desc, k * 32),
                                         (cfg.b_mn || cfg.mix) ? idesc_mn : idesc_k);
          }
        }
        __syncwarp();
      }
      if (elect_one()) mma_commit_cg2(&done_bar, 0x3);
      __syncwarp();
    }
    mbar_wait(&done_bar, 0);
    const long long t1 = clock64();
    stop_flag = 1;
    if (lane_id() == 0) out[blockIdx.x] = static_cast<unsigned long long>(t1 - t0);
  } else if (warp == 1) {
    const uint32_t idesc_k = make_idesc_bf16(128, cfg.n, false, false);
    const uint32_t idesc_mn = make_idesc_bf16(128, cfg.n, false, true);
    const uint64_t a_desc = make_smem_desc_sw128(smem_u32(sA), 16, 1024);
    const uint64_t bk_desc = make_smem_desc_sw128(smem_u32(sB), 16, 1024);
    const uint64_t bmn_desc = make_smem_desc_sw128(smem_u32(sB), 16384, 1024);
    const uint32_t a_tmem = tmem + 448;   // 32 columns of bf16 pairs = 4 K steps
    long long t0 = 0, t1 = 0;
    __syncwarp();
    t0 = clock64();
    for (int it = 0; it < cfg.iters; ++it) {
      const uint32_t acc = tmem + (cfg.nacc == 2 ? (it & 1) * 128 : 0);
      if (cfg.mix) {
        // 8 MMAs "Q.K^T" (SS, K-major B) then 8 MMAs "P.V" (TS, MN-major B), separate accumulators
        const bool pv = (it >> 1) & 1;
        if (elect_one()) {
#pragma unroll
          for (int k = 0; k < 4; ++k) {
            if (!pv) mma_bf16_ss(tmem, desc_advance(a_desc, k * 32), desc_advance(bk_desc, k * 32), idesc_k, 1);
            else mma_bf16_ts(tmem + 256, a_tmem + k * 8, desc_advance(bmn_desc, k * 2048), idesc_mn, 1);
          }
        }
      } else if (elect_one()) {
#pragma unroll
        for (int k = 0; k < 4; ++k) {
          const uint64_t bd = cfg.b_mn ? desc_advance(bmn_desc, k * 2048) : desc_advance(bk_desc, k * 32);
          const uint32_t id = cfg.b_mn ? idesc_mn : idesc_k;
          if (cfg.a_tmem) mma_bf16_ts(acc, a_tmem + k * 8, bd, id, 1);
          else mma_bf16_ss(acc, desc_advance(a_desc, k * 32), bd, id, 1);
        }
      }
      __syncwarp();
    }
    if (elect_one()) mma_commit(&done_bar);
    __syncwarp();
    mbar_wait(&done_bar, 0);
    t1 = clock64();
    stop_flag = 1;
    if (lane_id() == 0) out[blockIdx.x] = static_cast<unsigned long long>(t1 - t0);
  } else if (warp >= 4 && cfg.bg != 0) {
    const uint32_t lane_base = tmem + (static_cast<uint32_t>((warp & 3) * 32) << 16);
    uint32_t r[32];
#pragma unroll
    for (int i = 0; i < 32; ++i) r[i] = threadIdx.x + i;
    float f = static_cast<float>(threadIdx.x) * 1e-3f;
    float g0 = f, g1 = f + 1.f, g2 = f + 2.f, g3 = f + 3.f;
    uint4* sx = reinterpret_cast<uint4*>(sX) + (threadIdx.x - 128);
    while (stop_flag == 0) {
      if (cfg.bg == 1) {
#pragma unroll 1
        for (int j = 0; j < 8; ++j) { tmem_ld_32x32(lane_base + 256 + 32 * (j & 3), r); tmem_wait_ld(); }
      } else if (cfg.bg == 2) {
#pragma unroll
        for (int j = 0; j < 32; ++j) {
          g0 = exp2_approx(g0 * 0.5f - 1.f); g1 = exp2_approx(g1 * 0.5f - 1.f);
          g2 = exp2_approx(g2 * 0.5f - 1.f); g3 = exp2_approx(g3 * 0.5f - 1.f);
        }
      } else if (cfg.bg == 3) {
#pragma unroll
        for (int j = 0; j < 8; ++j) { uint4 v = sx[j * 256]; v.x += 1; sx[j * 256] = v; }
      } else if (cfg.bg == 4) {
#pragma unroll 1
        for (int j = 0; j < 8; ++j) { tmem_st_32x32(lane_base + 384 + 32 * (j & 1), r); tmem_wait_st(); }
      }
    }
    if (g0 + g1 + g2 + g3 + static_cast<float>(r[0]) == 12345.678f) out[0] = 0;
  }
  tc_fence_before();
  __syncthreads();
  if (kPair) asm volatile("barrier.cluster.arrive.release.aligned;\n\tbarrier.cluster.wait.acquire.aligned;" ::: "memory");
  if (warp == 2) { if (kPair) tmem_dealloc_cg2<512>(tmem); else tmem_dealloc<512>(tmem); }
}

static const char* bg_name(int bg) {
  switch (bg) { case 0: return "none"; case 1: return "tmem.ld x8w"; case 2: return "ex2 x8w"; case 3: return "smem ld/st x8w";
                case 4: return "tmem.st x8w"; }
  return "?";
}

int main() {
  int dev = 0; cudaSetDevice(dev);
  cudaDeviceProp prop; cudaGetDeviceProperties(&prop, dev);
  printf("device %s, %d SMs\n", prop.name, prop.multiProcessorCount);
  const int smem_bytes = 49152 + 49152 + 1024;
  cudaFuncSetAttribute(mma_rate_kernel<false>, cudaFuncAttributeMaxDynamicSharedMemorySize, smem_bytes);
  cudaFuncSetAttribute(mma_rate_kernel<true>, cudaFuncAttributeMaxDynamicSharedMemorySize, smem_bytes);
  unsigned long long* d_out; cudaMalloc(&d_out, sizeof(unsigned long long) * 1024);
  unsigned long long h_out[1024];
  const RateCfg cfgs[] = {
      {128, 0, 0, 1, 0, 0, 512, 0}, {128, 0, 1, 1, 0, 0, 512, 0}, {128, 1, 0, 1, 0, 0, 512, 0}, {128, 1, 1, 1, 0, 0, 512, 0},
      {128, 0, 0, 2, 0, 0, 512, 0}, {128, 1, 1, 2, 0, 0, 512, 0}, {128, 0, 0, 1, 1, 0, 512, 0},
      {256, 0, 0, 1, 0, 0, 512, 0}, {256, 1, 0, 1, 0, 0, 512, 0}, {64, 0, 0, 1, 0, 0, 512, 0}, {64, 1, 1, 1, 0, 0, 512, 0},
      {128, 0, 0, 1, 0, 1, 512, 0}, {128, 0, 0, 1, 0, 2, 512, 0}, {128, 0, 0, 1, 0, 3, 512, 0}, {128, 0, 0, 1, 0, 4, 512, 0},
      {128, 1, 1, 1, 0, 1, 512, 0}, {128, 1, 1, 1, 0, 2, 512, 0}, {128, 1, 1, 1, 0, 3, 512, 0}, {128, 1, 1, 1, 0, 4, 512, 0},
      {128, 0, 0, 1, 1, 1, 512, 0}, {128, 0, 0, 1, 1, 2, 512, 0}, {128, 0, 0, 1, 1, 4, 512, 0},
      {256, 0, 0, 1, 0, 0, 512, 1}, {128, 0, 0, 1, 0, 0, 512, 1}, {128, 1, 1, 1, 0, 0, 512, 1}, {128, 0, 0, 1, 1, 0, 512, 1},
      {128, 0, 0, 1, 1, 1, 512, 1}, {128, 0, 0, 1, 1, 2, 512, 1}, {64, 1, 1, 1, 0, 0, 512, 1}, {256, 1, 1, 1, 0, 0, 512, 1},
  };
  printf("%-5s %-6s %-6s %-5s %-4s %-16s | %-22s | %-22s\n", "N", "A", "B", "nacc", "mix", "background", "1 CTA: cycles/MMA",
         "148 CTAs: cyc/MMA (max)");
  for (const RateCfg& c : cfgs) {
    double res[2] = {0, 0}, resmax = 0;
    for (int g = 0; g < 2; ++g) {
      const int grid = g == 0 ? 1 : prop.multiProcessorCount;
      for (int rep = 0; rep < 2; ++rep) {
        {
          cudaLaunchConfig_t lc = {};
          lc.gridDim = dim3(c.cg2 ? (g == 0 ? 2 : (prop.multiProcessorCount / 2) * 2) : grid);
          lc.blockDim = dim3(384);
          lc.dynamicSmemBytes = smem_bytes;
          cudaLaunchAttribute at[1];
          at[0].id = cudaLaunchAttributeClusterDimension;
          at[0].val.clusterDim.x = c.cg2 ? 2 : 1; at[0].val.clusterDim.y = 1; at[0].val.clusterDim.z = 1;
          lc.attrs = at; lc.numAttrs = 1;
          if (c.cg2) cudaLaunchKernelEx(&lc, mma_rate_kernel<true>, c, d_out);
          else cudaLaunchKernelEx(&lc, mma_rate_kernel<false>, c, d_out);
        }
        cudaError_t e = cudaGetLastError();
        if (e != cudaSuccess) { printf("launch error: %s\n", cudaGetErrorString(e)); return 1; }
        e = cudaDeviceSynchronize();
        if (e != cudaSuccess) { printf("CUDA error: %s\n", cudaGetErrorString(e)); return 1; }
      }
      const int nblk = c.cg2 ? (g == 0 ? 2 : (prop.multiProcessorCount / 2) * 2) : grid;
      cudaMemcpy(h_out, d_out, sizeof(unsigned long long) * nblk, cudaMemcpyDeviceToHost);
      double sum = 0, mx = 0;
      for (int i = 0; i < nblk; ++i) { sum += h_out[i]; if (h_out[i] > mx) mx = h_out[i]; }
      res[g] = sum / nblk / (c.iters * 4.0);
      if (g == 1) resmax = mx / (c.iters * 4.0);
    }
    printf("%s%-4d %-6s %-6s %-5d %-4d %-16s | %22.1f | %12.1f (%8.1f)\n", c.cg2 ? "2x" : "  ", c.n, c.a_tmem ? "tmem" : "smem", c.b_mn ? "MN" : "K",
           c.nacc, c.mix, bg_name(c.bg), res[0], res[1], resmax);
  }
  cudaFree(d_out);
  return 0;
}
