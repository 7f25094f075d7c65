// True-fp32 batched GEMM (FFMA accumulate, no TF32 / bf16 splitting) for the fp32
// sharded-matmul cases, whose parity bar is rtol 1e-5 against the unsharded product:
// jax.lax.dot on sharded (4,16)x(16,4) operands (reference case1a.py:49, case1b.py:47,
// case2.py:44, case3_fully_sharded.py:46, case4_gspmd_ff.py:52) and
// jnp.einsum("ABC,ACD->ABD") (case4_gspmd_ff.py:30).  At the literal shapes the
// per-device product is a few hundred FLOPs, i.e. launch-latency bound; the tiled
// kernel below also serves the larger fp32 measurement shape.
#include "host_common.h"

namespace b200 {

constexpr int kF32Tile = 64;    // C tile per block: 64 x 64
constexpr int kF32K = 16;       // K slice per smem stage
constexpr int kF32Threads = 256;  // 16 x 16 threads, 4 x 4 outputs each

__global__ void __launch_bounds__(kF32Threads)
gemm_f32_kernel(const float* __restrict__ A, int64_t lda, int64_t strideA, const float* __restrict__ B,
                int64_t ldb, int64_t strideB, float* __restrict__ C, int64_t ldc, int64_t strideC, int M, int N,
                int K, int accumulate) {
  __shared__ float sA[kF32K][kF32Tile + 4];   // [k][m]  (+4: keeps rows 16-byte aligned, breaks bank stride)
  __shared__ float sB[kF32K][kF32Tile + 4];   // [k][n]
  const int batch = blockIdx.z;
  A += batch * strideA;
  B += batch * strideB;
  C += batch * strideC;
  const int m0 = blockIdx.y * kF32Tile;
  const int n0 = blockIdx.x * kF32Tile;
  const int tx = threadIdx.x & 15;   // column group
  const int ty = threadIdx.x >> 4;   // row group
  float acc[4][4];
#pragma unroll
  for (int i = 0; i < 4; ++i)
#pragma unroll
    for (int j = 0; j < 4; ++j) acc[i][j] = 0.f;

  for (int k0 = 0; k0 < K; k0 += kF32K) {
    // A tile: 64 rows x 16 k, coalesced along k; stored transposed
    for (int idx = threadIdx.x; idx < kF32Tile * kF32K; idx += kF32Threads) {
      const int r = idx / kF32K, kk = idx % kF32K;
      const int gm = m0 + r, gk = k0 + kk;
      sA[kk][r] = (gm < M && gk < K) ? A[static_cast<int64_t>(gm) * lda + gk] : 0.f;
    }
    // B tile: 16 k x 64 cols, coalesced along n
    for (int idx = threadIdx.x; idx < kF32Tile * kF32K; idx += kF32Threads) {
      const int kk = idx / kF32Tile, c = idx % kF32Tile;
      const int gk = k0 + kk, gn = n0 + c;
      sB[kk][c] = (gk < K && gn < N) ? B[static_cast<int64_t>(gk) * ldb + gn] : 0.f;
    }
    __syncthreads();
#pragma unroll
    for (int kk = 0; kk < kF32K; ++kk) {
      const float4 a = *reinterpret_cast<const float4*>(&sA[kk][ty * 4]);
      const float4 b = *reinterpret_cast<const float4*>(&sB[kk][tx * 4]);
      const float av[4] = {a.x, a.y, a.z, a.w};
      const float bv[4] = {b.x, b.y, b.z, b.w};
#pragma unroll
      for (int i = 0; i < 4; ++i)
#pragma unroll
        for (int j = 0; j < 4; ++j) acc[i][j] = fmaf(av[i], bv[j], acc[i][j]);
    }
    __syncthreads();
  }
#pragma unroll
  for (int i = 0; i < 4; ++i) {
    const int gm = m0 + ty * 4 + i;
    if (gm >= M) continue;
#pragma unroll
    for (int j = 0; j < 4; ++j) {
      const int gn = n0 + tx * 4 + j;
      if (gn >= N) continue;
      float* dst = C + static_cast<int64_t>(gm) * ldc + gn;
      *dst = accumulate ? (*dst + acc[i][j]) : acc[i][j];
    }
  }
}

}  // namespace b200

using namespace b200;

extern "C" int b200_gemm_f32(const float* A, int64_t lda, int64_t strideA, const float* B, int64_t ldb,
                             int64_t strideB, float* C, int64_t ldc, int64_t strideC, int batch, int M, int N, int K,
                             int accumulate, void* stream) {
  B200_REQUIRE(A && B && C, "b200_gemm_f32: null operand pointer");
  B200_REQUIRE(batch > 0 && M > 0 && N > 0 && K > 0, "b200_gemm_f32: batch, M, N, K must be positive");
  B200_REQUIRE(lda >= K && ldb >= N && ldc >= N, "b200_gemm_f32: leading dimension smaller than row");
  B200_REQUIRE(batch <= 65535, "b200_gemm_f32: batch > 65535");
  dim3 grid((N + kF32Tile - 1) / kF32Tile, (M + kF32Tile - 1) / kF32Tile, batch);
  B200_REQUIRE(grid.y <= 65535, "b200_gemm_f32: M too large for one launch");
  gemm_f32_kernel<<<grid, kF32Threads, 0, reinterpret_cast<cudaStream_t>(stream)>>>(
      A, lda, strideA, B, ldb, strideB, C, ldc, strideC, M, N, K, accumulate);
  B200_CUDA_CHECK(cudaGetLastError());
  launch_counter().fetch_add(1);
  return B200_OK;
}
