// How fast can one GPU WRITE into a peer's HBM over NVLink, and with which store pattern?  (bring-up microbenchmark)
// One process, GPU 0 writes into a buffer of GPU 1 (peer access) while GPU 1 writes into GPU 0's, like the fused
// reduce-scatter push where every GPU sends and receives at once.  148 CTAs x 128 threads = the GEMM's epilogue warps.
//   rows128 / rows256 / rows512 : 16-byte stores per lane; a warp instruction covers 4 / 2 / 1 row segments of 128 / 256 /
//                                 512 contiguous bytes, rows 8 KB apart (the C matrix of the FFN's second GEMM)
//   bulkN                        : cp.async.bulk.global.shared::cta of N bytes per lane from a shared-memory tile
// nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o nvlink_store tools/diag/nvlink_store.cu
#include <cuda_runtime.h>
#include <stdint.h>
#include <stdio.h>

#define CK(x) do { cudaError_t e = (x); if (e != cudaSuccess) { printf("%s: %s\n", #x, cudaGetErrorString(e)); return 1; } } while (0)

constexpr int kRowBytes = 8192;      // ldc of a [*, 4096] bf16 matrix

template <int SEG>   // contiguous bytes per row segment: 128, 256, 512
__global__ void __launch_bounds__(128) rows_kernel(uint8_t* dst, long long rows) {
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  constexpr int lanes_per_seg = SEG / 16, segs = 32 / lanes_per_seg;
  const long long warps = (long long)gridDim.x * 4, w = (long long)blockIdx.x * 4 + warp;
  const uint4 v = make_uint4(lane, warp, blockIdx.x, 7);
  // a warp owns 32 consecutive rows at a time and walks the 8 KB row in SEG-byte steps, `segs` rows per instruction
  for (long long r0 = w * 32; r0 < rows; r0 += warps * 32)
    for (int col = 0; col < kRowBytes; col += SEG)
#pragma unroll 4
      for (int rr = 0; rr < 32; rr += segs) {
        const long long row = r0 + rr + lane / lanes_per_seg;
        *reinterpret_cast<uint4*>(dst + row * kRowBytes + col + (lane % lanes_per_seg) * 16) = v;
      }
}

template <int N>   // bytes per bulk copy (one per lane and iteration)
__global__ void __launch_bounds__(128) bulk_kernel(uint8_t* dst, long long rows) {
  extern __shared__ __align__(128) uint8_t tile[];      // 128 threads x N bytes
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  for (int i = threadIdx.x * 16; i < 128 * N; i += 128 * 16) *reinterpret_cast<uint4*>(tile + i) = make_uint4(i, 1, 2, 3);
  __syncthreads();
  asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
  const uint32_t src = (uint32_t)__cvta_generic_to_shared(tile + threadIdx.x * N);
  const long long warps = (long long)gridDim.x * 4, w = (long long)blockIdx.x * 4 + warp;
  for (long long r0 = w * 32; r0 < rows; r0 += warps * 32) {
    for (int col = 0; col < kRowBytes; col += N) {
      uint8_t* g = dst + (r0 + lane) * kRowBytes + col;
      asm volatile("cp.async.bulk.global.shared::cta.bulk_group [%0], [%1], %2;" ::"l"(g), "r"(src), "r"(N) : "memory");
      asm volatile("cp.async.bulk.commit_group;" ::: "memory");
      asm volatile("cp.async.bulk.wait_group.read 4;" ::: "memory");
    }
  }
  asm volatile("cp.async.bulk.wait_group 0;" ::: "memory");
}

template <typename F>
static float run_both(F launch, uint8_t* on0, uint8_t* on1, cudaStream_t s0, cudaStream_t s1) {
  cudaEvent_t a, b;
  cudaSetDevice(0); cudaEventCreate(&a); cudaEventCreate(&b);
  for (int it = 0; it < 2; ++it) { cudaSetDevice(0); launch(on1, s0); cudaSetDevice(1); launch(on0, s1); }
  cudaSetDevice(0); cudaStreamSynchronize(s0); cudaSetDevice(1); cudaStreamSynchronize(s1);
  cudaSetDevice(0); cudaEventRecord(a, s0);
  for (int it = 0; it < 5; ++it) { cudaSetDevice(0); launch(on1, s0); cudaSetDevice(1); launch(on0, s1); }
  cudaSetDevice(0); cudaEventRecord(b, s0); cudaStreamSynchronize(s0);
  cudaSetDevice(1); cudaStreamSynchronize(s1);
  float ms = 0; cudaEventElapsedTime(&ms, a, b);
  return ms / 5;
}

int main() {
  int n = 0; CK(cudaGetDeviceCount(&n));
  if (n < 2) { printf("needs 2 GPUs\n"); return 0; }
  const long long rows = 16384, bytes = rows * kRowBytes;   // 134 MB = one partial y
  uint8_t *b0, *b1, *l0; cudaStream_t s0, s1;
  CK(cudaSetDevice(0)); CK(cudaDeviceEnablePeerAccess(1, 0)); CK(cudaMalloc(&b0, bytes)); CK(cudaMalloc(&l0, bytes)); CK(cudaStreamCreate(&s0));
  CK(cudaSetDevice(1)); CK(cudaDeviceEnablePeerAccess(0, 0)); CK(cudaMalloc(&b1, bytes)); CK(cudaStreamCreate(&s1));
  auto report = [&](const char* name, float ms) { printf("%-10s %.4f ms  %.0f GB/s per direction\n", name, ms, bytes / ms / 1e6); };
#define ROWS(SEG) report("rows" #SEG, run_both([&](uint8_t* d, cudaStream_t s) { rows_kernel<SEG><<<148, 128, 0, s>>>(d, rows); }, b0, b1, s0, s1));
  ROWS(128) ROWS(256) ROWS(512)
#define BULK(N) { cudaSetDevice(0); cudaFuncSetAttribute(bulk_kernel<N>, cudaFuncAttributeMaxDynamicSharedMemorySize, 128 * N); \
                  cudaSetDevice(1); cudaFuncSetAttribute(bulk_kernel<N>, cudaFuncAttributeMaxDynamicSharedMemorySize, 128 * N); \
    report("bulk" #N, run_both([&](uint8_t* d, cudaStream_t s) { bulk_kernel<N><<<148, 128, 128 * N, s>>>(d, rows); }, b0, b1, s0, s1)); }
  BULK(128) BULK(256) BULK(512) BULK(1024)
  // the same into LOCAL memory, for scale
  cudaSetDevice(0);
  cudaEvent_t a, b; cudaEventCreate(&a); cudaEventCreate(&b);
  rows_kernel<128><<<148, 128, 0, s0>>>(l0, rows);
  cudaEventRecord(a, s0);
  for (int it = 0; it < 5; ++it) rows_kernel<128><<<148, 128, 0, s0>>>(l0, rows);
  cudaEventRecord(b, s0); cudaStreamSynchronize(s0);
  float ms; cudaEventElapsedTime(&ms, a, b); report("local128", ms / 5);
  CK(cudaGetLastError());
  return 0;
}
