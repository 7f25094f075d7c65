// Peer-memory plumbing for the fused "GEMM -> reduce-scatter" of row-parallel layers (one process per
// GPU, NVLink 5 / NVSwitch): receive buffers shared through CUDA IPC, a cross-GPU barrier on flags in
// those buffers, and the local reduction of the received bf16 partials.
//
// Step of a row-parallel layer on rank r (W ranks, T rows, R = T / W rows per owner):
//   1. b200_gemm_bf16_ex with c_block_ptrs[p] = rank p's receive slot [r]: every finished tile of the
//      partial product is stored straight into its owner's memory from the GEMM epilogue;
//   2. b200_peer_barrier: all ranks' pushes are complete and visible;
//   3. b200_sum_bf16_partials: y[R, N] = bf16( sum_s recv[s][R, N] ) (+ bias), f32 accumulation.
// This replaces the all-reduce / reduce-scatter XLA inserts after a contraction-sharded dot
// (reference pattern: case1a.py:24-49; FFN second Dense, case6_attention.py:36-40).
#include <string.h>

#include "host_common.h"
#include "sm100_ptx.cuh"

namespace b200 {

__global__ void peer_barrier_kernel(uint32_t* const* flag_ptrs, int rank, int world, uint32_t epoch) {
  // thread p signals rank p, then waits for rank p's signal in this rank's own flag block
  const int peer = threadIdx.x;
  if (peer >= world) return;
  __threadfence_system();   // this GPU's earlier peer stores (previous kernels in the stream) are ordered before the flag
  uint32_t* remote = flag_ptrs[peer] + rank;
  asm volatile("st.release.sys.global.u32 [%0], %1;" ::"l"(remote), "r"(epoch) : "memory");
  const uint32_t* mine = flag_ptrs[rank] + peer;
  const long long t0 = clock64();
  uint32_t seen;
  do {
    asm volatile("ld.acquire.sys.global.u32 %0, [%1];" : "=r"(seen) : "l"(mine) : "memory");
    if (clock64() - t0 > 4000000000ll) {   // ~2 s: a rank is missing -- fail instead of hanging the GPU
      printf("b200: peer barrier timeout rank=%d waiting for rank %d epoch=%u seen=%u\n", rank, peer, epoch, seen);
      __trap();
    }
  } while (static_cast<int32_t>(seen - epoch) < 0);
}

// dst[i] = bf16(sum_r src[r*stride + i]) (+ bias[i % cols]); 8 elements per thread, f32 accumulation
__global__ void __launch_bounds__(256) sum_bf16_partials_kernel(const __nv_bfloat16* __restrict__ src,
                                                               int64_t src_stride, int nsrc,
                                                               const __nv_bfloat16* __restrict__ bias, int64_t cols,
                                                               __nv_bfloat16* __restrict__ dst, int64_t n) {
  const int64_t nvec = n / 8;
  const int64_t stride = static_cast<int64_t>(gridDim.x) * blockDim.x;
  for (int64_t i = static_cast<int64_t>(blockIdx.x) * blockDim.x + threadIdx.x; i < nvec; i += stride) {
    float acc[8];
#pragma unroll
    for (int t = 0; t < 8; ++t) acc[t] = 0.f;
    for (int r = 0; r < nsrc; ++r) {
      const uint4 v = reinterpret_cast<const uint4*>(src + r * src_stride)[i];
      const __nv_bfloat162* h = reinterpret_cast<const __nv_bfloat162*>(&v);
#pragma unroll
      for (int t = 0; t < 4; ++t) {
        const float2 f = __bfloat1622float2(h[t]);
        acc[2 * t] += f.x;
        acc[2 * t + 1] += f.y;
      }
    }
    if (bias != nullptr) {   // Dense semantics: round the sum to bf16, then add the bf16 bias in bf16
      const uint4 bv = *reinterpret_cast<const uint4*>(bias + (i * 8) % cols);
      const __nv_bfloat162* hb = reinterpret_cast<const __nv_bfloat162*>(&bv);
#pragma unroll
      for (int t = 0; t < 4; ++t) {
        const float2 f = __bfloat1622float2(hb[t]);
        acc[2 * t] = bf16_round(acc[2 * t]) + f.x;
        acc[2 * t + 1] = bf16_round(acc[2 * t + 1]) + f.y;
      }
    }
    uint4 o;
    o.x = pack_bf16x2(acc[0], acc[1]);
    o.y = pack_bf16x2(acc[2], acc[3]);
    o.z = pack_bf16x2(acc[4], acc[5]);
    o.w = pack_bf16x2(acc[6], acc[7]);
    reinterpret_cast<uint4*>(dst)[i] = o;
  }
}

// The reduction of sum_bf16_partials_kernel, one raster group of rows at a time, each as soon as every sender's progress flag for
// that group shows this step's epoch (see GemmParams::done_*): runs beside the pushing GEMMs, so that only the last group's
// share of the reduction is left when they finish.
__global__ void __launch_bounds__(256) progressive_sum_kernel(const uint32_t* __restrict__ flags, uint32_t epoch, int senders,
                                                             int groups, int64_t group_elems,
                                                             const __nv_bfloat16* __restrict__ src, int64_t src_stride,
                                                             const __nv_bfloat16* __restrict__ bias, int64_t cols,
                                                             __nv_bfloat16* __restrict__ dst) {
  const int64_t nvec = group_elems / 8;
  const int64_t stride = static_cast<int64_t>(gridDim.x) * blockDim.x;
  for (int g = 0; g < groups; ++g) {
    if (threadIdx.x < senders) {
      const uint32_t* f = flags + threadIdx.x * groups + g;
      const long long t0 = clock64();
      uint32_t seen;
      for (;;) {
        asm volatile("ld.acquire.sys.global.u32 %0, [%1];" : "=r"(seen) : "l"(f) : "memory");
        if (static_cast<int32_t>(seen - epoch) >= 0) break;
        __nanosleep(200);
        if (clock64() - t0 > 8000000000ll) {   // ~4 s: a sender is missing -- fail instead of hanging the GPU
          printf("b200: progressive sum timeout waiting for sender %d group %d epoch=%u seen=%u\n", (int)threadIdx.x, g, epoch, seen);
          __trap();
        }
      }
    }
    __syncthreads();
    const int64_t base = static_cast<int64_t>(g) * nvec;
    for (int64_t i = static_cast<int64_t>(blockIdx.x) * blockDim.x + threadIdx.x; i < nvec; i += stride) {
      float acc[8];
#pragma unroll
      for (int t = 0; t < 8; ++t) acc[t] = 0.f;
      for (int r = 0; r < senders; ++r) {
        const uint4 v = reinterpret_cast<const uint4*>(src + r * src_stride)[base + i];
        const __nv_bfloat162* h = reinterpret_cast<const __nv_bfloat162*>(&v);
#pragma unroll
        for (int t = 0; t < 4; ++t) {
          const float2 f = __bfloat1622float2(h[t]);
          acc[2 * t] += f.x;
          acc[2 * t + 1] += f.y;
        }
      }
      if (bias != nullptr) {
        const uint4 bv = *reinterpret_cast<const uint4*>(bias + ((base + i) * 8) % cols);
        const __nv_bfloat162* hb = reinterpret_cast<const __nv_bfloat162*>(&bv);
#pragma unroll
        for (int t = 0; t < 4; ++t) {
          const float2 f = __bfloat1622float2(hb[t]);
          acc[2 * t] = bf16_round(acc[2 * t]) + f.x;
          acc[2 * t + 1] = bf16_round(acc[2 * t + 1]) + f.y;
        }
      }
      uint4 o;
      o.x = pack_bf16x2(acc[0], acc[1]);
      o.y = pack_bf16x2(acc[2], acc[3]);
      o.z = pack_bf16x2(acc[4], acc[5]);
      o.w = pack_bf16x2(acc[6], acc[7]);
      reinterpret_cast<uint4*>(dst)[base + i] = o;
    }
  }
}

// Pull one column chunk of every rank's weight shard into the local gathered matrix: W shards of
// [rows, src_pitch] (this rank's own included), destination rows p*rows .. (p+1)*rows of a [W*rows, dst_pitch]
// matrix.  16-byte loads straight from the peers' HBM over NVLink, eight in flight per thread; each CTA starts
// at a different peer so that all links are busy at once.
__global__ void __launch_bounds__(256) peer_gather_kernel(const uint8_t* const* __restrict__ src_bases, int world,
                                                         int rank, int64_t src_off_bytes, int64_t src_pitch,
                                                         uint8_t* __restrict__ dst, int64_t dst_pitch, int rows,
                                                         int width16) {
  const int64_t units = static_cast<int64_t>(rows) * width16;            // 16-byte units of one shard's chunk
  const int64_t per_cta = (units + gridDim.x - 1) / gridDim.x;
  const int64_t u_begin = per_cta * blockIdx.x;
  const int64_t u_end = u_begin + per_cta < units ? u_begin + per_cta : units;
  for (int pp = 0; pp < world; ++pp) {
    const int p = (rank + 1 + pp + blockIdx.x) % world;
    const uint8_t* src = src_bases[p] + src_off_bytes;
    uint8_t* out = dst + static_cast<int64_t>(p) * rows * dst_pitch;
    for (int64_t u0 = u_begin + threadIdx.x; u0 < u_end; u0 += 8 * 256) {
      uint4 v[8];
#pragma unroll
      for (int k = 0; k < 8; ++k) {
        const int64_t u = u0 + k * 256;
        if (u < u_end) {
          const int64_t r = u / width16, c = u % width16;
          v[k] = *reinterpret_cast<const uint4*>(src + r * src_pitch + c * 16);
        }
      }
#pragma unroll
      for (int k = 0; k < 8; ++k) {
        const int64_t u = u0 + k * 256;
        if (u < u_end) {
          const int64_t r = u / width16, c = u % width16;
          *reinterpret_cast<uint4*>(out + r * dst_pitch + c * 16) = v[k];
        }
      }
    }
  }
}

// The PUSH form of the weight all-gather: this rank's contiguous shard goes to the same offset of every rank's gathered
// buffer (its own included) with plain 16-byte stores.  Stores to peer memory are posted -- they reach the ~677 GB/s per
// direction that NVLink gives peer writes here with a few CTAs (tools/diag/nvlink_store.cu) -- whereas the pull form above
// is bound by the round-trip latency of its loads.  Each CTA starts at a different peer so that all links carry traffic
// from the first instruction on.
__global__ void __launch_bounds__(256) peer_push_kernel(uint8_t* const* __restrict__ dst_bases, int world, int rank,
                                                       int64_t dst_off_bytes, const uint8_t* __restrict__ src,
                                                       int64_t units) {
  const int64_t per_cta = (units + gridDim.x - 1) / gridDim.x;
  const int64_t u_begin = per_cta * blockIdx.x;
  const int64_t u_end = u_begin + per_cta < units ? u_begin + per_cta : units;
  const uint4* s4 = reinterpret_cast<const uint4*>(src);
  for (int64_t u0 = u_begin + threadIdx.x; u0 < u_end; u0 += 4 * 256) {
    uint4 v[4];
#pragma unroll
    for (int k = 0; k < 4; ++k)
      if (u0 + k * 256 < u_end) v[k] = s4[u0 + k * 256];
    for (int pp = 0; pp < world; ++pp) {
      const int p = (rank + 1 + pp + blockIdx.x) % world;
      uint4* d4 = reinterpret_cast<uint4*>(dst_bases[p] + dst_off_bytes);
#pragma unroll
      for (int k = 0; k < 4; ++k)
        if (u0 + k * 256 < u_end) d4[u0 + k * 256] = v[k];
    }
  }
}

}  // namespace b200

using namespace b200;

extern "C" int b200_peer_push(void* const* dst_bases_dev, int world, int rank, int64_t dst_off_bytes, const void* src,
                              int64_t bytes, int ctas, void* stream) {
  B200_REQUIRE(dst_bases_dev && src && world >= 1 && rank >= 0 && rank < world && bytes > 0, "b200_peer_push: bad arguments");
  B200_REQUIRE(bytes % 16 == 0 && dst_off_bytes % 16 == 0 && reinterpret_cast<uintptr_t>(src) % 16 == 0,
               "b200_peer_push: offset, size and source must be multiples of 16 bytes");
  B200_REQUIRE(ctas >= 1 && ctas <= 1024, "b200_peer_push: ctas must be in [1, 1024]");
  peer_push_kernel<<<ctas, 256, 0, reinterpret_cast<cudaStream_t>(stream)>>>(
      reinterpret_cast<uint8_t* const*>(dst_bases_dev), world, rank, dst_off_bytes, reinterpret_cast<const uint8_t*>(src),
      bytes / 16);
  B200_CUDA_CHECK(cudaGetLastError());
  launch_counter().fetch_add(1);
  return B200_OK;
}

extern "C" int b200_peer_gather(void* const* src_bases_dev, int world, int rank, int64_t src_off_bytes,
                                int64_t src_pitch_bytes, void* dst, int64_t dst_pitch_bytes, int rows,
                                int64_t width_bytes, int ctas, void* stream) {
  B200_REQUIRE(src_bases_dev && dst && world >= 1 && rank >= 0 && rank < world && rows > 0 && width_bytes > 0,
               "b200_peer_gather: bad arguments");
  B200_REQUIRE(width_bytes % 16 == 0 && src_pitch_bytes % 16 == 0 && dst_pitch_bytes % 16 == 0 && src_off_bytes % 16 == 0,
               "b200_peer_gather: offsets, pitches and width must be multiples of 16 bytes");
  B200_REQUIRE(ctas >= 1 && ctas <= 1024, "b200_peer_gather: ctas must be in [1, 1024]");
  peer_gather_kernel<<<ctas, 256, 0, reinterpret_cast<cudaStream_t>(stream)>>>(
      reinterpret_cast<const uint8_t* const*>(src_bases_dev), world, rank, src_off_bytes, src_pitch_bytes,
      reinterpret_cast<uint8_t*>(dst), dst_pitch_bytes, rows, static_cast<int>(width_bytes / 16));
  B200_CUDA_CHECK(cudaGetLastError());
  launch_counter().fetch_add(1);
  return B200_OK;
}

extern "C" int b200_peer_alloc(uint64_t bytes, void** dev_ptr, void* handle64) {
  B200_REQUIRE(bytes > 0 && dev_ptr && handle64, "b200_peer_alloc: bad arguments");
  void* p = nullptr;
  B200_CUDA_CHECK(cudaMalloc(&p, bytes));
  cudaError_t e = cudaMemset(p, 0, bytes);
  if (e == cudaSuccess) e = cudaDeviceSynchronize();
  cudaIpcMemHandle_t h;
  if (e == cudaSuccess) e = cudaIpcGetMemHandle(&h, p);
  if (e != cudaSuccess) {
    cudaFree(p);
    return fail(B200_ERR_CUDA, "b200_peer_alloc: %s", cudaGetErrorString(e));
  }
  static_assert(sizeof(cudaIpcMemHandle_t) == 64, "IPC handle is 64 bytes");
  memcpy(handle64, &h, 64);
  *dev_ptr = p;
  return B200_OK;
}

extern "C" int b200_peer_open(const void* handle64, void** dev_ptr) {
  B200_REQUIRE(handle64 && dev_ptr, "b200_peer_open: bad arguments");
  cudaIpcMemHandle_t h;
  memcpy(&h, handle64, 64);
  void* p = nullptr;
  B200_CUDA_CHECK(cudaIpcOpenMemHandle(&p, h, cudaIpcMemLazyEnablePeerAccess));
  *dev_ptr = p;
  return B200_OK;
}

extern "C" int b200_peer_close(void* dev_ptr) {
  B200_CUDA_CHECK(cudaIpcCloseMemHandle(dev_ptr));
  return B200_OK;
}

extern "C" int b200_peer_free(void* dev_ptr) {
  B200_CUDA_CHECK(cudaFree(dev_ptr));
  return B200_OK;
}

extern "C" int b200_peer_copy2d(void* dst, int64_t dst_pitch_bytes, const void* src, int64_t src_pitch_bytes,
                                int64_t width_bytes, int64_t rows, void* stream) {
  B200_REQUIRE(dst && src && width_bytes > 0 && rows > 0, "b200_peer_copy2d: bad arguments");
  B200_REQUIRE(dst_pitch_bytes >= width_bytes && src_pitch_bytes >= width_bytes, "b200_peer_copy2d: pitch smaller than row");
  // device-to-device (local or IPC-mapped peer) strided copy on the copy engines: no SM time
  B200_CUDA_CHECK(cudaMemcpy2DAsync(dst, static_cast<size_t>(dst_pitch_bytes), src, static_cast<size_t>(src_pitch_bytes),
                                    static_cast<size_t>(width_bytes), static_cast<size_t>(rows), cudaMemcpyDefault,
                                    reinterpret_cast<cudaStream_t>(stream)));
  return B200_OK;
}

extern "C" int b200_peer_barrier(void* const* flag_ptrs_dev, int rank, int world, uint32_t epoch, void* stream) {
  B200_REQUIRE(flag_ptrs_dev && world >= 1 && world <= 64 && rank >= 0 && rank < world,
               "b200_peer_barrier: bad arguments (world must be 1..64)");
  peer_barrier_kernel<<<1, 64, 0, reinterpret_cast<cudaStream_t>(stream)>>>(
      reinterpret_cast<uint32_t* const*>(flag_ptrs_dev), rank, world, epoch);
  B200_CUDA_CHECK(cudaGetLastError());
  launch_counter().fetch_add(1);
  return B200_OK;
}

extern "C" int b200_sum_bf16_partials(const void* src, int64_t src_stride, int nsrc, const void* bias, int64_t cols,
                                      void* dst, int64_t n, void* stream) {
  B200_REQUIRE(src && dst && nsrc >= 1 && n >= 0, "b200_sum_bf16_partials: bad arguments");
  B200_REQUIRE(n % 8 == 0 && src_stride % 8 == 0, "b200_sum_bf16_partials: n and src_stride must be multiples of 8");
  B200_REQUIRE(!bias || (cols > 0 && cols % 8 == 0 && n % cols == 0),
               "b200_sum_bf16_partials: cols must be a positive multiple of 8 dividing n when bias is given");
  B200_REQUIRE((reinterpret_cast<uintptr_t>(src) | reinterpret_cast<uintptr_t>(dst) | reinterpret_cast<uintptr_t>(bias)) % 16 == 0,
               "b200_sum_bf16_partials: buffers must be 16-byte aligned");
  if (n == 0) return B200_OK;
  const int sms = num_sms();
  int64_t blocks = (n / 8 + 255) / 256;
  if (blocks > static_cast<int64_t>(sms) * 8) blocks = static_cast<int64_t>(sms) * 8;
  sum_bf16_partials_kernel<<<static_cast<int>(blocks), 256, 0, reinterpret_cast<cudaStream_t>(stream)>>>(
      reinterpret_cast<const __nv_bfloat16*>(src), src_stride, nsrc, reinterpret_cast<const __nv_bfloat16*>(bias),
      cols > 0 ? cols : 8, reinterpret_cast<__nv_bfloat16*>(dst), n);
  B200_CUDA_CHECK(cudaGetLastError());
  launch_counter().fetch_add(1);
  return B200_OK;
}

extern "C" int b200_peer_progressive_sum(const uint32_t* flags, uint32_t epoch, int senders, int groups, int64_t rows_per_group,
                                         const void* src, int64_t src_stride, const void* bias, int64_t cols, void* dst,
                                         int ctas, void* stream) {
  B200_REQUIRE(flags && src && dst && senders >= 1 && senders <= 64 && groups >= 1 && rows_per_group > 0 && cols > 0,
               "b200_peer_progressive_sum: bad arguments (1..64 senders)");
  B200_REQUIRE(cols % 8 == 0 && src_stride % 8 == 0, "b200_peer_progressive_sum: cols and src_stride must be multiples of 8");
  B200_REQUIRE((reinterpret_cast<uintptr_t>(src) | reinterpret_cast<uintptr_t>(dst) | reinterpret_cast<uintptr_t>(bias)) % 16 == 0,
               "b200_peer_progressive_sum: buffers must be 16-byte aligned");
  B200_REQUIRE(ctas >= 1 && ctas <= 4096, "b200_peer_progressive_sum: ctas must be in [1, 4096]");
  // Launched as a PROGRAMMATIC DEPENDENT of the kernel before it on the stream (the pushing GEMM): it starts when every CTA
  // of that grid has executed griddepcontrol.launch_dependents -- all of the persistent GEMM is resident by then -- and runs
  // beside it.  (Started any earlier, its CTAs could take the registers or thread slots of an SM that a GEMM CTA still needs,
  // and then wait for ever for the tiles of that very CTA.)  It never executes griddepcontrol.wait: the data it reads is
  // ordered by the per-group flags, not by the GEMM's completion.
  cudaLaunchConfig_t cfg = {};
  cfg.gridDim = dim3(ctas);
  cfg.blockDim = dim3(256);
  cfg.dynamicSmemBytes = 0;
  cfg.stream = reinterpret_cast<cudaStream_t>(stream);
  cudaLaunchAttribute attrs[1];
  attrs[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
  attrs[0].val.programmaticStreamSerializationAllowed = 1;
  cfg.attrs = attrs;
  cfg.numAttrs = 1;
  const int64_t group_elems = rows_per_group * cols;
  B200_CUDA_CHECK(cudaLaunchKernelEx(&cfg, progressive_sum_kernel, flags, epoch, senders, groups, group_elems,
                                     reinterpret_cast<const __nv_bfloat16*>(src), src_stride,
                                     reinterpret_cast<const __nv_bfloat16*>(bias), cols, reinterpret_cast<__nv_bfloat16*>(dst)));
  launch_counter().fetch_add(1);
  return B200_OK;
}
