// Host-side helpers shared by the C-ABI translation units: thread-local error
// text, CUDA status checks, launch counter and TMA tensor-map construction
// (cuTensorMapEncodeTiled resolved at run time through cudaGetDriverEntryPoint, so
// the library has no link-time dependency on libcuda and loads on a GPU-less box).
#pragma once
#include <cuda.h>
#include <cuda_runtime.h>
#include <stdarg.h>
#include <stdint.h>
#include <stdio.h>

#include <atomic>

#include "../../include/b200_sharded.h"

namespace b200 {

char* error_buffer();              // thread-local, 512 bytes
int fail(int status, const char* fmt, ...);
std::atomic<int64_t>& launch_counter();

#define B200_CUDA_CHECK(expr)                                                                       \
  do {                                                                                              \
    cudaError_t _e = (expr);                                                                        \
    if (_e != cudaSuccess)                                                                          \
      return ::b200::fail(B200_ERR_CUDA, "%s failed: %s (%s:%d)", #expr, cudaGetErrorString(_e),    \
                          __FILE__, __LINE__);                                                      \
  } while (0)

#define B200_REQUIRE(cond, ...)                                              \
  do {                                                                       \
    if (!(cond)) return ::b200::fail(B200_ERR_INVALID_ARGUMENT, __VA_ARGS__); \
  } while (0)

// 2-D .. 4-D bf16/f32 tiled tensor map. dims/strides innermost first; strides[0] is implied
// (contiguous), strides[i>0] in BYTES. swizzle: 0 none, 1 = 32B, 2 = 64B, 3 = 128B.
int make_tensor_map(CUtensorMap* out, const void* base, int elem_bytes, int rank, const uint64_t* dims,
                    const uint64_t* strides_bytes, const uint32_t* box, int swizzle);

int num_sms();   // SM count of the current device (cached per device)

// Device buffer for the clock64 phase profiles of the attention kernels.  Always NULL unless the library was built with
// -DB200_PROFILE_HOOKS (a bring-up build: it adds b200_attention_set_profile_buffer, which is process-wide state).
long long* profile_buffer();

}  // namespace b200
