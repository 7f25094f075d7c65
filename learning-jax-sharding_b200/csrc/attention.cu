#include "host_common.h"
using namespace b200;
extern "C" int b200_attention_fwd(const void* Q, const void* K, const void* V, void* O, int B, int Sq, int Sk, int H, int D,
                       int64_t q_stride_b, int64_t q_stride_s, int64_t q_stride_h, int64_t k_stride_b,
                       int64_t k_stride_s, int64_t k_stride_h, int64_t v_stride_b, int64_t v_stride_s,
                       int64_t v_stride_h, int64_t o_stride_b, int64_t o_stride_s, int64_t o_stride_h, float scale,
                       int causal, void* stream) {
  return fail(B200_ERR_UNSUPPORTED, "b200_attention_fwd: not built yet");
}
