// One device's share of the feed-forward block y = act(x.W1 + b1).W2 (+ b2): two launches of
// the tcgen05 GEMM, the first with the bias+activation epilogue fused (h never leaves bf16),
// the second with the optional output bias or a raw f32 store for row-parallel partial sums.
// Reference: the FFN formula in the comment at case6_attention.py:36-40; W1 column-parallel is
// the layout of case4_gspmd_ff.py:46-52, W2 row-parallel the contraction-sharded layout of
// case1a.py:24-49.
#include "host_common.h"

using namespace b200;

extern "C" int b200_ffn_fwd(const void* x, const void* W1, const void* b1, const void* W2, const void* b2, void* h,
                            void* y, int y_dtype, int T, int d_model, int F, int act, void* stream) {
  B200_REQUIRE(x && W1 && W2 && h && y, "b200_ffn_fwd: null operand pointer");
  B200_REQUIRE(T > 0 && d_model > 0 && F > 0, "b200_ffn_fwd: T, d_model, F must be positive");
  B200_REQUIRE(act == B200_EPI_BIAS_RELU || act == B200_EPI_BIAS_GELU_TANH || act == B200_EPI_RELU ||
                   act == B200_EPI_GELU_TANH,
               "b200_ffn_fwd: act must be one of the RELU / GELU_TANH epilogues (got %d)", act);
  const bool act_bias = act == B200_EPI_BIAS_RELU || act == B200_EPI_BIAS_GELU_TANH;
  B200_REQUIRE(!act_bias || b1, "b200_ffn_fwd: activation %d needs b1", act);
  B200_REQUIRE(!(y_dtype == B200_F32 && b2), "b200_ffn_fwd: f32 partial output takes no b2 (add it after the reduction)");
  int rc = b200_gemm_bf16(x, d_model, W1, F, B200_B_KN, act_bias ? b1 : nullptr, h, F, B200_BF16, T, F, d_model, act,
                          stream);
  if (rc != B200_OK) return rc;
  return b200_gemm_bf16(h, F, W2, d_model, B200_B_KN, b2, y, d_model, y_dtype, T, d_model, F,
                        b2 ? B200_EPI_BIAS : B200_EPI_NONE, stream);
}
