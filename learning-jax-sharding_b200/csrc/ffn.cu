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

// ---------------------------------------------------------------------------------------------------------------
// Tensor-parallel FFN with the reduction over the 'model' axis done here (comm handle, reduce
// mode, chunks): GEMM1 on all tokens, then per token chunk GEMM2 -> f32 partial -> NCCL all-reduce / reduce-scatter on a
// side stream -> bf16 (+ b2), so the collective of chunk c runs under GEMM2 of chunk c+1.  The side stream and its events
// are created once per device (first call = set-up; make it before any stream capture) and joined to `stream` with events,
// so the call is stream-ordered from outside.  Calls on one device are serialised on the host (a per-device mutex around the
// enqueue), because they share that side stream and its events.  The throughput path of this repo is the fused variant (csrc/peer.cu +
// b200_gemm_bf16_ex); this entry point is the NCCL form for binders without CUDA-IPC.
#include <mutex>

namespace {
constexpr int kMaxChunks = 16;
struct TpSide {
  std::mutex call_mu;   // held for the whole enqueue of one call: two host threads on one device cannot interleave their
                        // cudaEventRecord / cudaStreamWaitEvent pairs on the shared events
  cudaStream_t stream = nullptr;
  cudaEvent_t ready[kMaxChunks] = {nullptr};
  cudaEvent_t done = nullptr;
};
int tp_side(TpSide** out) {
  static std::mutex mu;
  static TpSide sides[64];
  int dev = 0;
  B200_CUDA_CHECK(cudaGetDevice(&dev));
  B200_REQUIRE(dev >= 0 && dev < 64, "b200_ffn_fwd_tp: device ordinal %d out of range", dev);
  std::lock_guard<std::mutex> lock(mu);
  TpSide& s = sides[dev];
  if (!s.stream) {
    B200_CUDA_CHECK(cudaStreamCreateWithFlags(&s.stream, cudaStreamNonBlocking));
    for (int i = 0; i < kMaxChunks; ++i) B200_CUDA_CHECK(cudaEventCreateWithFlags(&s.ready[i], cudaEventDisableTiming));
    B200_CUDA_CHECK(cudaEventCreateWithFlags(&s.done, cudaEventDisableTiming));
  }
  *out = &s;
  return B200_OK;
}
}  // namespace

extern "C" int b200_ffn_fwd_tp(b200_comm* comm, const void* x, const void* W1, const void* b1, const void* W2,
                               const void* b2, void* h, float* partial, void* y, int reduce_mode, int T, int d_model, int F,
                               int act, int chunks, void* stream) {
  B200_REQUIRE(comm && x && W1 && W2 && h && partial && y, "b200_ffn_fwd_tp: null argument");
  B200_REQUIRE(reduce_mode == B200_REDUCE_ALL || reduce_mode == B200_REDUCE_SCATTER, "b200_ffn_fwd_tp: bad reduce_mode %d",
               reduce_mode);
  B200_REQUIRE(chunks >= 1 && chunks <= kMaxChunks, "b200_ffn_fwd_tp: chunks must be in [1, %d]", kMaxChunks);
  B200_REQUIRE(T > 0 && d_model > 0 && F > 0 && d_model % 4 == 0, "b200_ffn_fwd_tp: bad sizes");
  const int size = b200_comm_size(comm), rank = b200_comm_rank(comm);
  const int blocks = reduce_mode == B200_REDUCE_SCATTER ? size : 1;      // sequence blocks the output is split into
  B200_REQUIRE(T % (blocks * chunks) == 0, "b200_ffn_fwd_tp: T = %d must divide into %d block(s) x %d chunk(s)", T, blocks,
               chunks);
  const int R = T / (blocks * chunks);                                   // rows of one chunk inside one block
  B200_REQUIRE(reduce_mode == B200_REDUCE_ALL || chunks == 1 || R % 128 == 0,
               "b200_ffn_fwd_tp: reduce-scatter in chunks needs chunk rows (%d) to be a multiple of 128", R);
  const bool act_bias = act == B200_EPI_BIAS_RELU || act == B200_EPI_BIAS_GELU_TANH;
  B200_REQUIRE(!act_bias || b1, "b200_ffn_fwd_tp: activation %d needs b1", act);
  TpSide* side = nullptr;
  if (int rc = tp_side(&side)) return rc;
  std::lock_guard<std::mutex> call_lock(side->call_mu);
  cudaStream_t main_st = reinterpret_cast<cudaStream_t>(stream);
  int rc = b200_gemm_bf16(x, d_model, W1, F, B200_B_KN, act_bias ? b1 : nullptr, h, F, B200_BF16, T, F, d_model, act, stream);
  if (rc != B200_OK) return rc;
  const int64_t chunk_rows = static_cast<int64_t>(blocks) * R;           // rows GEMM2 produces per chunk
  for (int c = 0; c < chunks; ++c) {
    float* pc = partial + static_cast<int64_t>(c) * chunk_rows * d_model;
    if (blocks == 1) {
      const char* hc = static_cast<const char*>(h) + static_cast<int64_t>(c) * R * F * 2;
      rc = b200_gemm_bf16(hc, F, W2, d_model, B200_B_KN, nullptr, pc, d_model, B200_F32, R, d_model, F, B200_EPI_NONE, stream);
    } else {
      // rows [c R, (c+1) R) of EVERY rank's sequence block as one GEMM: block p of the result belongs to rank p
      const char* hc = static_cast<const char*>(h) + static_cast<int64_t>(c) * R * F * 2;
      rc = b200_gemm_bf16_rowblocked(hc, F, chunks == 1 ? 0 : R, chunks == 1 ? 0 : T / blocks, W2, d_model, B200_B_KN, nullptr,
                                     pc, d_model, B200_F32, static_cast<int>(chunk_rows), d_model, F, B200_EPI_NONE, stream);
    }
    if (rc != B200_OK) return rc;
    B200_CUDA_CHECK(cudaEventRecord(side->ready[c], main_st));
    B200_CUDA_CHECK(cudaStreamWaitEvent(side->stream, side->ready[c], 0));
    const int64_t n_out = static_cast<int64_t>(R) * d_model;
    char* yc = static_cast<char*>(y) + static_cast<int64_t>(c) * n_out * 2;
    if (blocks == 1) {
      rc = b200_all_reduce(comm, pc, pc, n_out, B200_F32, side->stream);
      if (rc == B200_OK) rc = b200_sum_partials_to_bf16(pc, 0, 1, b2, d_model, yc, n_out, side->stream);
    } else {
      float* mine = pc + static_cast<int64_t>(rank) * n_out;              // NCCL's in-place reduce-scatter position
      rc = b200_reduce_scatter(comm, pc, mine, n_out, B200_F32, side->stream);
      if (rc == B200_OK) rc = b200_sum_partials_to_bf16(mine, 0, 1, b2, d_model, yc, n_out, side->stream);
    }
    if (rc != B200_OK) return rc;
  }
  B200_CUDA_CHECK(cudaEventRecord(side->done, side->stream));
  B200_CUDA_CHECK(cudaStreamWaitEvent(main_st, side->done, 0));
  return B200_OK;
}
