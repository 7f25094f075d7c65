/*
 * b200_sharded.h -- C ABI of libb200sl.so, the B200 (sm_100a) per-shard compute
 * library behind the sharded matmul / Dense / FFN / attention layers that the
 * entrpn/learning-jax-sharding case scripts run under jax.jit.
 *
 * Every entry point is what an XLA FFI custom-call handler (or a ctypes / cffi
 * binding) for that layer would call once per device with that device's shard
 * buffers -- INTEGRATION.md shows both bindings.  Conventions:
 *   - all buffers are DEVICE pointers owned by the caller; results are
 *     pre-allocated by the caller; nothing is retained, freed or resized here;
 *   - calls only ENQUEUE work on `stream` (a cudaStream_t passed as void*; NULL =
 *     legacy default stream) of the CURRENT device and never synchronise it;
 *   - no exceptions cross the boundary: the return value is a b200_status and
 *     b200_last_error() returns a thread-local message for the last failure;
 *   - no torch / C++ types in any signature.
 * The reference has no native code; each comment cites the reference Python call
 * site (file:line under the upstream repo) whose per-device lowering the entry
 * point replaces.
 */
#ifndef B200_SHARDED_H_
#define B200_SHARDED_H_

#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define B200_VERSION 100 /* major*10000 + minor*100 + patch */

typedef enum {
  B200_OK = 0,
  B200_ERR_INVALID_ARGUMENT = 1, /* shape / alignment / enum out of range            */
  B200_ERR_CUDA = 2,             /* a CUDA runtime or driver call failed              */
  B200_ERR_UNSUPPORTED = 3       /* valid request this build has no kernel for        */
} b200_status;

typedef enum { B200_BF16 = 0, B200_F32 = 1 } b200_dtype;

/* Epilogue fused into the GEMM's TMEM->global path.  Rounding follows two chained
 * bf16 flax.linen.Dense layers: bf16(acc), then bf16(+bias), then bf16(act(.)). */
typedef enum {
  B200_EPI_NONE = 0,
  B200_EPI_BIAS = 1,
  B200_EPI_BIAS_RELU = 2,
  B200_EPI_BIAS_GELU_TANH = 3,
  B200_EPI_RELU = 4,
  B200_EPI_GELU_TANH = 5
} b200_epilogue;

/* Layout of the B operand of b200_gemm_bf16. */
typedef enum {
  B200_B_KN = 0, /* B is [K,N] row-major: a flax Dense kernel (in_features, out_features) */
  B200_B_NK = 1  /* B is [N,K] row-major: C = A . B^T (e.g. Q.K^T, pre-transposed weights) */
} b200_b_layout;

/* Layout of the A operand (b200_gemm_args.a_layout). */
typedef enum {
  B200_A_MK = 0, /* A is [M,K] row-major */
  B200_A_KM = 1  /* A is stored [K,M] row-major: C = A^T . B (weight gradients dW = x^T . dy) */
} b200_a_layout;

/* Library version (B200_VERSION of the build). */
int b200_version(void);
/* Message for the last non-OK status returned on this host thread ("" if none). */
const char* b200_last_error(void);
/* Number of kernels this library has launched in this process (all threads). */
int64_t b200_launch_count(void);

/*
 * C[M,N] = epilogue(A[M,K] . B + bias), bf16 inputs, f32 accumulation in TMEM
 * (tcgen05.mma), output bf16 or f32.  A row-major with row stride lda (elements),
 * B per b_layout with row stride ldb, C row-major with row stride ldc.
 * Requirements: K % 8 == 0, N % 8 == 0, lda/ldb/ldc % 8 == 0, 16-byte aligned bases.
 * bias: bf16[N] or NULL (required for the *_BIAS_* epilogues).
 * Replaces the per-device lowering of flax nn.Dense(dtype=bf16)
 *   (case6_attention.py:61-90,96-99,139; case5_attention_dense.py:59-67)
 * and of each half of the FFN  y = act(x.Win).Wout  (case6_attention.py:36-40;
 * column-parallel half case4_gspmd_ff.py:46-52, row-parallel half case1a.py:24-49).
 */
int b200_gemm_bf16(const void* A, int64_t lda, const void* B, int64_t ldb, int b_layout, const void* bias,
                   void* C, int64_t ldc, int out_dtype, int M, int N, int K, int epilogue, void* stream);

/*
 * b200_gemm_bf16 with a row-blocked view of A: logical row m of A is read from physical row
 *   (m / a_rows_per_block) * a_block_stride + m % a_rows_per_block
 * (a_rows_per_block a multiple of 128; 0 = plain).  Lets a tensor-parallel layer take the token
 * chunk "rows [c*R, (c+1)*R) of every rank's sequence block" as ONE GEMM, so that the
 * reduce-scatter of that chunk lands in contiguous sequence blocks (the layout
 * PartitionSpec('data','model') gives x and y in case6_attention.py:161,230).
 */
int b200_gemm_bf16_rowblocked(const void* A, int64_t lda, int a_rows_per_block, int a_block_stride, const void* B,
                              int64_t ldb, int b_layout, const void* bias, void* C, int64_t ldc, int out_dtype, int M,
                              int N, int K, int epilogue, void* stream);

/*
 * General form of b200_gemm_bf16.  Zero-initialise the struct, fill what you need.
 * Besides the plain operands and the row-blocked A view it offers a ROW-BLOCK SCATTER of the
 * output, the compute half of a fused "GEMM -> reduce-scatter" for row-parallel layers (the
 * all-reduce XLA inserts after a contraction-sharded dot, case1a.py:49; the FFN's second Dense,
 * case6_attention.py:36-40): with m_rows_per_block = R (a multiple of 256 dividing M), logical
 * row m belongs to row block (m / R + m_block_rot) % (M / R); A is read from that block's rows and
 * the bf16 result is written to c_block_ptrs[block] + (m % R) * ldc, where c_block_ptrs is a
 * DEVICE array of M / R base pointers -- typically peer GPUs' receive slots mapped through CUDA
 * IPC (b200_peer_*), so partial sums cross NVLink straight out of the GEMM epilogue, tile by
 * tile, in 128-byte row segments.  m_block_rot staggers the ranks (rank r starts with block r+1)
 * so that no two ranks push to the same destination at once.  c_block_ptrs = NULL keeps the
 * result in C (rotation only).  The scatter may be combined with the row-blocked A view when
 * a_rows_per_block == m_rows_per_block: block b of the product is then read from rows
 * [b * a_block_stride, b * a_block_stride + R) of A -- a token CHUNK of every owner's row block, so
 * that the reduction of chunk c can run beside the GEMM of chunk c+1.
 */
/* Launch-heuristic overrides carried by the call itself (the library keeps no tuning state between calls, so two host
 * threads or two devices can never see each other's settings).  Zero in every field = the library's own choice. */
typedef struct {
  int cta_group;    /* 1 = one CTA per tile, 2 = CTA pairs (tcgen05 cta_group::2, 256-row tiles) */
  int block_n;      /* 64, 128 or 256 columns per tile */
  int group_m;      /* raster group: row-blocks that sweep all column-blocks together (L2 reuse of the A panels) */
  int max_stages;   /* cap on the shared-memory ring depth */
  int max_ctas;     /* cap on the persistent grid, e.g. to leave SMs to a collective kernel running beside the GEMM */
  int k_serpentine; /* 1 = every other tile of a CTA walks K downwards */
} b200_gemm_tuning;

typedef struct {
  const void* A;
  int64_t lda;
  int a_rows_per_block, a_block_stride; /* row-blocked A view, 0 = off (see b200_gemm_bf16_rowblocked) */
  const void* B;
  int64_t ldb;
  int b_layout;
  const void* bias;
  void* C;
  int64_t ldc;
  int out_dtype;
  void* const* c_block_ptrs;            /* device array of row-block base pointers, or NULL */
  int m_rows_per_block, m_block_rot;    /* 0 = off */
  int M, N, K, epilogue;
  int a_layout;                         /* B200_A_MK (0): A is [M,K] row-major; B200_A_KM: A is stored [K,M] row-major,
                                           i.e. the product is (stored A)^T . B -- the weight-gradient GEMM dW = x^T . dy of
                                           jax.grad through nn.Dense (case5_attention_dense.py:130-139) */
  b200_gemm_tuning tune;                /* per-call overrides of the launch heuristics; all zero = library choice */
  int a_blocks_inner, a_outer_stride;   /* two-level row-blocked A view (0 = off): block q of a_rows_per_block rows starts
                                           at row (q % a_blocks_inner) * a_block_stride + (q / a_blocks_inner) *
                                           a_outer_stride.  With a [batch, seq, K] activation, a_rows_per_block = seq / W,
                                           a_blocks_inner = batch, a_block_stride = seq, a_outer_stride = seq / W the
                                           product comes out in "owner j: its seq block of every batch item" order, which
                                           together with the row-block scatter (m_rows_per_block = batch * seq / W) is the
                                           reduce-scatter ALONG SEQ that out_shardings = P('data','model') asks for
                                           (case6_attention.py:161,230), whatever the batch size */
  int m_block_interleave;               /* row-block scatter only.  0: the kernel finishes one owner's block after the
                                           other (rank r: r+1, r+2, ..., its own last -- one sender per destination at a
                                           time).  1: consecutive tiles cycle through the owner blocks, so every wave
                                           sends to all peers at once and the NVLink traffic is spread evenly over the
                                           kernel; same result, bit for bit */
  /* Progress signals of a row-block scatter (all zero = off): the kernel produces its tiles raster group by raster group
   * (tune.group_m row-blocks of 128 * tune.cta_group rows x all column-blocks); when every tile of group g is stored and
   * visible system-wide it writes done_epoch into flag [done_sender * done_groups + g] of EVERY owner's flag array
   * (done_flag_ptrs: device array of M / m_rows_per_block pointers into peer memory), so that b200_peer_progressive_sum
   * on the owner reduces the rows of a group while the later groups are still being produced.  With m_block_interleave
   * group g holds rows [g, g + 1) * (group_m / owners) * 128 * cta_group of every owner's block.  done_counters: local
   * device array of done_groups uint32, zero-initialised once, never reset (it counts across launches).  Needs explicit
   * tune.cta_group / block_n / group_m. */
  unsigned int* done_counters;
  unsigned int* const* done_flag_ptrs;
  int done_sender, done_groups;
  unsigned int done_epoch;
} b200_gemm_args;
int b200_gemm_bf16_ex(const b200_gemm_args* args, void* stream);

/*
 * Batched true-fp32 GEMM (FFMA, no TF32): for b in [0,batch):
 *   C_b[M,N] = A_b[M,K] . B_b[K,N], all row-major, element strides lda/ldb/ldc
 *   between rows and strideA/strideB/strideC between batch items.
 * If accumulate != 0 the product is added to C (used to sum K-partials locally).
 * Replaces the per-device dot of jax.lax.dot on sharded operands
 *   (case1a.py:49; case1b.py:47; case2.py:44; case3_fully_sharded.py:46;
 *    case4_gspmd_ff.py:52) and jnp.einsum("ABC,ACD->ABD") (case4_gspmd_ff.py:30).
 */
int b200_gemm_f32(const float* A, int64_t lda, int64_t strideA, const float* B, int64_t ldb, int64_t strideB,
                  float* C, int64_t ldc, int64_t strideC, int batch, int M, int N, int K, int accumulate,
                  void* stream);

/*
 * Multi-head attention core, forward:  O = softmax(scale * Q.K^T [+ causal mask]) . V
 * Q: [B,Sq,H,D], K,V: [B,Sk,H,D], O: [B,Sq,H,D]; bf16; head dim innermost, heads
 * second-innermost (the reference's "b s n h" layout, so the (b,S,N*D) projection
 * output is consumed without a transpose).  Row strides (elements) are given per
 * tensor as (batch, seq, head); the D elements of a head are contiguous.
 * Scores and softmax are f32, P is rounded to bf16 before P.V, O is bf16 --
 * the order of case6_attention.py:121-135 (online softmax, never materialises S).
 * causal != 0 masks keys t > f + (Sk - Sq) (the BASELINE.json case6 config; the
 * reference itself has no mask, case6_attention.py:125-127).  D in {64, 128}.
 */
int b200_attention_fwd(const void* Q, const void* K, const void* V, void* O, int B, int Sq, int Sk, int H, int D,
                       int64_t q_stride_b, int64_t q_stride_s, int64_t q_stride_h, int64_t k_stride_b,
                       int64_t k_stride_s, int64_t k_stride_h, int64_t v_stride_b, int64_t v_stride_s,
                       int64_t v_stride_h, int64_t o_stride_b, int64_t o_stride_s, int64_t o_stride_h, float scale,
                       int causal, void* stream);

/*
 * General form of b200_attention_fwd (zero-initialise, fill what you need): strides[12] = (batch, seq, head) element
 * strides of Q, K, V, O in that order; lse2 (optional, f32 [B, H, Sq]) as in b200_attention_fwd_lse; variant picks the
 * kernel for this call only -- 0 = library choice, 1 = one query tile per CTA (no lse2), 8 = persistent single-CTA
 * kernel, 9 = persistent CTA-pair kernel (D = 128).  The library keeps no variant setting between calls.
 */
typedef struct {
  const void *Q, *K, *V;
  void* O;
  float* lse2;
  int B, Sq, Sk, H, D;
  int64_t strides[12];
  float scale;
  int causal;
  int variant;
} b200_attention_args;
int b200_attention_fwd_ex(const b200_attention_args* args, void* stream);

/*
 * b200_attention_fwd that also returns lse2[B, H, Sq] (f32) = log2(sum_t exp2(log2(e) * scale * s[t])) per query row: the
 * residual the backward needs.  Always runs the persistent kernels (attention9.cu for D = 128, attention8.cu for D = 64).
 */
int b200_attention_fwd_lse(const void* Q, const void* K, const void* V, void* O, float* lse2, int B, int Sq, int Sk, int H,
                           int D, int64_t q_stride_b, int64_t q_stride_s, int64_t q_stride_h, int64_t k_stride_b,
                           int64_t k_stride_s, int64_t k_stride_h, int64_t v_stride_b, int64_t v_stride_s,
                           int64_t v_stride_h, int64_t o_stride_b, int64_t o_stride_s, int64_t o_stride_h, float scale,
                           int causal, void* stream);

/*
 * Blockwise ("ring") attention over a sequence-sharded K / V -- the literal layout of the reference, where the activations
 * stay split along S over 'model' (case6_attention.py:105-107,114-116) and XLA would all-gather K and V for the score
 * einsum (:125,133).  Instead every device runs b200_attention_fwd_lse on its query block against ONE key/value block at a
 * time (its own, then the blocks passed round the ring) and folds each result into a running pair with this call:
 *   lse' = log2(2^lse_acc + 2^lse_blk),  acc' = acc 2^(lse_acc - lse') + o_blk 2^(lse_blk - lse')
 * acc [B, S, H, D] f32 and lse_acc [B, H, S] f32 are the running pair (first != 0 initialises them from the block),
 * o_blk [B, S, H, D] bf16 / lse_blk [B, H, S] f32 one block's result; all contiguous.  After the last block
 * b200_cast_f32_to_bf16(acc) is the attention output.  D in {64, 128}.
 */
int b200_attention_combine(float* acc, float* lse_acc, const void* o_blk, const float* lse_blk, int B, int S, int H, int D,
                           int first, void* stream);

/*
 * Attention backward: dQ, dK, dV of O = softmax(scale Q K^T [+ causal mask]) V given dO, O and the forward's lse2 --
 * the part of jax.grad through FlaxAttention.__call__ (case6_attention.py:121-135) that the reference's train_step runs
 * (case6_attention.py:206-220).  Layouts as in b200_attention_fwd; strides (elements) for
 * q, k, v, o, dO, dQ, dK, dV as (batch, seq, head) triples in that order; dQ must be contiguous [B, Sq, H, D].
 * workspace: f32, B*Sq*H*D + B*H*Sq elements (dQ accumulator + row sums of dO*O), caller-provided.
 * P and dS are rounded to bf16 before the tensor-core products (as the forward rounds P); accumulation is f32.
 */
typedef struct {
  const void *Q, *K, *V, *O, *dO;
  const float* lse2;
  void *dQ, *dK, *dV;
  float* workspace;
  int B, Sq, Sk, H, D;
  int64_t strides[24];
  float scale;
  int causal;
} b200_attention_bwd_args;
int b200_attention_bwd(const b200_attention_bwd_args* args, void* stream);

/*
 * optax.adam update of one f32 parameter tensor, in place (scale_by_adam, then scale(-lr), then apply_updates):
 *   mu = b1 mu + (1-b1) g;  nu = b2 nu + (1-b2) g^2;
 *   param -= lr * (mu / (1 - b1^step)) / (sqrt(nu / (1 - b2^step)) + eps)
 * grad is f32 or bf16 (grad_dtype), step the 1-based update count after this call.  Replaces what
 * TrainState.apply_gradients does with optax.adam(learning_rate=0.001) in the reference's train_step
 *   (case5_attention_dense.py:107,130-139; case6_attention.py:181,206-220) -- SURVEY.md section 8(f) row 1.
 */
int b200_adam_step(float* param, const void* grad, int grad_dtype, float* mu, float* nu, int64_t n, float lr, float b1,
                   float b2, float eps, int step, void* stream);

/*
 * One device's share of the feed-forward block  y = act(x.W1 + b1).W2 (+ b2):
 * x [T,d_model] bf16, W1 [d_model,F] ([K,N] row-major), b1 [F] or NULL,
 * W2 [F,d_model], b2 [d_model] or NULL, h [T,F] bf16 caller-provided workspace,
 * y [T,d_model] bf16 (or f32 when y_dtype = B200_F32, for partial sums that a
 * row-parallel reduction will add).  F is the LOCAL d_ff (d_ff / model-axis size):
 * W1 column-parallel, W2 row-parallel; with F < d_ff the caller must all-reduce
 * or reduce-scatter y over the 'model' axis and pass b2 on one rank only.
 * act: B200_EPI_BIAS_RELU / B200_EPI_BIAS_GELU_TANH / B200_EPI_RELU / B200_EPI_GELU_TANH.
 * Reference: the FFN comment at case6_attention.py:36-40 (= case5_attention_dense.py:35-39).
 */
int b200_ffn_fwd(const void* x, const void* W1, const void* b1, const void* W2, const void* b2, void* h, void* y,
                 int y_dtype, int T, int d_model, int F, int act, void* stream);

/*
 * b200_ffn_fwd for hosts that cannot allocate device memory around the call.  (This entry point
 * has no counterpart in the reference; it exists for integrators of this library.)
 * Who needs it: a custom-call handler running under a framework whose own allocator has already
 * taken almost all of HBM (XLA preallocates it), where a cudaMalloc for the intermediate h either
 * fails or forces a device-wide synchronisation, and bindings whose operand list was fixed without
 * a scratch buffer.  Same arguments and result as b200_ffn_fwd minus `h`: the intermediate lives in
 * a 32 MiB array compiled into the library (B200_FFN_STATIC_SCRATCH_BYTES), so the call performs no
 * cudaMalloc / cudaFree and no allocating library call.
 *  - Limit: 2 * T * F bytes must fit in 32 MiB (e.g. T = 1024 tokens at F = 16384); larger requests
 *    return B200_ERR_INVALID_ARGUMENT with the sizes in b200_last_error() -- split the tokens.
 *  - Every GPU whose context loads the library carries those 32 MiB until the process exits.
 *  - The first call on each device looks up the array's device address (once, with that device
 *    current); make that first call before starting any stream capture.
 *  - All calls on a device share the one array and it keeps whatever the previous call wrote;
 *    nothing needs clearing because the first GEMM overwrites all of h that the second reads.
 *    Two calls that overlap on the GPU (different streams, or host threads racing on one device)
 *    would both write it and corrupt each other: keep calls on a device stream-ordered.
 */
#define B200_FFN_STATIC_SCRATCH_BYTES (32ull << 20)
int b200_ffn_fwd_static_scratch(const void* x, const void* W1, const void* b1, const void* W2, const void* b2, void* y,
                                int y_dtype, int T, int d_model, int F, int act, void* stream);

/*
 * Random fill of ONE DEVICE'S SHARD with the values jax.random assigns to the whole array ("born sharded": no host
 * generation, no placement copy).  key = the two uint32 words of a jax threefry key (jax.random.PRNGKey(s) = {0, s});
 * global_shape / shard_start / shard_shape describe the row-major array and this device's slice of it (ndim <= 4);
 * element i of the flattened global array is word i of jax's threefry2x32 bit stream for (key, size), transformed as
 *   B200_RNG_BITS              raw uint32 words                                   (out: uint32)
 *   B200_RNG_UNIFORM           jax.random.uniform(key, shape, minval=lo, maxval=hi)   (out: f32)
 *   B200_RNG_NORMAL            jax.random.normal(key, shape)                      (case1a.py:16-18; case6_attention.py:152)
 *   B200_RNG_TRUNCATED_NORMAL  jax.random.truncated_normal(key, lo, hi, shape); with lo = -2, hi = 2 and
 *                              scale = sqrt(1 / fan_in) / 0.87962566103423978 this is nn.initializers.lecun_normal(),
 *                              the initialiser of every Dense kernel under model.init (case6_attention.py:57,86,171-196)
 * each multiplied by `scale` (0 is read as 1).  The integer stages are bit-exact with jax; the float stages follow the
 * published formulas in f32 (XLA may differ by an ulp).  Arrays of 2^32 or more elements are not supported.
 */
typedef enum { B200_RNG_BITS = 0, B200_RNG_UNIFORM = 1, B200_RNG_NORMAL = 2, B200_RNG_TRUNCATED_NORMAL = 3 } b200_rng_distribution;
typedef struct {
  uint32_t key[2];
  int ndim;
  int64_t global_shape[4], shard_start[4], shard_shape[4];
  int distribution;
  float lo, hi;
  float scale;
} b200_random_args;
int b200_random_fill(const b200_random_args* args, void* out, void* stream);

/* dst[i] = bf16(src[i]) -- the per-call f32->bf16 parameter/activation cast flax
 * promote_dtype performs inside nn.Dense (case6_attention.py:61-90). n elements. */
int b200_cast_f32_to_bf16(const float* src, void* dst, int64_t n, void* stream);
/* dst[i] = f32(src[i]), n elements. */
int b200_cast_bf16_to_f32(const void* src, float* dst, int64_t n, void* stream);
/* dst[i] = bf16(sum_r src_r[i]) over `nsrc` f32 partial buffers laid out back to back
 * with stride `src_stride` elements; optional bf16 bias[cols] added per row.
 * Finishes a row-parallel reduction whose partials were exchanged in f32. */
int b200_sum_partials_to_bf16(const float* src, int64_t src_stride, int nsrc, const void* bias, int64_t cols,
                              void* dst, int64_t n, void* stream);

/*
 * Peer-memory plumbing of the fused GEMM -> reduce-scatter (one process per GPU on one NVSwitch box).
 * SET-UP calls (they allocate / map device memory; not for the hot path):
 *   b200_peer_alloc   cudaMalloc + zero-fill a receive buffer and return its 64-byte CUDA IPC handle;
 *   b200_peer_open    map another rank's buffer from its handle (exchanged by the host layer, e.g.
 *                     torch.distributed.all_gather); b200_peer_close / b200_peer_free undo them.
 * HOT-PATH calls (enqueue only):
 *   b200_peer_barrier every rank writes `epoch` into slot [rank] of each rank's flag block
 *                     (flag_ptrs_dev: DEVICE array of `world` pointers to uint32[world] blocks, one per
 *                     rank, inside the shared buffers) and waits until its own block shows >= epoch from
 *                     everybody; epochs must increase by one per call.  All of this GPU's earlier peer
 *                     stores on the stream are visible to the peers that pass the barrier.
 *   b200_sum_bf16_partials  dst[i] = bf16(sum_r src[r*src_stride + i]) (+ bf16 bias[i % cols]), bf16 in,
 *                     f32 accumulation: the local half of the reduce-scatter.
 * No counterpart in the reference: XLA:CPU's in-process collectives between virtual devices play
 * this role there.
 */
int b200_peer_alloc(uint64_t bytes, void** dev_ptr, void* handle64);
int b200_peer_open(const void* handle64, void** dev_ptr);
int b200_peer_close(void* dev_ptr);
int b200_peer_free(void* dev_ptr);
/* Strided device-to-device copy (rows x width_bytes) between local and/or IPC-mapped peer memory on the
 * copy engines (cudaMemcpy2DAsync): the pull half of the FSDP weight gather (all-gather before the GEMM,
 * BASELINE.json case3; reference pattern case5_attention_dense.py:59-67 with rules :109-112) -- weights are
 * read-only, so pulling needs no cross-rank synchronisation. */
int b200_peer_copy2d(void* dst, int64_t dst_pitch_bytes, const void* src, int64_t src_pitch_bytes, int64_t width_bytes,
                     int64_t rows, void* stream);
/* The same pull as ONE kernel: column chunk [src_off_bytes, +width_bytes) of every rank's [rows, src_pitch] shard
 * (src_bases_dev: DEVICE array of `world` base pointers, own shard included) into rows p*rows.. of dst.  16-byte
 * loads from peer HBM over NVLink, eight in flight per thread, `ctas` CTAs of 256 threads that co-reside with
 * the persistent GEMM.  Reaches NVLink bandwidth where strided copy-engine copies do not. */
int b200_peer_gather(void* const* src_bases_dev, int world, int rank, int64_t src_off_bytes, int64_t src_pitch_bytes,
                     void* dst, int64_t dst_pitch_bytes, int rows, int64_t width_bytes, int ctas, void* stream);
/* The PUSH form of the same all-gather: this rank's contiguous shard (`bytes` at `src`) is stored at dst_off_bytes of EVERY
 * rank's gathered buffer (dst_bases_dev: DEVICE array of `world` base pointers, own buffer included).  Peer stores are posted,
 * so a few CTAs reach NVLink's write bandwidth where pulls are bound by load latency.  The caller synchronises the ranks
 * around it (b200_peer_barrier: the destination buffers are free before, every push has landed after). */
int b200_peer_push(void* const* dst_bases_dev, int world, int rank, int64_t dst_off_bytes, const void* src, int64_t bytes,
                   int ctas, void* stream);
int b200_peer_barrier(void* const* flag_ptrs_dev, int rank, int world, uint32_t epoch, void* stream);
/* The local half of the fused reduce-scatter, running WHILE the pushing GEMMs still produce: for g = 0 .. groups-1 wait until
 * flags[s * groups + g] >= epoch for every sender s (the progress signals of b200_gemm_args), then
 * dst rows [g, g + 1) * rows_per_group = bf16(sum_s src[s * src_stride + those rows]) (+ bias).  `ctas` CTAs of 256 threads
 * that co-reside with the persistent GEMM (no shared memory).  Enqueue it on the SAME stream RIGHT AFTER the pushing
 * b200_gemm_bf16_ex: it is launched as a programmatic dependent of the kernel before it, i.e. it starts once that whole grid
 * is resident and runs beside it (starting earlier could starve a GEMM CTA of its SM and deadlock on that CTA's tiles). */
int b200_peer_progressive_sum(const uint32_t* flags, uint32_t epoch, int senders, int groups, int64_t rows_per_group,
                              const void* src, int64_t src_stride, const void* bias, int64_t cols, void* dst, int ctas,
                              void* stream);
int b200_sum_bf16_partials(const void* src, int64_t src_stride, int nsrc, const void* bias, int64_t cols, void* dst,
                           int64_t n, void* stream);

/*
 * Mesh-axis communicators and collectives for a binder that runs one process per GPU.
 * Replaces, at the device level, Mesh(create_device_mesh((2,4)), ('data','model')) (case1a.py:15;
 * case6_attention.py:155-156) and the psum / all_gather / psum_scatter / all_to_all XLA inserts around sharded
 * dots and at with_logical_constraint boundaries (case1a.py:49; case6_attention.py:137,141,229-237).
 * Transport is NCCL (libnccl.so.2 resolved with dlopen at first use; B200_ERR_UNSUPPORTED if absent).
 *   b200_comm_unique_id  rank 0 fills a 128-byte id and hands it to the other ranks by any host-side channel.
 *   b200_mesh_create     SET-UP, collective over all ranks: rank r sits at the row-major coordinates of r in
 *                        `shape` (last axis fastest: device id = 4x + y on a 2x4 mesh, as create_device_mesh).
 *   b200_mesh_axis_comm  communicator of the ranks that differ from this one only along `axis`, ordered by that
 *                        coordinate; the first call per axis is collective over the mesh (ncclCommSplit).
 *   b200_all_reduce / reduce_scatter / all_gather / all_to_all   enqueue-only on `stream`, sum reduction,
 *                        dtype B200_BF16 or B200_F32, counts in elements (see csrc/mesh_comm.cu for block layouts).
 * The hot FFN / attention layers do not use these: their reduce-scatter is fused into the GEMM epilogue (b200_peer_*).
 */
typedef struct b200_mesh b200_mesh;
typedef struct b200_comm b200_comm;
int b200_comm_unique_id(void* out128);
int b200_mesh_create(const void* unique_id128, int world, int rank, const int* shape, int naxes, b200_mesh** out);
int b200_mesh_axis_comm(b200_mesh* mesh, int axis, b200_comm** out);
int b200_comm_size(const b200_comm* comm);
int b200_comm_rank(const b200_comm* comm);
int b200_comm_release(b200_comm* comm);
int b200_mesh_destroy(b200_mesh* mesh);
int b200_all_reduce(b200_comm* comm, const void* send, void* recv, int64_t count, int dtype, void* stream);
int b200_reduce_scatter(b200_comm* comm, const void* send, void* recv, int64_t recv_count, int dtype, void* stream);
int b200_all_gather(b200_comm* comm, const void* send, void* recv, int64_t send_count, int dtype, void* stream);
int b200_all_to_all(b200_comm* comm, const void* send, void* recv, int64_t count, int dtype, void* stream);

/*
 * The tensor-parallel FFN with its reduction: b200_ffn_fwd's two GEMMs on this device's W1 columns / W2 rows, then
 * the sum over the mesh axis of `comm` -- reduce_mode B200_REDUCE_ALL: y [T, d_model] bf16 on every rank (psum);
 * B200_REDUCE_SCATTER: y [T / size, d_model] = this rank's sequence block (psum_scatter, the out_shardings of
 * case6_attention.py:229-232).  `chunks` token chunks: the collective of chunk c runs on an internal stream under
 * GEMM2 of chunk c+1 (reduce-scatter in chunks needs T / (size * chunks) % 128 == 0).  h [T, F] bf16 and partial
 * [T, d_model] f32 are caller-provided workspaces; b2 is added after the reduction (pass it on every rank).
 * The first call on a device creates the internal stream and events (set-up: before any stream capture).
 * Reference: the FFN comment case6_attention.py:36-40 with the all-reduce XLA inserts after a contraction-sharded dot
 * (case1a.py:49).  This repo's throughput path is the fused peer-memory variant (b200_gemm_bf16_ex + b200_peer_*).
 */
typedef enum { B200_REDUCE_ALL = 0, B200_REDUCE_SCATTER = 1 } b200_reduce_mode;
int b200_ffn_fwd_tp(b200_comm* comm, const void* x, const void* W1, const void* b1, const void* W2, const void* b2,
                    void* h, float* partial, void* y, int reduce_mode, int T, int d_model, int F, int act, int chunks,
                    void* stream);

#ifdef __cplusplus
}
#endif
#endif /* B200_SHARDED_H_ */
