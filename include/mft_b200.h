/*
 * mft_b200.h — C ABI of the B200-native training-step hot path of MobileFineTuner.
 *
 * Every entry point is `extern "C"`, takes plain DEVICE pointers and sizes (no C++/torch types),
 * enqueues work on the given CUDA stream and does not synchronise. Tensors are dense row-major fp32
 * unless a leading dimension (`ld*`, in elements) is given; token ids / labels are int32.
 * Return value: 0 = ok, non-zero = error (message via mft_last_error()). There is NO CPU fallback:
 * the library is sm_100a-only and every call fails loudly when no such device is present.
 *
 * Each declaration cites the reference code it replaces (paths relative to the reference's
 * operators/finetune_ops/ unless stated otherwise). The host-side C++ mirror of the reference's
 * operator API (ops::Tensor, ops::matmul, LoRALinear, Adam, GPT2Model ...) lives in
 * mobilefinetuner_b200/csrc/host and calls only these functions.
 */
#ifndef MFT_B200_H_
#define MFT_B200_H_

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

typedef void* mft_stream_t; /* a cudaStream_t (NULL = default stream) */

/* ------------------------------------------------------------------ library / errors */
int mft_version(void);
int64_t mft_launch_count(void);            /* kernels enqueued by this library so far (process-wide) */
const char* mft_last_error(void);          /* thread-local, valid until the next failing call */
int mft_device_check(void);                /* 0 iff the current device is sm_100 (B200)        */
int mft_use_tensor_cores(int enable);      /* 1 (default): tcgen05 TF32 GEMM when eligible; 0: SIMT fp32 GEMM
                                              everywhere (bit-faithful fp32, used by parity tests). Returns previous. */
int mft_tensor_cores_enabled(void);     /* current setting of mft_use_tensor_cores, unchanged */

/* ------------------------------------------------------------------ GEMM
 * C[m,n] = alpha * ( opA(A) * opB(B)  +  A2 * B2^T )  (+ bias[n])  (+ beta * C), optional GELU aux.
 *   trans_a = 0: A stored [m,k] (lda >= k)      trans_a = 1: A stored [k,m] (lda >= m)
 *   trans_b = 0: B stored [k,n] (ldb >= n)      trans_b = 1: B stored [n,k] (ldb >= k)
 *   optional second K-segment (LoRA rank-r update fused into the same accumulator):
 *       A2 stored [m,k2] (lda2), B2 stored [n,k2] (ldb2); k2 = 0 disables it.
 *   batch > 1: strided batches (element strides), bias/second segment not batched.
 *   gelu_out != NULL: C receives the pre-activation, gelu_out (ldc) receives gelu_tanh(C).
 *   Epilogue order: v = alpha*acc + bias + beta*C + addend;  v *= gelu_tanh'(gelu_grad_x);  C = v;  gelu_out = gelu_tanh(v).
 *   addend fuses the residual add that follows a projection (ops::add after LoRALinear::forward, graph/gpt2_model.cpp
 *   block forward); gelu_grad_x fuses GeluBackward (core/backward_functions.cpp:155-177) into the dX product of the
 *   layer that consumed the activation. Both are element-wise on the fp32 result, so the fused value equals the
 *   separate kernels' bit for bit. Not available with split-K.
 * Tensor-core path (tcgen05.mma kind::tf32, fp32 accumulate in TMEM, TMA-fed) is taken when both
 * operands are K-major (trans_a = 0, trans_b = 1), 16-byte aligned with ld % 4 == 0; otherwise a
 * SIMT fp32 kernel runs. Replaces ops::matmul + MatmulBackward's three products
 * (core/ops.cpp:486-651, core/backward_functions.cpp:94-138; cblas_sgemm call sites
 * core/ops.cpp:563,613,692,759) and the add/mul nodes of LoRALinear::forward (nn/lora_linear.cpp:47-108).
 */
typedef struct mft_gemm_args {
    int32_t m, n, k;
    const float* a; int64_t lda; int32_t trans_a;
    const float* b; int64_t ldb; int32_t trans_b;
    float* c; int64_t ldc;
    float alpha, beta;
    const float* bias;                 /* [n] or NULL */
    float* gelu_out;                   /* [m,n] (ldc) or NULL */
    int32_t k2; const float* a2; int64_t lda2; const float* b2; int64_t ldb2;
    int32_t batch; int64_t stride_a, stride_b, stride_c;
    const float* addend; int64_t ld_addend;            /* [m,n] added after bias/beta (residual stream), or NULL   */
    const float* gelu_grad_x; int64_t ld_gelu_grad_x;  /* [m,n]: result *= d gelu_tanh / dx at this tensor, or NULL */
    void* gelu_out_bf16; int64_t ld_gelu_out_bf16;     /* bf16 copy of gelu_out, or NULL; on the tensor-core path gelu_out itself
                                                          may then be NULL (the activation leaves as bf16 only)      */
    void* c_bf16; int64_t ldc_bf16;                    /* bf16 copy of C (after alpha / bias / beta), or NULL. On the tensor-core path
                                                          c may then be NULL: the bf16 copy is the only result (its consumers are
                                                          bf16 GEMM / attention operands), which halves the epilogue's HBM writes */
} mft_gemm_args;
int mft_gemm(const mft_gemm_args* args, mft_stream_t stream);
/* bf16-operand form of mft_gemm: a, b, a2, b2 point to bf16 arrays (leading dimensions in elements, multiples of 8),
 * both operands K-major (trans_a = 0, trans_b = 1); accumulation and C stay fp32 (tcgen05.mma kind::f16). Twice the
 * tensor throughput of the TF32 form for the same bytes staged; inputs are rounded to bf16 by mft_cast_bf16 /
 * mft_transpose_cast_bf16 (weights: once, cached by the host layer). No SIMT fallback: ineligible arguments are an error. */
int mft_gemm_bf16(const mft_gemm_args* args, mft_stream_t stream);
/* out (bf16, ldo) = in (fp32, ldi), rows x cols, round-to-nearest-even; and the same with a transpose: out[c,r] = in[r,c]. */
int mft_cast_bf16(const float* in, int64_t ldi, void* out, int64_t ldo, int64_t rows, int64_t cols, mft_stream_t stream);
int mft_transpose_cast_bf16(const float* in, int64_t ldi, void* out, int64_t ldo, int64_t rows, int64_t cols, mft_stream_t stream);
/* Split-K form for products with a tiny output and a long contraction (LoRA's dA = x^T d(xA) and dB = (xA)^T dY, where the
 * contraction runs over all tokens): the fp32 SIMT kernel computes `splits` partial products over disjoint K ranges into
 * `workspace` (mft_gemm_splitk_workspace_bytes) and a second kernel sums them in a fixed order and applies alpha/bias/beta
 * (deterministic). Same operation as mft_gemm otherwise; k2 must be 0 when splits > 1. */
size_t mft_gemm_splitk_workspace_bytes(int32_t m, int32_t n, int32_t batch, int32_t splits);
int mft_gemm_splitk(const mft_gemm_args* args, int32_t splits, void* workspace, size_t workspace_bytes, mft_stream_t stream);
/* Rank-r LoRA factor products (r <= 64), streaming kernels bound by HBM bandwidth; they replace the matmul / MatmulBackward
 * calls LoRALinear::forward issues for slice.A / slice.B (nn/lora_linear.cpp:66-72, core/backward_functions.cpp:94-138).
 *   mft_lora_down: y[T,r] (ldy) = alpha * x[T,K] (ldx) * A, with A stored [K,r] (a_is_kr = 1) or [r,K] (a_is_kr = 0).
 *                  Used for x*A (forward) and dY*B^T (backward). x must be 16-byte aligned, K % 4 == 0, ldx % 4 == 0.
 *   mft_lora_tn  : out = alpha * S^T * W with S[T,r] (lds), W[T,q] (ldw); out is [r,q] (out_is_rq = 1) or [q,r] (0), row pitch
 *                  ldo. Used for dB = (xA)^T dY and dA = x^T d(xA). Partial sums over 64-token slabs go to `workspace`
 *                  (mft_lora_tn_workspace_bytes) and are reduced in a fixed order (deterministic). */
int mft_lora_down(const float* x, int64_t ldx, const float* a, int64_t lda, int a_is_kr, float* y, int64_t ldy,
                  int64_t T, int64_t K, int32_t r, float alpha, mft_stream_t stream);
/* same, plus a bf16 copy of y (NULL = off; row pitch ld_bf16 elements) for the rank-r segment of the bf16 GEMM */
int mft_lora_down_ex(const float* x, int64_t ldx, const float* a, int64_t lda, int a_is_kr, float* y, int64_t ldy, void* y_bf16,
                     int64_t ld_bf16, int64_t T, int64_t K, int32_t r, float alpha, mft_stream_t stream);
/* The same product on bf16 operands (bf16 operand mode): x as the bf16 copy its producer wrote, A as its K-major bf16 shadow
 * [r][K] (pitch lda) — the array the rank-r segment of mft_gemm_bf16 reads. r <= 64; K a multiple of 128, up to 6144 (r <= 16),
 * 3072 (r <= 32) or 1536 (r <= 64): the whole shadow is staged in shared memory; pitches
 * multiples of 8, 16-byte aligned bases. y is fp32 [T, r]; y_bf16 (NULL = off) its bf16 copy. */
int mft_lora_down_bf16(const void* x_bf16, int64_t ldx, const void* a_kmajor_bf16, int64_t lda, float* y, int64_t ldy, void* y_bf16,
                       int64_t ld_bf16, int64_t T, int64_t K, int32_t r, float alpha, mft_stream_t stream);
size_t mft_lora_tn_workspace_bytes(int64_t T, int64_t q, int32_t r);
int mft_lora_tn(const float* s, int64_t lds, const float* w, int64_t ldw, float* out, int64_t ldo, int out_is_rq,
                int64_t T, int64_t q, int32_t r, float alpha, void* workspace, size_t workspace_bytes, mft_stream_t stream);
/* 1 if mft_gemm would take the tcgen05 path for these arguments (host-side predicate, no launch). */
/* Grouped form of mft_lora_tn: many independent products (every adapter's dA / dB of one backward pass, deferred to its end)
 * in ONE launch. The problem table lives in DEVICE memory. Per problem: the mft_lora_tn arguments, its own `partial` scratch
 * (mft_lora_tn_workspace_bytes), colblocks = mft_lora_tn_group_colblocks(q), splits = mft_lora_tn_group_splits(T),
 * use_tickets = 1 when splits <= 64 (the last-arriving CTA of a column block sums the splits in index order) and then a
 * ticket_base such that [ticket_base, ticket_base + colblocks) ranges of different problems do not overlap and stay below
 * mft_lora_tn_group_ticket_capacity(). Only problems with mft_lora_tn_group_eligible(...) = 1 may be listed. */
typedef struct mft_lora_tn_problem {
    const float* s; int64_t lds; const float* w; int64_t ldw; float* partial; float* out; int64_t ldo;
    int64_t T, q; int32_t r, out_is_rq; float alpha; int32_t use_tickets, colblocks, splits, ticket_base;
} mft_lora_tn_problem;
int mft_lora_tn_group_eligible(int64_t T, int64_t q, int32_t r, const float* w, int64_t ldw);
int32_t mft_lora_tn_group_colblocks(int64_t q);
int32_t mft_lora_tn_group_splits(int64_t T);
int32_t mft_lora_tn_group_ticket_capacity(void);
int mft_lora_tn_group(const mft_lora_tn_problem* dev_problems, int32_t num_problems, int32_t max_colblocks, int32_t max_splits,
                      int32_t any_long_split, int64_t max_rq, mft_stream_t stream);
int mft_gemm_uses_tensor_cores(const mft_gemm_args* args);
/* Measurement aid (no reference counterpart): while enabled, every tensor-core GEMM launch is bracketed by CUDA events on
 * its stream; read returns the summed algorithmic FLOPs (2*m*n*(k+k2)*batch), summed launch milliseconds and launch count. */
int mft_gemm_profile(int enable);
int mft_gemm_profile_read(double* flops, double* ms, int64_t* launches);

/* Out-of-place 2-D transpose: out[c,r] (ldo) = in[r,c] (ldi). Replaces ops::transpose's per-element
 * index loop for the 2-D case (core/ops.cpp:844-865) — used to keep K-major shadows of weights. */
int mft_transpose2d(const float* in, int64_t ldi, float* out, int64_t ldo, int64_t rows, int64_t cols,
                    int64_t batch, int64_t stride_in, int64_t stride_out, mft_stream_t stream);
/* out[b,d2,d1,d3] = in[b,d1,d2,d3]  (ops::permute {0,2,1,3}, core/ops.cpp:886-961; its own inverse). */
int mft_permute0213(const float* in, float* out, int64_t d0, int64_t d1, int64_t d2, int64_t d3, mft_stream_t stream);
/* General permute (ops::permute's index loop, core/ops.cpp:886-961) for the layouts without a fast path: out is dense with
 * shape out_shape[ndim]; src_stride[i] is the INPUT stride (elements) of output axis i. ndim <= 8. */
int mft_permute_nd(const float* in, float* out, int32_t ndim, const int64_t* out_shape, const int64_t* src_stride, mft_stream_t stream);
/* Data-preparation helpers that the reference runs as host loops: ops::cast fp32<->int32 (core/tensor.cpp cast,
 * truncation toward zero), create_causal_mask's lower-triangular ones (core/ops.cpp:1976-1992), and the one-hot rows
 * behind nll_loss / cross_entropy_loss (core/ops.cpp:1177-1402; targets int32, or float-coded when target_is_i32 = 0). */
int mft_cast_i32(const void* in, void* out, int64_t n, int to_i32, mft_stream_t stream);
int mft_tril_ones(float* out, int64_t S, mft_stream_t stream);
int mft_one_hot(const void* target, int target_is_i32, float* out, int64_t N, int64_t C, mft_stream_t stream);

/* ------------------------------------------------------------------ elementwise (HBM streaming)
 * mft_binary: out = a (op) b with numpy broadcasting of b over a described by (outer, mid, inner):
 *   a, out are [outer, mid, inner]; b_mode 0: b same shape; 1: b is [inner] (bias); 2: b is [mid, inner]
 *   (mask [S,S] onto [B*H,S,S]); 3: b is [outer_b, 1, inner] with outer index a_outer / b_div
 *   (padding mask [B,1,1,S]). op: 0 add, 1 sub, 2 mul, 3 div (core/ops.cpp:166-484); 4 eq, 5 ne, 6 gt, 7 lt,
 *   8 ge, 9 le with 1.0/0.0 results (core/ops.cpp:1941-1992). */
int mft_binary(int op, const float* a, const float* b, float* out, int64_t outer, int64_t mid, int64_t inner,
               int b_mode, int64_t b_div, mft_stream_t stream);
/* out = a*alpha + beta   (ops::mul/add/sub/div with a scalar, core/ops.cpp:380-484; ScaleBackward :970-974) */
int mft_affine(const float* a, float* out, int64_t n, float alpha, float beta, mft_stream_t stream);
/* acc += g   (Engine::accumulate_grad, core/autograd_engine.cpp:116-140) */
int mft_accumulate(float* acc, const float* g, int64_t n, mft_stream_t stream);
int mft_fill(float* out, int64_t n, float value, mft_stream_t stream);
/* out[n] = sum over rows of g[rows,n] (ldg) — AddBackward's sum_to_shape for a bias
 * (core/backward_functions.cpp:36-64). workspace: mft_colsum_workspace_bytes(rows,n). */
size_t mft_colsum_workspace_bytes(int64_t rows, int64_t n);
int mft_colsum(const float* g, int64_t ldg, float* out, int64_t rows, int64_t n, void* workspace,
               size_t workspace_bytes, mft_stream_t stream);
/* unary ops: kind 0 gelu_tanh fwd, 1 relu, 2 sigmoid, 3 silu, 4 tanh, 5 exp, 6 log, 7 sqrt, 8 abs
 * (core/ops.cpp:1033-1117). */
int mft_unary(int kind, const float* x, float* y, int64_t n, mft_stream_t stream);
/* parameterised unary ops: kind 0 pow(x, p0), kind 1 clamp(x, p0, p1) (ops::pow / ops::clamp, core/ops.cpp). */
int mft_unary_param(int kind, const float* x, float* y, int64_t n, float p0, float p1, mft_stream_t stream);
/* GELU (tanh form, constants 0.7978845608f / 0.044715f): core/ops.cpp:1055-1079, GeluBackward
 * core/backward_functions.cpp:155-177. */
int mft_gelu_fwd(const float* x, float* y, int64_t n, mft_stream_t stream);
int mft_gelu_bwd(const float* x, const float* gy, float* gx, int64_t n, mft_stream_t stream);
/* MulBackward for same-shape operands: ga = gy*b, gb = gy*a (core/backward_functions.cpp:74-83). */
/* mft_gelu_bwd plus a dense bf16 copy of gx (NULL = off) for the bf16 GEMM that consumes it; no reference counterpart. */
int mft_gelu_bwd_ex(const float* x, const float* gy, float* gx, void* gx_bf16, int64_t n, mft_stream_t stream);
/* GeGLU of the Gemma MLP in one pass: y = gelu_tanh(gate) * up (ops::gelu + ops::mul, graph/gemma_model.cpp:534-551) and its
 * backward d gate = (gy * up) * gelu'(gate), d up = gy * gelu(gate) (MulBackward + GeluBackward, core/backward_functions.cpp:66-83,
 * 155-177). The *_bf16 pointers (NULL = off) receive bf16 copies for the GEMMs that consume them. n % 4 == 0, 16-byte aligned. */
int mft_geglu_fwd(const float* gate, const float* up, float* y, void* y_bf16, int64_t n, mft_stream_t stream);
int mft_geglu_bwd(const float* gate, const float* up, const float* gy, float* dgate, float* dup, void* dgate_bf16, void* dup_bf16, int64_t n,
                  mft_stream_t stream);
/* GeluBackward (core/backward_functions.cpp:155-177) for the bf16 operand mode, with the incoming gradient and the result as
 * bf16 arrays (they are a bf16-only GEMM result and a bf16 GEMM operand): x is fp32, or bf16 when x_is_bf16; gx (fp32) is
 * optional. n % 8 == 0, 16-byte aligned pointers. 8 (or 6) bytes per element instead of 14. */
int mft_gelu_bwd_lowp(const void* x, int x_is_bf16, const void* gy_bf16, float* gx, void* gx_bf16, int64_t n, mft_stream_t stream);
/* bf16 -> fp32 (row pitches in elements): materialises the fp32 values of a tensor that a kernel produced as bf16 only. */
int mft_cast_from_bf16(const void* in, int64_t ldi, float* out, int64_t ldo, int64_t rows, int64_t cols, mft_stream_t stream);
int mft_mul_bwd(const float* a, const float* b, const float* gy, float* ga, float* gb, int64_t n, mft_stream_t stream);

/* ------------------------------------------------------------------ normalisation
 * LayerNorm over the last dim (rows x dim): y = (x-mu)*rstd*gamma + beta, biased variance, fp32
 * accumulation. Forward: ops::layer_norm core/ops.cpp:1404-1458. Backward recomputes mu/rstd and
 * always produces dx, dgamma, dbeta: LayerNormBackward core/backward_functions.cpp:425-481.
 * The backward needs mft_norm_bwd_workspace_bytes(rows, dim) bytes of scratch for the column
 * reductions (dgamma/dbeta partial sums); dgamma/dbeta are overwritten, not accumulated. dgamma and dbeta may both be
 * NULL (frozen LayerNorm parameters, LoRA runs): only dx is produced and no workspace is needed. */
int mft_layernorm_fwd(const float* x, const float* gamma, const float* beta, float* y, int64_t rows,
                      int64_t dim, float eps, mft_stream_t stream);
size_t mft_norm_bwd_workspace_bytes(int64_t rows, int64_t dim);
int mft_layernorm_bwd(const float* x, const float* gamma, const float* gy, float* gx, float* dgamma,
                      float* dbeta, int64_t rows, int64_t dim, float eps, void* workspace,
                      size_t workspace_bytes, mft_stream_t stream);
/* The residual adds on either side of a LayerNorm folded into it (same fp32 values as the separate ops::add / gradient
 * accumulation, bit for bit):
 *   mft_add_layernorm_fwd : sum_out = x + add (ops::add, graph/gpt2_model.cpp block forward) and y = LayerNorm(sum_out)
 *   mft_layernorm_bwd_add : gx = dX + addend, where addend is the gradient that already reached x through the residual
 *                           branch (Engine::accumulate_grad, core/autograd_engine.cpp:116-140). */
int mft_add_layernorm_fwd(const float* x, const float* add, float* sum_out, const float* gamma, const float* beta, float* y,
                          int64_t rows, int64_t dim, float eps, mft_stream_t stream);
int mft_layernorm_bwd_add(const float* x, const float* gamma, const float* gy, const float* addend, float* gx, float* dgamma,
                          float* dbeta, int64_t rows, int64_t dim, float eps, void* workspace, size_t workspace_bytes,
                          mft_stream_t stream);
/* Supersets for the bf16 GEMM mode (no reference counterpart: the reference is fp32 throughout): `add`/`sum_out`/`addend` may
 * be NULL; y_bf16 / gx_bf16 (NULL = off) receive a bf16 copy of the fp32 result, row pitch ld_bf16 elements (multiple of 4,
 * base 8-byte aligned), written by the same kernel so the tensor-core GEMM that consumes the result needs no cast pass. */
int mft_layernorm_fwd_ex(const float* x, const float* add, float* sum_out, const float* gamma, const float* beta, float* y,
                         void* y_bf16, int64_t ld_bf16, int64_t rows, int64_t dim, float eps, mft_stream_t stream);
int mft_layernorm_bwd_ex(const float* x, const float* gamma, const float* gy, const float* addend, float* gx, void* gx_bf16,
                         int64_t ld_bf16, float* dgamma, float* dbeta, int64_t rows, int64_t dim, float eps, void* workspace,
                         size_t workspace_bytes, mft_stream_t stream);
/* LayerNorm backward for callers that cannot allocate device memory. It has no counterpart in the reference.
 * Who needs it: a front end over this ABI (cgo, JNI, ctypes) that merely forwards tensors owned by its host
 * application has no device allocator to take workspace from, and a steady-state loop may ban cudaMalloc because
 * it stalls the whole device. Same arguments and results as mft_layernorm_bwd minus the workspace: the partial
 * sums go to an array built into the library (MFT_LN_BWD_STATIC_SCRATCH_BYTES, resident on every GPU that loads the
 * library until the process exits). Limit: mft_norm_bwd_workspace_bytes(rows, dim) must not exceed that size
 * (e.g. dim <= 2048 with rows <= 18944, dim <= 1024 with rows <= 37888); larger problems are rejected with an error
 * and must use mft_layernorm_bwd. Nothing is allocated or freed by this call. The first call on each device also
 * looks up the array's device address, so make that first call before starting any stream capture. All calls share
 * the one array (its previous contents are simply overwritten, nothing needs clearing), so two calls that may run at
 * the same time on the GPU - different streams, or different host threads - race on it: serialise them. */
#define MFT_LN_BWD_STATIC_SCRATCH_BYTES 4849664
int mft_layernorm_bwd_static(const float* x, const float* gamma, const float* gy, float* gx, float* dgamma,
                             float* dbeta, int64_t rows, int64_t dim, float eps, mft_stream_t stream);
/* RMSNorm with (1+w) scale, double-precision row statistics: y = x * rsqrt(mean(x^2)+eps) * (1+w)
 * (ops::rms_norm core/ops.cpp:1489-1537, RMSNormBackward core/backward_functions.cpp:483-523).
 * plus_one = 0 gives rms_norm_affine (y = x_hat * w). dw may be NULL in mft_rmsnorm_bwd (frozen weight): only dx is
 * produced and no workspace is needed. */
int mft_rmsnorm_fwd(const float* x, const float* w, float* y, int64_t rows, int64_t dim, float eps,
                    int plus_one, mft_stream_t stream);
/* same, plus a bf16 copy of y (NULL = off; row pitch ld_bf16 >= dim) for the bf16 GEMM operand mode; no reference counterpart */
int mft_rmsnorm_fwd_ex(const float* x, const float* w, float* y, void* y_bf16, int64_t ld_bf16, int64_t rows, int64_t dim,
                       float eps, int plus_one, mft_stream_t stream);
int mft_rmsnorm_bwd(const float* x, const float* w, const float* gy, float* gx, float* dw, int64_t rows,
                    int64_t dim, float eps, int plus_one, void* workspace, size_t workspace_bytes,
                    mft_stream_t stream);
/* same, plus a bf16 copy of gx (NULL = off) for the bf16 dX products that read the gradient */
int mft_rmsnorm_bwd_ex(const float* x, const float* w, const float* gy, float* gx, void* gx_bf16, int64_t ld_bf16, float* dw, int64_t rows,
                       int64_t dim, float eps, int plus_one, void* workspace, size_t workspace_bytes, mft_stream_t stream);

/* ------------------------------------------------------------------ softmax / attention / rope
 * Softmax over the last dim: core/ops.cpp:1119-1175; backward dx = y*(gy - sum(gy*y)) with the row
 * dot accumulated in double: SoftmaxBackward core/backward_functions.cpp:297-340. */
int mft_softmax_fwd(const float* x, float* y, int64_t rows, int64_t dim, mft_stream_t stream);
int mft_softmax_bwd(const float* y, const float* gy, float* gx, int64_t rows, int64_t dim, mft_stream_t stream);
/* Causal self-attention forward, GPT-2 semantics (scores*scale, j>i masked with -1e10, softmax, P*V;
 * no gradient): memory_efficient_attention core/memory_efficient_attention.cpp:40-185.
 * q/k/v are addressed as  base + b*stride_b + s*stride_s + h*stride_h + d  so the kernel reads the
 * fused [B,S,3C] qkv buffer directly (replacing the zeros+memcpy split and the two permutes of
 * graph/gpt2_model.cpp:616-664) and writes ctx as [B,S,H*Hd] (replacing :739-740).
 * window > 0 additionally masks i-j >= window (Gemma sliding layers, graph/gemma_model.cpp:272-285). */
/* Precision of both attention products: 0 (default) = TF32 operands like the tcgen05 GEMMs (relative error ~1e-3);
 * 1 = 3xTF32 split, fp32-accurate (used by the exact-mode parity tests; also selected by MFT_USE_TENSOR_CORES=0).
 * Returns the previous setting. */
int mft_attention_precise(int enable);
/* 1: head_dim-64 attention runs on bf16 operands (mma.sync m16n8k16, fp32 accumulate and softmax) — the attention side of the
 * bf16 GEMM operand mode; ignored while the precise mode is on. Returns the previous setting. No reference counterpart. */
int mft_attention_bf16(int enable);
/* Current settings without changing them: bit 0 = precise, bit 1 = bf16 operands. */
int mft_attention_mode(void);
int mft_causal_attention_fwd(const float* q, const float* k, const float* v, float* ctx, int32_t B, int32_t H,
                             int32_t S, int32_t Hd, int64_t q_stride_b, int64_t q_stride_s, int64_t q_stride_h,
                             int64_t kv_stride_b, int64_t kv_stride_s, int64_t kv_stride_h,
                             int64_t o_stride_b, int64_t o_stride_s, int64_t o_stride_h,
                             float scale, int32_t window, mft_stream_t stream);
/* same, plus a bf16 copy of ctx (NULL = off; same element strides as ctx) for the bf16 GEMM of the output projection */
int mft_causal_attention_fwd_ex(const float* q, const float* k, const float* v, float* ctx, void* ctx_bf16, int32_t B, int32_t H,
                                int32_t S, int32_t Hd, int64_t q_stride_b, int64_t q_stride_s, int64_t q_stride_h,
                                int64_t kv_stride_b, int64_t kv_stride_s, int64_t kv_stride_h,
                                int64_t o_stride_b, int64_t o_stride_s, int64_t o_stride_h,
                                float scale, int32_t window, mft_stream_t stream);
/* The same causal attention forward (core/memory_efficient_attention.cpp:92-162) on the tcgen05 tensor cores, for the bf16
 * operand mode and head_dim 64: q/k/v are the three column blocks of ONE bf16 [B, S, 3C] buffer (the qkv projection's bf16
 * output; row pitch ld elements, a multiple of 8), loaded by TMA; S = Q K^T and P V run as tcgen05.mma with fp32 accumulators
 * in TMEM; scores, softmax and the output stay fp32. ctx (fp32) and ctx_bf16 are [B, S, C]; either may be NULL. */
int mft_causal_attention_tc_fwd(const void* qkv_bf16, int64_t ld, float* ctx, void* ctx_bf16, int32_t B, int32_t H, int32_t S,
                                int32_t Hd, float scale, mft_stream_t stream);
/* Grouped-query attention with a backward, forward and gradients without the [B,H,S,S] tensor: the whole chain of
 * graph/gemma_model.cpp:481-505 — repeat_kv_heads (core/ops.cpp:2072-2149: query head h uses kv head h / (H/KV)), Q K^T,
 * * scale, + causal / sliding mask (finite -1e10, j <= i and, for window > 0, i - j < window: graph/gemma_model.cpp:261-285),
 * + key_bias[b, j] (the [B,1,1,S] padding mask; NULL = none), softmax (core/ops.cpp:1119-1175), P V — and its backward
 * (MatmulBackward, SoftmaxBackward core/backward_functions.cpp:297-340, ScaleBackward :970-974, RepeatKVHeadsBackward
 * :683-716 = sum over the group). q [B,S,H,Hd], k/v [B,S,KV,Hd], ctx [B,S,H,Hd] by (batch, seq, head) strides in elements
 * (multiples of 4, 16-byte aligned bases); Hd in {16,32,64,128,256}. lse [B,H,S] receives the per-row log-sum-exp the
 * backward needs (NULL when no backward will run); ctx_bf16 (NULL = off) is a bf16 copy with ctx's strides. With KV == H and
 * head_dim 64 this is also the true-gradient attention of GPT-2 (core/memory_efficient_attention.cpp:92-162 has no backward).
 * Precision follows mft_attention_precise. The backward is deterministic (no atomics); its workspace holds rowsum(dO*O) and,
 * for grouped heads, the per-query-head dK / dV that a last kernel sums in head order. */
int mft_gqa_attention_fwd(const float* q, const float* k, const float* v, const float* key_bias, float* ctx, float* lse, void* ctx_bf16,
                          int32_t B, int32_t H, int32_t KV, int32_t S, int32_t Hd,
                          int64_t q_stride_b, int64_t q_stride_s, int64_t q_stride_h,
                          int64_t kv_stride_b, int64_t kv_stride_s, int64_t kv_stride_h,
                          int64_t o_stride_b, int64_t o_stride_s, int64_t o_stride_h,
                          float scale, int32_t window, mft_stream_t stream);
size_t mft_gqa_attention_bwd_workspace_bytes(int32_t B, int32_t H, int32_t KV, int32_t S, int32_t Hd);
/* dctx has ctx's strides; dq / (dk, dv) are written with their own strides (every element of the three is overwritten). */
int mft_gqa_attention_bwd(const float* q, const float* k, const float* v, const float* key_bias, const float* ctx, const float* lse,
                          const float* dctx, float* dq, float* dk, float* dv,
                          int32_t B, int32_t H, int32_t KV, int32_t S, int32_t Hd,
                          int64_t q_stride_b, int64_t q_stride_s, int64_t q_stride_h,
                          int64_t kv_stride_b, int64_t kv_stride_s, int64_t kv_stride_h,
                          int64_t o_stride_b, int64_t o_stride_s, int64_t o_stride_h,
                          int64_t dq_stride_b, int64_t dq_stride_s, int64_t dq_stride_h,
                          int64_t dkv_stride_b, int64_t dkv_stride_s, int64_t dkv_stride_h,
                          float scale, int32_t window, void* workspace, size_t workspace_bytes, mft_stream_t stream);
/* Per-head RMSNorm (1 + w) then interleaved-pair RoPE for the query AND key projections of a layer in one launch, on the
 * projections' own [tokens, heads, Hd] layout (position of a row = token index % S): rms_norm core/ops.cpp:1489-1537 +
 * apply_rope core/ops.cpp:2151-2225 as called by graph/gemma_model.cpp:383-420 (the permutes around them disappear).
 * Either side may be NULL. do_norm = 0 applies RoPE only (DISABLE_QK_NORM=1). Backward for one tensor: inverse rotation
 * (ApplyRoPEBackward core/backward_functions.cpp:718-763) then RMSNormBackward (:483-523) for dx; the norm weight's own
 * gradient is not produced (use mft_rmsnorm_bwd + mft_rope when it trains). */
int mft_qk_norm_rope_fwd(const float* q, const float* wq, float* yq, int64_t q_tokens, int32_t q_heads,
                         const float* k, const float* wk, float* yk, int64_t k_tokens, int32_t k_heads,
                         int32_t S, int32_t Hd, float eps, float theta, int do_norm, mft_stream_t stream);
int mft_qk_norm_rope_bwd(const float* x, const float* w, const float* gy, float* gx, int64_t tokens, int32_t heads,
                         int32_t S, int32_t Hd, float eps, float theta, int do_norm, mft_stream_t stream);
/* Interleaved-pair RoPE on [B,H,S,Hd]: pair (2d,2d+1) rotated by pos*theta^(-2d/Hd)
 * (ops::apply_rope core/ops.cpp:2151-2225). inverse != 0 applies the transpose rotation
 * (ApplyRoPEBackward core/backward_functions.cpp:718-763). */
int mft_rope(const float* x, float* y, int64_t BH, int64_t S, int64_t Hd, float theta, int inverse, mft_stream_t stream);
/* GQA head repeat [B,KV,S,Hd] -> [B,KV*rep,S,Hd] (head h reads kv head h/rep) and its backward (sum over
 * the repeats): ops::repeat_kv_heads core/ops.cpp:2072-2149, RepeatKVHeadsBackward backward_functions.cpp:683-716. */
int mft_repeat_kv_fwd(const float* x, float* y, int64_t B, int64_t KV, int64_t rep, int64_t S, int64_t Hd, mft_stream_t stream);
int mft_repeat_kv_bwd(const float* gy, float* gx, int64_t B, int64_t KV, int64_t rep, int64_t S, int64_t Hd, mft_stream_t stream);

/* ------------------------------------------------------------------ embedding
 * GPT-2: x[t,:] = wte[ids[t],:] + wpe[t % S,:]  (graph/gpt2_model.cpp:346-381,500-528).
 * Gemma (wpe = NULL): x[t,:] = embed[ids[t],:] * scale (graph/gemma_model.cpp:222-259).
 * An out-of-range id sets *status (device int, may be NULL) to 1; the row is zero-filled. */
int mft_embedding_fwd(const int32_t* ids, const float* wte, const float* wpe, float* x, int64_t tokens,
                      int64_t S, int64_t C, int64_t vocab, float scale, int32_t* status, mft_stream_t stream);

/* ------------------------------------------------------------------ shifted LM cross-entropy
 * Row (b,s), s < S-1, is scored against labels[b,s+1]; rows whose label is ignore_index or outside
 * [0,V) are skipped; logsumexp with a double sum; mean over the valid rows
 * (ops::lm_cross_entropy core/lm_loss.cpp:106-210). logits has leading dimension ld (>= V).
 *   fwd: loss_out[0] = reduced loss ("mean": reduction=0, "sum": 1), valid_out[0] = number of valid
 *        rows (int32), row_lse[B*S] = max+log(sum) per row (saved for the backward).
 *   bwd: dlogits[r,v] = (exp(x-lse) - [v==y]) * gout[0] / valid  for valid rows, 0 elsewhere
 *        (LMCrossEntropyBackward core/lm_loss.cpp:19-104). dlogits has leading dimension ldg.
 * workspace: mft_lm_ce_workspace_bytes(B*S) for the per-row losses reduced in a fixed order. */
size_t mft_lm_ce_workspace_bytes(int64_t rows);
int mft_lm_ce_fwd(const float* logits, int64_t ld, const int32_t* labels, int32_t B, int32_t S, int32_t V,
                  int32_t ignore_index, int32_t reduction, float* loss_out, int32_t* valid_out, float* row_lse,
                  void* workspace, size_t workspace_bytes, mft_stream_t stream);
int mft_lm_ce_bwd(const float* logits, int64_t ld, const int32_t* labels, const float* row_lse,
                  const int32_t* valid, const float* gout, int32_t B, int32_t S, int32_t V, int32_t ignore_index,
                  int32_t reduction, float* dlogits, int64_t ldg, mft_stream_t stream);
/* same, plus a bf16 copy of dlogits (NULL = off; row pitch ld_bf16 elements) for the bf16 lm_head products */
int mft_lm_ce_bwd_ex(const float* logits, int64_t ld, const int32_t* labels, const float* row_lse, const int32_t* valid,
                     const float* gout, int32_t B, int32_t S, int32_t V, int32_t ignore_index, int32_t reduction,
                     float* dlogits, int64_t ldg, void* dlogits_bf16, int64_t ld_bf16, mft_stream_t stream);

/* ------------------------------------------------------------------ clip + Adam (multi-tensor)
 * A tensor list is an array of mft_tensor_slot in DEVICE memory (param/grad/m/v pointers + length).
 * mft_grad_sumsq: out[0] = sum over all slots of grad^2 (double accumulation, fixed reduction order),
 *   the quantity under the sqrt of the trainers' clip (gpt2_lora_finetune/main.cpp:491-516,
 *   optim/gemma_trainer.cpp:84-102). workspace: mft_grad_sumsq_workspace_bytes().
 * mft_clip_scale: scale_out[0] = (max_norm > 0 && norm > max_norm) ? max_norm/(norm+1e-6) : 1, norm_out[0] = norm,
 *   with norm = sqrt(sumsq[0]) * pre_scale (pre_scale = 1/world_size after a summed allreduce).
 *   scale_out already includes pre_scale, so it can be handed to mft_adam_step / mft_scale_grads as is.
 * mft_scale_grads: grad *= scale[0] in place (host-visible clip semantics).
 * mft_adam_step: Adam with L2-coupled weight decay exactly as optim/adam.cpp:25-90:
 *   g' = g*gscale + wd*p; m = b1*m+(1-b1)g'; v = b2*v+(1-b2)g'^2; p -= lr*(m/bc1)/(sqrt(v/bc2)+eps),
 *   bc = 1-beta^t with std::pow(float) semantics, t = step_per_slot (each slot carries its own step
 *   count, advanced by the host only for tensors that have a gradient, adam.cpp:35-43).
 *   gscale = grad_scale[0] (device scalar, NULL = 1) so the clip factor never visits the host. */
typedef struct mft_tensor_slot {
    float* param; float* grad; float* m; float* v;
    int64_t n;
    float bias_corr1, bias_corr2;     /* 1-beta1^t, 1-beta2^t for this tensor's step count */
    int64_t bucket_offset;            /* element offset of this tensor in a flat gradient bucket (mft_pack_grads) */
} mft_tensor_slot;
/* Data-parallel exchange: pack every gradient (zeros where grad == NULL) into one flat fp32 bucket that a single
 * all-reduce can sum; no counterpart in the reference (it has no distributed code). */
int mft_pack_grads(const mft_tensor_slot* slots, int32_t num_slots, float* bucket, mft_stream_t stream);
size_t mft_grad_sumsq_workspace_bytes(void);
int mft_grad_sumsq(const mft_tensor_slot* slots, int32_t num_slots, double* out, void* workspace,
                   size_t workspace_bytes, mft_stream_t stream);
int mft_clip_scale(const double* sumsq, float max_norm, float pre_scale, float* scale_out, float* norm_out,
                   mft_stream_t stream);
int mft_scale_grads(const mft_tensor_slot* slots, int32_t num_slots, const float* scale, mft_stream_t stream);
int mft_adam_step(const mft_tensor_slot* slots, int32_t num_slots, float lr, float beta1, float beta2, float eps,
                  float weight_decay, const float* grad_scale, mft_stream_t stream);
/* Same update with the per-tensor step counters kept in device memory (`steps`, one int32 per slot, zero-initialised by the
 * caller): they are advanced for every slot that has a gradient and the bias corrections 1-beta^t are evaluated on the
 * device, so a step needs no per-step host data (lets the whole step be replayed as a CUDA graph). `lr` is one float in
 * device memory for the same reason: a schedule rewrites it between steps without touching the captured launch. */
int mft_adam_step_dev(const mft_tensor_slot* slots, int32_t num_slots, const float* lr, float beta1, float beta2, float eps,
                      float weight_decay, const float* grad_scale, int32_t* steps, mft_stream_t stream);

/* bf16 shadows of trained matrices after the optimizer step, all tensors in ONE launch (bf16 operand mode): for each slot,
 * plain[r, c] = bf16(src[r, c]) and / or transposed[c, r] = bf16(src[r, c]) (either pointer may be NULL). The slot table is in
 * DEVICE memory. max_elems = the largest rows*cols of the table (sizes the grid). No counterpart in the reference (fp32 only). */
typedef struct mft_shadow_slot {
    const float* src; int64_t rows, cols, ld_src;
    void* plain; int64_t ld_plain;
    void* transposed; int64_t ld_transposed;
} mft_shadow_slot;
int mft_refresh_bf16_shadows(const mft_shadow_slot* slots, int32_t num_slots, int64_t max_elems, mft_stream_t stream);

#ifdef __cplusplus
}
#endif
#endif /* MFT_B200_H_ */
