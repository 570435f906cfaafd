// Helpers shared by the op translation units (not part of the mirrored API).
#pragma once
#include <vector>

#include "autograd_engine.h"
#include "tensor.h"

namespace ops {
namespace internal {

bool wants_grad(const TensorPtr& t);
// Registers `fn` as the producer of `out` when the engine is enabled and any input requires grad
// (the reference's `if (a->requires_grad() || ...) { set_requires_grad(true); register_backward(...) }`).
void track(const TensorPtr& out, const std::vector<TensorPtr>& inputs, BackwardFunctionPtr fn);
TensorPtr new_like(const TensorPtr& t);
void require_f32(const TensorPtr& t, const char* op);
int64_t out_ld(int64_t n);

// One GEMM in mft_gemm terms with device pointers; `b_weight` lets run_gemm reuse a cached K-major shadow.
struct GemmCall {
    int64_t m = 0, n = 0, k = 0;
    const float* a = nullptr; int64_t lda = 0; int trans_a = 0;
    const float* b = nullptr; int64_t ldb = 0; int trans_b = 0;
    float* c = nullptr; int64_t ldc = 0;
    float alpha = 1.0f, beta = 0.0f;
    const float* bias = nullptr;
    float* gelu_out = nullptr;
    const float* addend = nullptr; int64_t ld_addend = 0;             // + residual, in the epilogue
    const float* gelu_grad_x = nullptr; int64_t ld_gelu_grad_x = 0;   // * gelu'(x), in the epilogue
    int64_t k2 = 0; const float* a2 = nullptr; int64_t lda2 = 0; const float* b2 = nullptr; int64_t ldb2 = 0;
    int batch = 1; int64_t stride_a = 0, stride_b = 0, stride_c = 0;
    TensorPtr b_weight;   // tensor behind `b` (lets run_gemm reuse its cached K-major / bf16 shadows)
    TensorPtr a_tensor;   // tensor behind `a`, dense view [m or k rows] (bf16 shadow reuse); optional
    TensorPtr a2_tensor;  // tensor behind `a2` [m, k2] (bf16 shadow reuse); optional
    TensorPtr b2_tensor;  // tensor behind the rank-r operand: stored [n, k2] (b2_stored_kn = false) or [k2, n] (true); optional
    bool b2_stored_kn = false;
    TensorPtr gelu_tensor;   // tensor behind gelu_out: in bf16 mode the epilogue also writes its bf16 shadow
    // Result tensor (optional; `c` may then be left null and is resolved by run_gemm). On the bf16 tensor-core form:
    //   c_bf16_only    : the result leaves as bf16 only (Tensor::bf16_only_out) — for results whose consumers are bf16 GEMM /
    //                    attention operands; anything else that reads it gets an upcast on demand;
    //   gelu_bf16_only : same for the GELU aux output.
    // `a` may also be left null when `a_tensor` is given: the bf16 form then reads the tensor's bf16 copy without touching
    // (or materialising) its fp32 data.
    TensorPtr c_tensor;
    bool c_bf16_only = false, gelu_bf16_only = false;
};
void run_gemm(GemmCall g);

// Adapter-gradient products (dA, dB) recorded during a backward pass and launched together at its end (ops_linear.cpp).
void begin_deferred_lora_tn();
void flush_deferred_lora_tn();

// Backward node of y = gelu(x) (used by the fused linear+GELU op as well as ops::gelu).
BackwardFunctionPtr make_gelu_backward(const TensorPtr& x);

// Operand precision of the large (tensor-core) GEMMs: 0 = TF32 (default), 1 = bf16. MFT_GEMM_PRECISION=bf16|tf32.
int gemm_precision();
// true when large GEMMs run on bf16 operands: producers of GEMM inputs then also write a bf16 copy (Tensor::bf16_emit)
bool emit_bf16_operands();
// true when run_gemm will take the bf16 tensor-core form for an [m,k] x [k,n] product with a K-major A operand
bool gemm_takes_bf16(int64_t m, int64_t n, int64_t k);
void set_gemm_precision(int bf16);
// true when results consumed only as bf16 operands may skip their fp32 copy (bf16 operand mode and MFT_BF16_ONLY != 0)
bool bf16_only_results();
// True-gradient mode: GPT-2 attention with a backward + accumulating `.grad` (off by default; MFT_TRUE_ATTN_GRAD=1).
bool true_gradients();
void set_true_gradients(bool on);

}  // namespace internal
}  // namespace ops
