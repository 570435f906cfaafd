// ops:: free functions — API mirror of the reference's core/ops.h:24-189 (same names, argument meaning and
// error behaviour: shape errors throw ops::TensorError / std::runtime_error). Every body enqueues hand-written
// CUDA kernels through the C ABI (include/mft_b200.h); nothing computes on the host.
#pragma once
#include <utility>
#include <memory>
#include <string>
#include <vector>

#include "tensor.h"

namespace ops {

// arithmetic (numpy broadcasting for tensor-tensor forms)
TensorPtr add(const TensorPtr& a, const TensorPtr& b);
TensorPtr sub(const TensorPtr& a, const TensorPtr& b);
TensorPtr mul(const TensorPtr& a, const TensorPtr& b);
TensorPtr div(const TensorPtr& a, const TensorPtr& b);
TensorPtr add(const TensorPtr& t, float s);
TensorPtr sub(const TensorPtr& t, float s);
TensorPtr mul(const TensorPtr& t, float s);
TensorPtr div(const TensorPtr& t, float s);
TensorPtr add(float s, const TensorPtr& t);
TensorPtr sub(float s, const TensorPtr& t);
TensorPtr mul(float s, const TensorPtr& t);
TensorPtr div(float s, const TensorPtr& t);

// linear algebra
TensorPtr matmul(const TensorPtr& a, const TensorPtr& b);
TensorPtr matmul_rhs_T(const TensorPtr& a, const TensorPtr& b);   // a @ b^T without materialising b^T
TensorPtr transpose(const TensorPtr& t, int dim0, int dim1);
TensorPtr permute(const TensorPtr& t, const std::vector<int>& dims);
TensorPtr linear(const TensorPtr& input, const TensorPtr& weight, const TensorPtr& bias = nullptr);
TensorPtr lora_linear(const TensorPtr& input, const TensorPtr& weight, const TensorPtr& lora_A,
                      const TensorPtr& lora_B, float alpha = 1.0f, const TensorPtr& bias = nullptr);

// activations
TensorPtr relu(const TensorPtr& x);
TensorPtr gelu(const TensorPtr& x);
TensorPtr silu(const TensorPtr& x);
TensorPtr sigmoid(const TensorPtr& x);
TensorPtr tanh_op(const TensorPtr& x);
inline TensorPtr tanh(const TensorPtr& x) { return tanh_op(x); }
TensorPtr softmax(const TensorPtr& x, int dim = -1);
TensorPtr log_softmax(const TensorPtr& x, int dim = -1);

// losses
TensorPtr mse_loss(const TensorPtr& input, const TensorPtr& target, const std::string& reduction = "mean");
TensorPtr cross_entropy_loss(const TensorPtr& input, const TensorPtr& target, const std::string& reduction = "mean");
TensorPtr nll_loss(const TensorPtr& input, const TensorPtr& target, const std::string& reduction = "mean");

// normalisation
TensorPtr layer_norm(const TensorPtr& input, const TensorPtr& weight, const TensorPtr& bias, float eps = 1e-5f);
// B200 extension: as layer_norm, result possibly written as bf16 only (bf16 operand mode, consumer is a GEMM operand)
TensorPtr layer_norm_lowp(const TensorPtr& input, const TensorPtr& weight, const TensorPtr& bias, float eps, bool y_bf16_only);
TensorPtr rms_norm(const TensorPtr& input, const TensorPtr& weight, float eps = 1e-6f);         // scale (1 + w)
TensorPtr rms_norm_post(const TensorPtr& input, const TensorPtr& weight, float eps = 1e-6f);    // rms_norm of a projection's output
TensorPtr rms_norm_affine(const TensorPtr& input, const TensorPtr& weight, float eps = 1e-6f);  // scale w
TensorPtr batch_norm(const TensorPtr& input, const TensorPtr& weight, const TensorPtr& bias,
                     const TensorPtr& running_mean = nullptr, const TensorPtr& running_var = nullptr,
                     bool training = true, float momentum = 0.1f, float eps = 1e-5f);

// shape
TensorPtr reshape(const TensorPtr& t, const std::vector<int64_t>& shape);
TensorPtr view(const TensorPtr& t, const std::vector<int64_t>& shape);
TensorPtr flatten(const TensorPtr& t, int start_dim = 0, int end_dim = -1);
TensorPtr squeeze(const TensorPtr& t, int dim = -1);
TensorPtr unsqueeze(const TensorPtr& t, int dim);

// dtype
TensorPtr cast(const TensorPtr& t, DType target_dtype);
inline TensorPtr to_fp16(const TensorPtr& t) { return cast(t, DType::kFloat16); }
inline TensorPtr to_fp32(const TensorPtr& t) { return cast(t, DType::kFloat32); }

// reductions: dim == -1 means "all elements" (reference quirk)
TensorPtr sum(const TensorPtr& t, int dim = -1, bool keepdim = false);
TensorPtr mean(const TensorPtr& t, int dim = -1, bool keepdim = false);

// comparisons (1.0 / 0.0 results)
TensorPtr eq(const TensorPtr& a, const TensorPtr& b);
TensorPtr ne(const TensorPtr& a, const TensorPtr& b);
TensorPtr gt(const TensorPtr& a, const TensorPtr& b);
TensorPtr lt(const TensorPtr& a, const TensorPtr& b);
TensorPtr ge(const TensorPtr& a, const TensorPtr& b);
TensorPtr le(const TensorPtr& a, const TensorPtr& b);

TensorPtr dropout(const TensorPtr& t, float p = 0.5f, bool training = true);
TensorPtr uniform(const std::vector<int64_t>& shape, float low = 0.0f, float high = 1.0f, DType dtype = kFloat32,
                  Device device = kCPU);
inline TensorPtr unifor(const std::vector<int64_t>& shape, float low = 0.0f, float high = 1.0f, DType dtype = kFloat32,
                        Device device = kCPU) { return uniform(shape, low, high, dtype, device); }

TensorPtr abs(const TensorPtr& t);
TensorPtr sqrt(const TensorPtr& t);
TensorPtr exp(const TensorPtr& t);
TensorPtr log(const TensorPtr& t);
TensorPtr pow(const TensorPtr& t, float exponent);
TensorPtr clamp(const TensorPtr& t, float min_val, float max_val);

// attention helpers
TensorPtr create_causal_mask(int seq_len, DType dtype = kFloat32, Device device = kCPU);
TensorPtr apply_mask(const TensorPtr& input, const TensorPtr& mask, float mask_value = -1e9f);
TensorPtr repeat_kv_heads(const TensorPtr& kv, int repeat_factor);
TensorPtr apply_rope(const TensorPtr& x, int seq_len, int head_dim, float rope_theta = 10000.0f);
TensorPtr swiglu(const TensorPtr& gate, const TensorPtr& up);
TensorPtr geglu(const TensorPtr& gate, const TensorPtr& up);          // B200 extension: gelu(gate) * up as one node / one kernel

bool same_shape(const TensorPtr& a, const TensorPtr& b);
bool broadcastable(const TensorPtr& a, const TensorPtr& b);
std::vector<int64_t> broadcast_shape(const TensorPtr& a, const TensorPtr& b);
std::vector<int64_t> infer_broadcast_shape(const TensorPtr& a, const TensorPtr& b);

// ---- B200 extensions used by the model graphs (no counterpart in the reference API) ----
// Dense copy of a pitched tensor (identity for dense ones).
TensorPtr contiguous(const TensorPtr& t);
// One autograd node for y = x W (+b) + sum_i s_i (x A_i) B_i  — replaces the six-node chain LoRALinear::forward
// builds in the reference (nn/lora_linear.cpp:47-108). A/B accept both [in,r]/[r,out] and PEFT [r,in]/[out,r].
struct LoraTerm { TensorPtr A, B; float scale; };
// When `gelu_act` is given, the GEMM epilogue also writes gelu_tanh(y) there (its own autograd node, GeluBackward), which
// replaces the separate ops::gelu pass of the MLP (graph/gpt2_model.cpp:794-802).
TensorPtr fused_lora_linear(const TensorPtr& x, const TensorPtr& W, const TensorPtr& bias, const std::vector<LoraTerm>& terms,
                            TensorPtr* gelu_act = nullptr);
// Element-wise neighbours of the projection folded into its GEMM epilogues (same fp32 values as the separate ops):
//   residual       : y = linear(x) + residual — the ops::add that follows attn.proj / mlp.fc_out in the block forward
//                    (graph/gpt2_model.cpp:760-812); the node passes gy through to `residual` like AddBackward.
//   gelu_input_pre : x is gelu(pre) produced by a previous fused_lora_linear(..., &gelu_act). The node then takes `pre` as
//                    its autograd input and its dX product returns d pre = (gy W^T + ...) * gelu'(pre) directly
//                    (GeluBackward, core/backward_functions.cpp:155-177, applied in the epilogue), so the activation's
//                    gradient is never written to HBM. Only valid when this op is the activation's single consumer.
// sum = x + a; returns layer_norm(sum) and hands `sum` back (one kernel; same autograd nodes as add + layer_norm).
TensorPtr add_layer_norm(const TensorPtr& x, const TensorPtr& a, const TensorPtr& weight, const TensorPtr& bias, float eps, TensorPtr* sum_out,
                         bool y_bf16_only = false);
struct LinearFusion {
    TensorPtr* gelu_act = nullptr;
    TensorPtr residual;
    TensorPtr gelu_input_pre;
    // bf16 operand mode only: the caller knows that y (or the GELU activation) is consumed as a bf16 GEMM / attention operand,
    // so the epilogue may skip the fp32 copy (Tensor::bf16_only_out — any fp32 reader still gets the values, via an upcast).
    bool out_bf16_only = false, gelu_bf16_only = false;
};
TensorPtr fused_lora_linear(const TensorPtr& x, const TensorPtr& W, const TensorPtr& bias, const std::vector<LoraTerm>& terms,
                            const LinearFusion& fusion);
// logits = x @ E^T for an embedding table E [V,C] (tied lm_head) without materialising E^T
// (graph/gpt2_model.cpp:425-439 transposes wte on every forward).
TensorPtr lm_head_tied(const TensorPtr& x, const TensorPtr& table);
// GPT-2 attention over the fused qkv projection [B,S,3C] -> ctx [B,S,C]; forward only, result carries no
// autograd history exactly like the reference's zeros+memcpy split (graph/gpt2_model.cpp:616-676).
TensorPtr gpt2_causal_attention(const TensorPtr& qkv, int n_head);
// Same attention with a backward (true-gradient mode, MFT_TRUE_ATTN_GRAD=1): one node from the fused qkv to ctx.
TensorPtr gpt2_causal_attention_with_grad(const TensorPtr& qkv, int n_head);
// Fused Gemma attention pieces (graph/gemma_model.cpp:383-514) on the projections' own [B,S,heads*Hd] layout:
// per-head RMSNorm (1+w) + RoPE for q and k in one launch (norm weights frozen), and grouped-query attention
// (causal; window > 0 = sliding; key_bias = additive [B,S] padding mask or null) with a real backward.
struct GqaLayout { bool fused_qkv = false; int64_t q_off = 0, k_off = 0, v_off = 0; };
bool gqa_head_dim_supported(int64_t Hd);
std::pair<TensorPtr, TensorPtr> qk_norm_rope(const TensorPtr& q, const TensorPtr& k, const TensorPtr& wq, const TensorPtr& wk, int H, int KV, int Hd,
                                             float eps, float theta, bool do_norm);
TensorPtr gqa_attention(const TensorPtr& q, const TensorPtr& k, const TensorPtr& v, const TensorPtr& key_bias, int H, int KV, int Hd, float scale,
                        int window);
// x[b,s,:] = wte[ids[b,s],:] + wpe[s,:]   (wpe == nullptr: Gemma, times `scale`)
TensorPtr embedding_lookup(const TensorPtr& ids, const TensorPtr& wte, const TensorPtr& wpe, float scale = 1.0f);

}  // namespace ops
