#include "lora_linear.h"

#include <stdexcept>

#include "mft_b200.h"
#include "ops.h"
#include "ops_internal.h"
#include "runtime.h"

namespace ops {

using internal::GemmCall;

const TensorPtr& LoRALinear::base_W() const {
    static const TensorPtr none;
    return (W_ref_ && *W_ref_) ? *W_ref_ : none;
}
const TensorPtr& LoRALinear::base_b() const { return (b_ref_ && *b_ref_) ? *b_ref_ : b_cache_; }

// nn/lora_linear.cpp:19-40
void LoRALinear::attach_lora(const TensorPtr& A, const TensorPtr& B, float scale, int col0, int cols) {
    if (!A || !B) throw std::runtime_error("LoRALinear::attach_lora: A and B must not be null");
    if (A->ndim() != 2 || B->ndim() != 2) throw std::runtime_error("LoRALinear::attach_lora: A and B must be 2D");
    const int width = cols <= 0 ? int(B->shape()[1]) : cols;
    A->set_requires_grad(true);
    B->set_requires_grad(true);
    slices_.emplace_back(A, B, scale, col0, width);
}
void LoRALinear::clear_lora() { slices_.clear(); merged_ = false; }

TensorPtr LoRALinear::forward(const TensorPtr& x) const { return forward_impl(x, LinearFusion{}); }
TensorPtr LoRALinear::forward_gelu(const TensorPtr& x, TensorPtr* pre) const {
    TensorPtr act;
    LinearFusion f;
    f.gelu_act = &act;
    TensorPtr y = forward_impl(x, f);
    if (pre) *pre = y;
    return act;
}
TensorPtr LoRALinear::forward_fused(const TensorPtr& x, const LinearFusion& fusion) const { return forward_impl(x, fusion); }

// nn/lora_linear.cpp:47-108
TensorPtr LoRALinear::forward_impl(const TensorPtr& x, const LinearFusion& fusion) const {
    TensorPtr* gelu_act = fusion.gelu_act;
    const TensorPtr& W = base_W();
    if (!W) throw std::runtime_error("LoRALinear::forward: base weight is null (sharded and not loaded?)");
    const TensorPtr& bias = base_b();
    const int64_t out = W->shape()[1];
    std::vector<LoraTerm> full;                    // slices spanning every output column: fused + tracked
    std::vector<const LoRASlice*> partial;         // column sub-ranges: added in place, invisible to autograd (:85-102)
    if (!merged_) {
        for (const auto& s : slices_) {
            if (!s.A->requires_grad()) s.A->set_requires_grad(true);
            if (!s.B->requires_grad()) s.B->set_requires_grad(true);
            if (s.cols == out && s.col0 == 0) full.push_back(LoraTerm{s.A, s.B, s.scale});
            else partial.push_back(&s);
        }
    }
    LinearFusion f = fusion;
    if (!partial.empty()) f.gelu_act = nullptr;    // the activation is taken after the in-place slice updates below
    TensorPtr y = fused_lora_linear(x, W, bias, full, f);
    if (!partial.empty()) {
        TensorPtr xc = contiguous(x);
        const int64_t in = W->shape()[0], T = xc->numel() / in;
        for (const LoRASlice* s : partial) {
            const bool a_in_r = s->A->shape()[0] == in;
            const int64_t r = a_in_r ? s->A->shape()[1] : s->A->shape()[0];
            const bool b_r_out = s->B->shape()[0] == r && s->B->shape()[1] == s->cols;
            auto xa = Tensor::make({T, r}, kFloat32);
            GemmCall g;
            g.m = T; g.n = r; g.k = in; g.alpha = s->scale;
            g.a = xc->dev_f32(); g.lda = in;
            g.b = s->A->dev_f32(); g.ldb = s->A->ld(); g.trans_b = a_in_r ? 0 : 1;
            g.c = xa->dev_f32_out(); g.ldc = r;
            internal::run_gemm(g);
            GemmCall h;                            // y[:, col0:col0+cols] += xa B
            h.m = T; h.n = s->cols; h.k = r; h.beta = 1.0f;
            h.a = xa->dev_f32(); h.lda = r;
            h.b = s->B->dev_f32(); h.ldb = s->B->ld(); h.trans_b = b_r_out ? 0 : 1;
            h.c = y->dev_f32_mut() + s->col0; h.ldc = y->ld();
            internal::run_gemm(h);
        }
        if (gelu_act) *gelu_act = gelu(y);
    }
    return y;
}

// W[:, col0:col0+cols] += sign * scale * A B   (nn/lora_linear.cpp:110-178)
void LoRALinear::add_delta_to_base(float sign) {
    const TensorPtr& W = base_W();
    if (!W) throw std::runtime_error("LoRALinear::merge_to_base: base weight is null");
    const int64_t in = W->shape()[0], out = W->shape()[1];
    for (const auto& s : slices_) {
        const bool a_in_r = s.A->shape()[0] == in;
        const int64_t r = a_in_r ? s.A->shape()[1] : s.A->shape()[0];
        const bool b_r_out = s.B->shape()[0] == r;
        const int64_t width = b_r_out ? s.B->shape()[1] : s.B->shape()[0];
        if (width != s.cols || s.col0 + s.cols > out) throw std::runtime_error("LoRALinear::merge_to_base: shape mismatch");
        GemmCall g;
        g.m = in; g.n = s.cols; g.k = r; g.alpha = sign * s.scale; g.beta = 1.0f;
        g.a = s.A->dev_f32(); g.lda = s.A->ld(); g.trans_a = a_in_r ? 0 : 1;
        g.b = s.B->dev_f32(); g.ldb = s.B->ld(); g.trans_b = b_r_out ? 0 : 1;
        g.c = W->dev_f32_mut() + s.col0; g.ldc = W->ld();
        internal::run_gemm(g);
    }
}
void LoRALinear::merge_to_base() {
    if (merged_ || slices_.empty()) return;
    add_delta_to_base(+1.0f);
    merged_ = true;
}
void LoRALinear::unmerge_from_base() {
    if (!merged_ || slices_.empty()) return;
    add_delta_to_base(-1.0f);
    merged_ = false;
}

std::vector<TensorPtr> LoRALinear::trainable_parameters() const {
    std::vector<TensorPtr> p;
    for (const auto& s : slices_) { p.push_back(s.A); p.push_back(s.B); }
    return p;
}
std::vector<std::pair<std::string, TensorPtr>> LoRALinear::debug_params() const {
    std::vector<std::pair<std::string, TensorPtr>> r;
    for (size_t i = 0; i < slices_.size(); ++i) {
        const std::string prefix = debug_name_.empty() ? ("lora_" + std::to_string(i)) : debug_name_;
        r.emplace_back(prefix + "_lora_A_default_weight", slices_[i].A);
        r.emplace_back(prefix + "_lora_B_default_weight", slices_[i].B);
    }
    return r;
}

}  // namespace ops
