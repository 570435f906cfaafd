// Remaining API surface of core/ops.h (comparisons, pow/clamp, masks, losses that the trainers do not call,
// cast, dropout, batch_norm) plus the model-level device helpers (embedding gather, GPT-2 attention).
#include <algorithm>
#include <cmath>
#include <cstdlib>
#include <cstring>
#include <random>
#include <stdexcept>

#include "autograd_engine.h"
#include "mft_b200.h"
#include "ops.h"
#include "ops_internal.h"
#include "runtime.h"

namespace ops {

using namespace internal;

namespace {
void* st() { return mft_rt::stream(); }

struct NoGrad {
    bool was;
    NoGrad() : was(autograd::Engine::instance().is_enabled()) { autograd::Engine::instance().set_enabled(false); }
    ~NoGrad() { autograd::Engine::instance().set_enabled(was); }
};

TensorPtr compare(int op, const TensorPtr& a_in, const TensorPtr& b_in, const char* name) {   // core/ops.cpp:1941-1992
    require_f32(a_in, name);
    require_f32(b_in, name);
    TensorPtr a = contiguous(a_in), b = contiguous(b_in);
    if (a->shape() != b->shape()) throw TensorError(std::string(name) + ": shapes must match");
    auto out = new_like(a);
    mft_rt::check(mft_binary(op, a->dev_f32(), b->dev_f32(), out->dev_f32_out(), 1, 1, a->numel(), 0, 1, st()), name);
    return out;
}

class PowBackward : public BackwardFunction {
public:
    PowBackward(TensorPtr x, float e) : x_(std::move(x)), e_(e) {}
    std::vector<TensorPtr> apply(const TensorPtr& gy) override {
        NoGrad ng;
        return {mul(gy, mul(pow(x_, e_ - 1.0f), e_))};
    }
private:
    TensorPtr x_;
    float e_;
};
class ClampBackward : public BackwardFunction {
public:
    ClampBackward(TensorPtr x, float lo, float hi) : x_(std::move(x)), lo_(lo), hi_(hi) {}
    std::vector<TensorPtr> apply(const TensorPtr& gy) override {
        NoGrad ng;
        auto inside = mul(ge(x_, full(x_->shape(), lo_)), le(x_, full(x_->shape(), hi_)));
        return {mul(gy, inside)};
    }
private:
    TensorPtr x_;
    float lo_, hi_;
};
}  // namespace

TensorPtr eq(const TensorPtr& a, const TensorPtr& b) { return compare(4, a, b, "eq"); }
TensorPtr ne(const TensorPtr& a, const TensorPtr& b) { return compare(5, a, b, "ne"); }
TensorPtr gt(const TensorPtr& a, const TensorPtr& b) { return compare(6, a, b, "gt"); }
TensorPtr lt(const TensorPtr& a, const TensorPtr& b) { return compare(7, a, b, "lt"); }
TensorPtr ge(const TensorPtr& a, const TensorPtr& b) { return compare(8, a, b, "ge"); }
TensorPtr le(const TensorPtr& a, const TensorPtr& b) { return compare(9, a, b, "le"); }

TensorPtr pow(const TensorPtr& t_in, float exponent) {
    require_f32(t_in, "pow");
    TensorPtr t = contiguous(t_in);
    auto out = new_like(t);
    mft_rt::check(mft_unary_param(0, t->dev_f32(), out->dev_f32_out(), t->numel(), exponent, 0.f, st()), "pow");
    track(out, {t_in}, std::make_shared<PowBackward>(t, exponent));
    return out;
}
TensorPtr clamp(const TensorPtr& t_in, float lo, float hi) {
    require_f32(t_in, "clamp");
    TensorPtr t = contiguous(t_in);
    auto out = new_like(t);
    mft_rt::check(mft_unary_param(1, t->dev_f32(), out->dev_f32_out(), t->numel(), lo, hi, st()), "clamp");
    track(out, {t_in}, std::make_shared<ClampBackward>(t, lo, hi));
    return out;
}

// core/tensor.cpp / core/ops.cpp cast: fp32 <-> int32 (truncation), one device kernel;
// fp16 storage is not used by the fp32 training path and is rejected loudly.
TensorPtr cast(const TensorPtr& t, DType target) {
    if (!t) throw TensorError("cast: null tensor");
    if (t->dtype() == target) return t->clone();
    if (target == kFloat16 || t->dtype() == kFloat16) throw TensorError("cast: float16 tensors are not supported by the B200 fp32/TF32 path");
    const bool to_i32 = t->dtype() == kFloat32 && target == kInt32;
    if (!to_i32 && !(t->dtype() == kInt32 && target == kFloat32)) throw TensorError("cast: unsupported dtype pair");
    TensorPtr src = t->dtype() == kFloat32 ? contiguous(t) : t;
    auto out = Tensor::make(t->shape(), target);
    mft_rt::check(mft_cast_i32(src->dev(), out->dev_out(), t->numel(), to_i32 ? 1 : 0, st()), "mft_cast_i32");
    return out;
}

TensorPtr uniform(const std::vector<int64_t>& shape, float low, float high, DType dtype, Device device) {
    auto t = zeros(shape, dtype, device);
    static std::mt19937 g{std::random_device{}()};
    std::uniform_real_distribution<float> d(low, high);
    float* p = t->data<float>();
    for (int64_t i = 0; i < t->numel(); ++i) p[i] = d(g);
    return t;
}

// core/ops.cpp dropout: inverted dropout with a host-generated keep mask (LoRA dropout is never applied by
// LoRALinear::forward in the reference — SURVEY.md Appendix B10 — so this is off the training path). The uniform draws
// are initialisation data produced by the host RNG like every other random fill; the thresholding runs on the device.
TensorPtr dropout(const TensorPtr& t, float p, bool training) {
    if (!training || p <= 0.0f) return t;
    if (p >= 1.0f) return mul(t, 0.0f);
    TensorPtr keep_mask;
    {
        NoGrad ng;             // the mask is data: u >= p ? 1/(1-p) : 0, thresholded on the device
        keep_mask = mul(ge(uniform(t->shape(), 0.f, 1.f), full(t->shape(), p)), 1.0f / (1.0f - p));
    }
    return mul(t, keep_mask);
}

// core/ops.cpp:1976-1992 / 1994-2070
TensorPtr create_causal_mask(int seq_len, DType dtype, Device device) {
    (void)device;
    if (dtype != kFloat32) throw TensorError("create_causal_mask: float32 only");
    auto mask = Tensor::make({seq_len, seq_len}, kFloat32);
    mft_rt::check(mft_tril_ones(mask->dev_f32_out(), seq_len, st()), "mft_tril_ones");
    return mask;
}
TensorPtr apply_mask(const TensorPtr& input, const TensorPtr& mask, float mask_value) {
    // keep-mask semantics: out = input where mask != 0, mask_value elsewhere
    NoGrad ng;
    auto keep = ne(add(mul(input, 0.0f), mask), zeros(input->shape()));
    return add(mul(input, keep), mul(sub(1.0f, keep), mask_value));
}

// ---- losses the trainers do not call (core/ops.cpp:1177-1402); composed from device ops
TensorPtr mse_loss(const TensorPtr& input, const TensorPtr& target, const std::string& reduction) {
    auto d = sub(input, target);
    auto sq = mul(d, d);
    if (reduction == "none") return sq;
    return reduction == "sum" ? sum(sq) : mean(sq);
}
static TensorPtr one_hot_from_targets(const TensorPtr& target, int64_t N, int64_t Cn) {
    auto oh = Tensor::make({N, Cn}, kFloat32);
    mft_rt::check(mft_one_hot(target->dev(), target->dtype() == kInt32 ? 1 : 0, oh->dev_f32_out(), N, Cn, st()), "mft_one_hot");
    return oh;
}
TensorPtr nll_loss(const TensorPtr& input, const TensorPtr& target, const std::string& reduction) {
    if (input->ndim() != 2) throw TensorError("nll_loss: input must be [N, C]");
    const int64_t N = input->shape()[0], Cn = input->shape()[1];
    auto picked = mul(sum(mul(input, one_hot_from_targets(target, N, Cn)), 1), -1.0f);
    if (reduction == "none") return picked;
    return reduction == "sum" ? sum(picked) : mean(picked);
}
TensorPtr cross_entropy_loss(const TensorPtr& input, const TensorPtr& target, const std::string& reduction) {
    return nll_loss(log_softmax(input, -1), target, reduction);
}

// core/ops.cpp:1592-1700 — batch statistics over dim 0 of [N, C] (training) or the running statistics (eval)
TensorPtr batch_norm(const TensorPtr& input, const TensorPtr& weight, const TensorPtr& bias, const TensorPtr& running_mean,
                     const TensorPtr& running_var, bool training, float momentum, float eps) {
    (void)momentum;
    if (input->ndim() != 2) throw TensorError("batch_norm: [N, C] input expected");
    TensorPtr mu, var;
    if (training || !running_mean || !running_var) {
        mu = mean(input, 0, true);
        auto d = sub(input, mu);
        var = mean(mul(d, d), 0, true);
    } else { mu = running_mean->reshape({1, input->shape()[1]}); var = running_var->reshape({1, input->shape()[1]}); }
    auto xhat = div(sub(input, mu), sqrt(add(var, eps)));
    if (weight) xhat = mul(xhat, weight);
    if (bias) xhat = add(xhat, bias);
    return xhat;
}

// ---------------------------------------------------------------------------- model-level helpers
TensorPtr embedding_lookup(const TensorPtr& ids, const TensorPtr& wte, const TensorPtr& wpe, float scale) {
    if (!ids || ids->dtype() != kInt32) throw std::runtime_error("embedding_lookup expects int32 input_ids");
    if (ids->ndim() != 2) throw std::runtime_error("embedding_lookup expects [batch, seq_len]");
    const int64_t B = ids->shape()[0], S = ids->shape()[1], C = wte->shape()[1], V = wte->shape()[0];
    auto x = Tensor::make({B, S, C}, kFloat32);
    mft_rt::check(mft_embedding_fwd(static_cast<const int32_t*>(ids->dev()), wte->dev_f32(), wpe ? wpe->dev_f32() : nullptr,
                                    x->dev_f32_out(), B * S, S, C, V, scale, nullptr, st()), "mft_embedding_fwd");
    return x;
}

TensorPtr gpt2_causal_attention(const TensorPtr& qkv_in, int n_head) {
    require_f32(qkv_in, "gpt2_causal_attention");
    TensorPtr qkv = contiguous(qkv_in);
    const auto& s = qkv->shape();
    if (s.size() != 3 || s[2] % 3) throw TensorError("gpt2_causal_attention: qkv must be [B, S, 3C]");
    const int64_t B = s[0], S = s[1], C = s[2] / 3, Hd = C / n_head;
    auto ctx = Tensor::make({B, S, C}, kFloat32);
    float* cp = ctx->dev_f32_out();
    void* c16 = nullptr;
    if (internal::emit_bf16_operands() && C % 8 == 0) c16 = ctx->bf16_emit().ptr;     // ctx is the A operand of the output projection
    // bf16 operand mode, head_dim 64: tcgen05 kernel fed by TMA from the projection's bf16 copy (MFT_ATTENTION_TC=0: mma.sync kernel)
    static const bool tc_on = [] { const char* e = std::getenv("MFT_ATTENTION_TC"); return !(e && e[0] == '0'); }();
    if (tc_on && Hd == 64 && internal::gemm_precision() == 1 && !(mft_attention_mode() & 1) && (3 * C) % 8 == 0) {
        auto v16 = qkv->bf16_view();
        mft_rt::check(mft_causal_attention_tc_fwd(v16.ptr, v16.ld, cp, c16, int32_t(B), n_head, int32_t(S), int32_t(Hd), 1.0f / std::sqrt(float(Hd)), st()),
                      "mft_causal_attention_tc_fwd");
        return ctx;
    }
    const float* base = qkv->dev_f32();
    mft_rt::check(mft_causal_attention_fwd_ex(base, base + C, base + 2 * C, cp, c16, int32_t(B), n_head, int32_t(S), int32_t(Hd),
                                              S * 3 * C, 3 * C, Hd, S * 3 * C, 3 * C, Hd, S * C, C, Hd,
                                              1.0f / std::sqrt(float(Hd)), 0, st()), "mft_causal_attention_fwd");
    return ctx;   // no autograd node: the reference's split severs the graph here
}

}  // namespace ops
