// Activations, normalisation, softmax, RoPE, GQA repeat, shape ops, reductions and the remaining API surface.
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

// ---- unary ops with gradients expressed through other device kernels
// bf16 GEMM mode: the tensor's bf16 shadow is written by the producing kernel (call after dev_f32_out()). Returns a null
// pointer when the mode is off or the layout does not allow it (the GEMM then casts on demand as before).
struct Emit { void* ptr = nullptr; int64_t ld = 0; };
Emit emit16(const TensorPtr& t, bool need_dense_pitch) {
    Emit e;
    if (!emit_bf16_operands() || !t || !t->is_dense() || t->numel() == 0) return e;
    const int64_t cols = t->shape().empty() ? 1 : t->shape().back();
    if (need_dense_pitch ? (cols % 8 != 0) : (cols % 4 != 0)) return e;
    auto v = t->bf16_emit();
    e.ptr = v.ptr; e.ld = v.ld;
    return e;
}

// bf16-only result allowed? (bf16 operand mode, vectorisable layout, MFT_BF16_ONLY != 0)
bool lowp_result_ok(const TensorPtr& y, int64_t dim) {
    return bf16_only_results() && y->is_dense() && dim % 8 == 0;
}

class GeluBackward : public BackwardFunction {       // core/backward_functions.cpp:155-177
public:
    explicit GeluBackward(TensorPtr x) : x_(std::move(x)) {}
    std::vector<TensorPtr> apply(const TensorPtr& gy_in) override {
        TensorPtr gy = contiguous(gy_in);
        auto gx = new_like(x_);
        // the gradient arrived as a bf16-only GEMM result and leaves as a bf16 GEMM operand: stay in bf16 on both sides
        // (x = the fp32 pre-activation); any fp32 reader of gx (bias / adapter gradients of fc_in) gets an upcast on demand
        if (gy->f32_pending() && gy->has_bf16() && lowp_result_ok(gx, x_->shape().back()) && x_->numel() % 8 == 0) {
            auto gv = gy->bf16_view();
            if (gv.ld == gy->shape().back()) {
                auto e = gx->bf16_only_out();
                if (x_->f32_pending() && x_->has_bf16() && x_->bf16_view().ld == x_->shape().back())      // the pre-activation itself exists as bf16 only
                    mft_rt::check(mft_gelu_bwd_lowp(x_->bf16_view().ptr, 1, gv.ptr, nullptr, e.ptr, x_->numel(), st()), "mft_gelu_bwd_lowp");
                else
                    mft_rt::check(mft_gelu_bwd_lowp(x_->dev_f32(), 0, gv.ptr, nullptr, e.ptr, x_->numel(), st()), "mft_gelu_bwd_lowp");
                return {gx};
            }
        }
        float* out = gx->dev_f32_out();
        const Emit e = emit16(gx, true);      // consumed by fc_in's dX product
        if (e.ptr) mft_rt::check(mft_gelu_bwd_ex(x_->dev_f32(), gy->dev_f32(), out, e.ptr, x_->numel(), st()), "mft_gelu_bwd_ex");
        else mft_rt::check(mft_gelu_bwd(x_->dev_f32(), gy->dev_f32(), out, x_->numel(), st()), "mft_gelu_bwd");
        return {gx};
    }
private:
    TensorPtr x_;
};

// d/dx for the remaining activations, composed from elementwise kernels (none is on the training path).
class UnaryBackward : public BackwardFunction {
public:
    UnaryBackward(int kind, TensorPtr x, TensorPtr y) : kind_(kind), x_(std::move(x)), y_(std::move(y)) {}
    std::vector<TensorPtr> apply(const TensorPtr& gy) override {
        autograd::Engine& e = autograd::Engine::instance();
        const bool was = e.is_enabled();
        e.set_enabled(false);
        TensorPtr g;
        switch (kind_) {
            case 1: g = mul(gy, gt(x_, zeros(x_->shape()))); break;                           // relu
            case 2: g = mul(gy, mul(y_, sub(1.0f, y_))); break;                               // sigmoid
            case 3: { auto s = sigmoid(x_); g = mul(gy, mul(s, add(1.0f, mul(x_, sub(1.0f, s))))); break; }  // silu
            case 4: g = mul(gy, sub(1.0f, mul(y_, y_))); break;                               // tanh
            case 5: g = mul(gy, y_); break;                                                   // exp
            case 6: g = div(gy, x_); break;                                                   // log
            case 7: g = div(gy, mul(y_, 2.0f)); break;                                        // sqrt
            default: g = mul(gy, sub(mul(gt(x_, zeros(x_->shape())), 2.0f), 1.0f)); break;    // abs -> sign
        }
        e.set_enabled(was);
        return {g};
    }
private:
    int kind_;
    TensorPtr x_, y_;
};

TensorPtr unary(int kind, const TensorPtr& x_in, const char* name) {
    require_f32(x_in, name);
    TensorPtr x = contiguous(x_in);
    auto y = new_like(x);
    mft_rt::check(mft_unary(kind, x->dev_f32(), y->dev_f32_out(), x->numel(), st()), name);
    if (kind == 0) track(y, {x_in}, std::make_shared<GeluBackward>(x));
    else track(y, {x_in}, std::make_shared<UnaryBackward>(kind, x, y));
    return y;
}

}  // namespace

namespace internal {
BackwardFunctionPtr make_gelu_backward(const TensorPtr& x) { return std::make_shared<GeluBackward>(x); }
}  // namespace internal

// core/ops.cpp:1033-1117
TensorPtr gelu(const TensorPtr& x) { return unary(0, x, "gelu"); }
TensorPtr relu(const TensorPtr& x) { return unary(1, x, "relu"); }
TensorPtr sigmoid(const TensorPtr& x) { return unary(2, x, "sigmoid"); }
TensorPtr silu(const TensorPtr& x) { return unary(3, x, "silu"); }
TensorPtr tanh_op(const TensorPtr& x) { return unary(4, x, "tanh"); }
TensorPtr exp(const TensorPtr& x) { return unary(5, x, "exp"); }
TensorPtr log(const TensorPtr& x) { return unary(6, x, "log"); }
TensorPtr sqrt(const TensorPtr& x) { return unary(7, x, "sqrt"); }
TensorPtr abs(const TensorPtr& x) { return unary(8, x, "abs"); }
namespace {
class GegluBackward : public BackwardFunction {
public:
    GegluBackward(TensorPtr gate, TensorPtr up) : gate_(std::move(gate)), up_(std::move(up)) {}
    std::vector<TensorPtr> apply(const TensorPtr& gy_in) override {
        TensorPtr gy = contiguous(gy_in);
        auto dg = new_like(gate_), du = new_like(up_);
        float* dgp = dg->dev_f32_out();
        float* dup = du->dev_f32_out();
        const Emit eg = emit16(dg, true), eu = emit16(du, true);      // both feed the dX products of gate / up
        mft_rt::check(mft_geglu_bwd(gate_->dev_f32(), up_->dev_f32(), gy->dev_f32(), dgp, dup, eg.ptr, eu.ptr, gate_->numel(), st()), "mft_geglu_bwd");
        return {dg, du};
    }
private:
    TensorPtr gate_, up_;
};
}  // namespace

// gelu(gate) * up as one node (B200 extension; the Gemma MLP, graph/gemma_model.cpp:534-551)
TensorPtr geglu(const TensorPtr& gate_in, const TensorPtr& up_in) {
    require_f32(gate_in, "geglu");
    require_f32(up_in, "geglu");
    if (gate_in->shape() != up_in->shape() || gate_in->numel() % 4) return mul(gelu(gate_in), up_in);
    TensorPtr gate = contiguous(gate_in), up = contiguous(up_in);
    auto y = new_like(gate);
    float* yp = y->dev_f32_out();
    const Emit e = emit16(y, true);                                   // the down projection's A operand
    mft_rt::check(mft_geglu_fwd(gate->dev_f32(), up->dev_f32(), yp, e.ptr, gate->numel(), st()), "mft_geglu_fwd");
    track(y, {gate_in, up_in}, std::make_shared<GegluBackward>(gate, up));
    return y;
}

TensorPtr swiglu(const TensorPtr& gate, const TensorPtr& up) { return mul(silu(gate), up); }   // core/ops.cpp:2227-2264

// ---------------------------------------------------------------------------- normalisation
namespace {
class LayerNormBackward : public BackwardFunction {  // core/backward_functions.cpp:425-481
public:
    LayerNormBackward(TensorPtr x, TensorPtr w, TensorPtr b, float eps) : x_(std::move(x)), w_(std::move(w)), b_(std::move(b)), eps_(eps) {}
    std::vector<TensorPtr> run(const TensorPtr& gy_in, const TensorPtr& waiting) {
        TensorPtr gy = contiguous(gy_in);
        const int64_t dim = x_->shape().back(), rows = x_->numel() / dim;
        auto gx = new_like(x_);
        // the reference always computes dgamma/dbeta (core/backward_functions.cpp:443-470); nobody reads them when both
        // parameters are frozen (LoRA), so the column reductions are skipped then
        const bool wg = wants_grad(w_) || wants_grad(b_);
        TensorPtr gw, gb;
        if (wg) { gw = Tensor::make(w_->shape(), kFloat32); gb = Tensor::make(w_->shape(), kFloat32); }
        size_t wsb = wg ? mft_norm_bwd_workspace_bytes(rows, dim) : 0;
        float* gxp = gx->dev_f32_out();
        const Emit e = emit16(gx, false);     // the stream gradient feeds the dX products of the layer below
        if (waiting || e.ptr)
            mft_rt::check(mft_layernorm_bwd_ex(x_->dev_f32(), w_->dev_f32(), gy->dev_f32(), waiting ? waiting->dev_f32() : nullptr, gxp, e.ptr, e.ld,
                                               wg ? gw->dev_f32_out() : nullptr, wg ? gb->dev_f32_out() : nullptr, rows, dim, eps_,
                                               wg ? mft_rt::workspace(wsb) : nullptr, wsb, st()), "mft_layernorm_bwd_ex");
        else
            mft_rt::check(mft_layernorm_bwd(x_->dev_f32(), w_->dev_f32(), gy->dev_f32(), gxp, wg ? gw->dev_f32_out() : nullptr,
                                            wg ? gb->dev_f32_out() : nullptr, rows, dim, eps_, wg ? mft_rt::workspace(wsb) : nullptr, wsb, st()),
                          "mft_layernorm_bwd");
        return {gx, gw, gb};
    }
    std::vector<TensorPtr> apply(const TensorPtr& gy_in) override { return run(gy_in, nullptr); }
    // the gradient waiting at x (the residual branch's) is added inside the kernel that produces dX
    bool wants_pending() const override { return true; }
    std::vector<TensorPtr> apply_fused(const TensorPtr& gy_in, const std::vector<TensorPtr>& pending, std::vector<bool>& absorbed) override {
        TensorPtr w = !pending.empty() ? pending[0] : nullptr;
        if (w && (w->numel() != x_->numel() || !w->is_dense())) w = nullptr;
        if (w && !absorbed.empty()) absorbed[0] = true;
        return run(gy_in, w);
    }
private:
    TensorPtr x_, w_, b_;
    float eps_;
};
class RmsNormBackward : public BackwardFunction {    // core/backward_functions.cpp:483-523
public:
    RmsNormBackward(TensorPtr x, TensorPtr w, float eps, int plus_one, bool gx_feeds_gemm)
        : x_(std::move(x)), w_(std::move(w)), eps_(eps), plus_one_(plus_one), gx_feeds_gemm_(gx_feeds_gemm) {}
    std::vector<TensorPtr> apply(const TensorPtr& gy_in) override {
        TensorPtr gy = contiguous(gy_in);
        const int64_t dim = x_->shape().back(), rows = x_->numel() / dim;
        auto gx = new_like(x_);
        // the reference always accumulates dW (core/backward_functions.cpp:483-523); with a frozen weight (LoRA runs) nobody reads
        // it, so the column reduction and its second stage are skipped, like the LayerNorm case above
        const bool wg = wants_grad(w_);
        TensorPtr gw = wg ? Tensor::make(w_->shape(), kFloat32) : nullptr;
        size_t wsb = wg ? mft_norm_bwd_workspace_bytes(rows, dim) : 0;
        float* gxp = gx->dev_f32_out();
        const Emit e = gx_feeds_gemm_ ? emit16(gx, false) : Emit{};      // a post-norm's gradient is the dX operand of the projection behind it
        mft_rt::check(mft_rmsnorm_bwd_ex(x_->dev_f32(), w_->dev_f32(), gy->dev_f32(), gxp, e.ptr, e.ld, wg ? gw->dev_f32_out() : nullptr, rows, dim,
                                         eps_, plus_one_, wg ? mft_rt::workspace(wsb) : nullptr, wsb, st()), "mft_rmsnorm_bwd");
        return {gx, gw};
    }
private:
    TensorPtr x_, w_;
    float eps_;
    int plus_one_;
    bool gx_feeds_gemm_;
};
}  // namespace

// core/ops.cpp:1404-1458
TensorPtr layer_norm(const TensorPtr& x_in, const TensorPtr& weight, const TensorPtr& bias, float eps) {
    return layer_norm_lowp(x_in, weight, bias, eps, false);
}

TensorPtr layer_norm_lowp(const TensorPtr& x_in, const TensorPtr& weight, const TensorPtr& bias, float eps, bool y_bf16_only) {
    require_f32(x_in, "layer_norm");
    if (!weight || !bias) throw TensorError("layer_norm: weight and bias are required");
    TensorPtr x = contiguous(x_in);
    const int64_t dim = x->shape().back(), rows = x->numel() / std::max<int64_t>(dim, 1);
    if (weight->numel() != dim || bias->numel() != dim) throw TensorError("layer_norm: parameter size mismatch");
    auto y = new_like(x);
    if (y_bf16_only && dim <= 2048 && lowp_result_ok(y, dim)) {          // the normalised rows are read as a bf16 GEMM operand only
        auto e = y->bf16_only_out();
        mft_rt::check(mft_layernorm_fwd_ex(x->dev_f32(), nullptr, nullptr, weight->dev_f32(), bias->dev_f32(), nullptr, e.ptr, e.ld, rows, dim, eps, st()),
                      "mft_layernorm_fwd_ex");
    } else {
        float* yp = y->dev_f32_out();
        const Emit e = emit16(y, false);          // LayerNorm outputs are GEMM inputs (qkv, fc_in, lm_head)
        if (e.ptr) mft_rt::check(mft_layernorm_fwd_ex(x->dev_f32(), nullptr, nullptr, weight->dev_f32(), bias->dev_f32(), yp, e.ptr, e.ld, rows, dim, eps, st()),
                                 "mft_layernorm_fwd_ex");
        else mft_rt::check(mft_layernorm_fwd(x->dev_f32(), weight->dev_f32(), bias->dev_f32(), yp, rows, dim, eps, st()), "mft_layernorm_fwd");
    }
    track(y, {x_in, weight, bias}, std::make_shared<LayerNormBackward>(x, weight, bias, eps));
    return y;
}

namespace {
class ResidualAddBackward : public BackwardFunction {   // AddBackward for equal shapes (core/backward_functions.cpp:36-72)
public:
    std::vector<TensorPtr> apply(const TensorPtr& gy) override { return {gy, gy}; }
};
}  // namespace

// sum = x + a and y = layer_norm(sum) from one kernel; both results carry the same autograd nodes as ops::add followed by
// ops::layer_norm (graph/gpt2_model.cpp block forward: residual add, then ln_2 / the next block's ln_1 / ln_f).
TensorPtr add_layer_norm(const TensorPtr& x_in, const TensorPtr& a_in, const TensorPtr& weight, const TensorPtr& bias, float eps, TensorPtr* sum_out,
                         bool y_bf16_only) {
    require_f32(x_in, "add_layer_norm");
    require_f32(a_in, "add_layer_norm");
    if (!weight || !bias || !sum_out) throw TensorError("add_layer_norm: weight, bias and sum_out are required");
    if (x_in->shape() != a_in->shape()) throw TensorError("add_layer_norm: operands must have the same shape");
    TensorPtr x = contiguous(x_in), a = contiguous(a_in);
    const int64_t dim = x->shape().back(), rows = x->numel() / std::max<int64_t>(dim, 1);
    if (weight->numel() != dim || bias->numel() != dim) throw TensorError("layer_norm: parameter size mismatch");
    auto sum = new_like(x);
    auto y = new_like(x);
    if (y_bf16_only && dim <= 2048 && lowp_result_ok(y, dim)) {
        auto e = y->bf16_only_out();
        mft_rt::check(mft_layernorm_fwd_ex(x->dev_f32(), a->dev_f32(), sum->dev_f32_out(), weight->dev_f32(), bias->dev_f32(), nullptr, e.ptr, e.ld, rows,
                                           dim, eps, st()), "mft_layernorm_fwd_ex");
    } else {
        float* yp = y->dev_f32_out();
        const Emit e = emit16(y, false);
        mft_rt::check(mft_layernorm_fwd_ex(x->dev_f32(), a->dev_f32(), sum->dev_f32_out(), weight->dev_f32(), bias->dev_f32(), yp, e.ptr, e.ld, rows,
                                           dim, eps, st()), "mft_layernorm_fwd_ex");
    }
    track(sum, {x_in, a_in}, std::make_shared<ResidualAddBackward>());
    track(y, {sum, weight, bias}, std::make_shared<LayerNormBackward>(sum, weight, bias, eps));
    *sum_out = sum;
    return y;
}

static TensorPtr rms_impl(const TensorPtr& x_in, const TensorPtr& weight, float eps, int plus_one, bool gx_feeds_gemm = false) {
    require_f32(x_in, "rms_norm");
    if (!weight) throw TensorError("rms_norm: weight is required");
    TensorPtr x = contiguous(x_in);
    const int64_t dim = x->shape().back(), rows = x->numel() / std::max<int64_t>(dim, 1);
    if (weight->numel() != dim) throw TensorError("rms_norm: weight size mismatch");
    auto y = new_like(x);
    float* yp = y->dev_f32_out();
    const Emit e = emit16(y, false);          // normalised activations feed the q/k/v, gate/up and lm_head products
    if (e.ptr) mft_rt::check(mft_rmsnorm_fwd_ex(x->dev_f32(), weight->dev_f32(), yp, e.ptr, e.ld, rows, dim, eps, plus_one, st()), "mft_rmsnorm_fwd_ex");
    else mft_rt::check(mft_rmsnorm_fwd(x->dev_f32(), weight->dev_f32(), yp, rows, dim, eps, plus_one, st()), "mft_rmsnorm_fwd");
    // the reference always registers a node for rms_norm (core/ops.cpp:1521-1524)
    if (autograd::Engine::instance().is_enabled()) {
        y->set_requires_grad(true);
        autograd::Engine::instance().register_node(y, {x_in, weight}, std::make_shared<RmsNormBackward>(x, weight, eps, plus_one, gx_feeds_gemm));
    }
    return y;
}
TensorPtr rms_norm(const TensorPtr& x, const TensorPtr& w, float eps) { return rms_impl(x, w, eps, 1); }          // core/ops.cpp:1489-1537
// the same op for a norm applied to a projection's output (Gemma's post-attention / post-feedforward norms): its backward also
// writes the gradient as the bf16 operand of that projection's dX product
TensorPtr rms_norm_post(const TensorPtr& x, const TensorPtr& w, float eps) { return rms_impl(x, w, eps, 1, true); }
TensorPtr rms_norm_affine(const TensorPtr& x, const TensorPtr& w, float eps) { return rms_impl(x, w, eps, 0); }  // core/ops.cpp:1539-1590

// ---------------------------------------------------------------------------- softmax
namespace {
class SoftmaxBackward : public BackwardFunction {    // core/backward_functions.cpp:297-340
public:
    explicit SoftmaxBackward(TensorPtr y) : y_(std::move(y)) {}
    std::vector<TensorPtr> apply(const TensorPtr& gy_in) override {
        TensorPtr gy = contiguous(gy_in);
        const int64_t dim = y_->shape().back(), rows = y_->numel() / dim;
        auto gx = new_like(y_);
        mft_rt::check(mft_softmax_bwd(y_->dev_f32(), gy->dev_f32(), gx->dev_f32_out(), rows, dim, st()), "mft_softmax_bwd");
        return {gx};
    }
private:
    TensorPtr y_;
};
class LogSoftmaxBackward : public BackwardFunction { // gx = gy - softmax * sum(gy)
public:
    explicit LogSoftmaxBackward(TensorPtr y) : y_(std::move(y)) {}
    std::vector<TensorPtr> apply(const TensorPtr& gy) override {
        autograd::Engine& e = autograd::Engine::instance();
        const bool was = e.is_enabled();
        e.set_enabled(false);
        auto p = exp(y_);
        const int64_t dim = y_->shape().back(), rows = y_->numel() / dim;
        auto s = reshape(sum(reshape(gy, {rows, dim}), 1, true), {rows, 1});
        auto g = sub(reshape(gy, {rows, dim}), mul(reshape(p, {rows, dim}), matmul(s, ones({1, dim}))));
        e.set_enabled(was);
        return {reshape(g, y_->shape())};
    }
private:
    TensorPtr y_;
};
}  // namespace

// core/ops.cpp:1119-1175 (last-dim softmax; other dims go through a transpose)
TensorPtr softmax(const TensorPtr& x_in, int dim) {
    require_f32(x_in, "softmax");
    const int nd = int(x_in->ndim());
    if (dim < 0) dim += nd;
    if (dim != nd - 1) return transpose(softmax(transpose(x_in, dim, nd - 1), -1), dim, nd - 1);
    TensorPtr x = contiguous(x_in);
    const int64_t d = x->shape().back(), rows = x->numel() / std::max<int64_t>(d, 1);
    auto y = new_like(x);
    mft_rt::check(mft_softmax_fwd(x->dev_f32(), y->dev_f32_out(), rows, d, st()), "mft_softmax_fwd");
    track(y, {x_in}, std::make_shared<SoftmaxBackward>(y));
    return y;
}
TensorPtr log_softmax(const TensorPtr& x_in, int dim) {
    autograd::Engine& e = autograd::Engine::instance();
    const bool was = e.is_enabled();
    e.set_enabled(false);
    auto y = log(softmax(x_in, dim));
    e.set_enabled(was);
    track(y, {x_in}, std::make_shared<LogSoftmaxBackward>(y));
    return y;
}

// ---------------------------------------------------------------------------- RoPE / GQA repeat
namespace {
class RopeBackward : public BackwardFunction {       // core/backward_functions.cpp:718-763
public:
    RopeBackward(int64_t bh, int64_t s, int64_t hd, float theta) : bh_(bh), s_(s), hd_(hd), theta_(theta) {}
    std::vector<TensorPtr> apply(const TensorPtr& gy_in) override {
        TensorPtr gy = contiguous(gy_in);
        auto gx = new_like(gy);
        mft_rt::check(mft_rope(gy->dev_f32(), gx->dev_f32_out(), bh_, s_, hd_, theta_, 1, st()), "mft_rope");
        return {gx};
    }
private:
    int64_t bh_, s_, hd_;
    float theta_;
};
class RepeatKvBackward : public BackwardFunction {   // core/backward_functions.cpp:683-716
public:
    RepeatKvBackward(std::vector<int64_t> in_shape, int rep) : s_(std::move(in_shape)), rep_(rep) {}
    std::vector<TensorPtr> apply(const TensorPtr& gy_in) override {
        TensorPtr gy = contiguous(gy_in);
        auto gx = Tensor::make(s_, kFloat32);
        mft_rt::check(mft_repeat_kv_bwd(gy->dev_f32(), gx->dev_f32_out(), s_[0], s_[1], rep_, s_[2], s_[3], st()), "mft_repeat_kv_bwd");
        return {gx};
    }
private:
    std::vector<int64_t> s_;
    int rep_;
};
}  // namespace

// core/ops.cpp:2151-2225
TensorPtr apply_rope(const TensorPtr& x_in, int seq_len, int head_dim, float rope_theta) {
    require_f32(x_in, "apply_rope");
    if (x_in->ndim() != 4) throw std::runtime_error("apply_rope expects 4D tensor: [batch, heads, seq_len, head_dim]");
    TensorPtr x = contiguous(x_in);
    const auto& s = x->shape();
    if (s[3] != head_dim || s[2] != seq_len) throw std::runtime_error("RoPE dimension mismatch");
    auto y = new_like(x);
    mft_rt::check(mft_rope(x->dev_f32(), y->dev_f32_out(), s[0] * s[1], s[2], s[3], rope_theta, 0, st()), "mft_rope");
    track(y, {x_in}, std::make_shared<RopeBackward>(s[0] * s[1], s[2], s[3], rope_theta));
    return y;
}

// core/ops.cpp:2072-2149
TensorPtr repeat_kv_heads(const TensorPtr& kv_in, int repeat_factor) {
    require_f32(kv_in, "repeat_kv_heads");
    if (kv_in->ndim() != 4) throw std::runtime_error("repeat_kv_heads expects [batch, kv_heads, seq_len, head_dim]");
    if (repeat_factor == 1) return kv_in;
    if (repeat_factor < 1) throw std::runtime_error("repeat_kv_heads: repeat_factor must be >= 1");
    TensorPtr kv = contiguous(kv_in);
    const auto& s = kv->shape();
    auto y = Tensor::make({s[0], s[1] * repeat_factor, s[2], s[3]}, kFloat32);
    mft_rt::check(mft_repeat_kv_fwd(kv->dev_f32(), y->dev_f32_out(), s[0], s[1], repeat_factor, s[2], s[3], st()), "mft_repeat_kv_fwd");
    track(y, {kv_in}, std::make_shared<RepeatKvBackward>(s, repeat_factor));
    return y;
}

// ---------------------------------------------------------------------------- shape ops (aliasing views)
namespace {
class ReshapeBackward : public BackwardFunction {
public:
    explicit ReshapeBackward(std::vector<int64_t> s) : s_(std::move(s)) {}
    std::vector<TensorPtr> apply(const TensorPtr& gy) override { return {gy->reshape(s_)}; }
private:
    std::vector<int64_t> s_;
};
}  // namespace

TensorPtr reshape(const TensorPtr& t, const std::vector<int64_t>& shape) {   // core/ops.cpp:1594-1640
    if (!t) throw TensorError("reshape: null tensor");
    auto out = t->reshape(shape);
    out->set_requires_grad(false);
    track(out, {t}, std::make_shared<ReshapeBackward>(t->shape()));
    return out;
}
TensorPtr view(const TensorPtr& t, const std::vector<int64_t>& shape) { return reshape(t, shape); }
TensorPtr flatten(const TensorPtr& t, int start_dim, int end_dim) {
    const int nd = int(t->ndim());
    if (start_dim < 0) start_dim += nd;
    if (end_dim < 0) end_dim += nd;
    if (start_dim < 0 || end_dim >= nd || start_dim > end_dim) throw TensorError("flatten: bad dims");
    std::vector<int64_t> s(t->shape().begin(), t->shape().begin() + start_dim);
    int64_t f = 1;
    for (int i = start_dim; i <= end_dim; ++i) f *= t->shape()[i];
    s.push_back(f);
    s.insert(s.end(), t->shape().begin() + end_dim + 1, t->shape().end());
    return reshape(t, s);
}
TensorPtr squeeze(const TensorPtr& t, int dim) { return reshape(t, t->squeeze(dim)->shape()); }
TensorPtr unsqueeze(const TensorPtr& t, int dim) { return reshape(t, t->unsqueeze(dim)->shape()); }

// ---------------------------------------------------------------------------- reductions
namespace {
class SumBackward : public BackwardFunction {        // core/backward_functions.cpp (SumBackward): broadcast gy back
public:
    SumBackward(std::vector<int64_t> in_shape, int dim, float scale) : s_(std::move(in_shape)), dim_(dim), scale_(scale) {}
    std::vector<TensorPtr> apply(const TensorPtr& gy_in) override {
        autograd::Engine& e = autograd::Engine::instance();
        const bool was = e.is_enabled();
        e.set_enabled(false);
        TensorPtr gy = contiguous(gy_in);
        TensorPtr g;
        int64_t n = 1;
        for (auto d : s_) n *= d;
        if (dim_ < 0) {                               // reduced over everything: g = gy[0] * scale everywhere
            auto onesv = full(s_, scale_);
            g = mul(onesv, gy->reshape({1}));
        } else {
            int64_t outer = 1, inner = 1;
            for (int i = 0; i < dim_; ++i) outer *= s_[i];
            for (size_t i = dim_ + 1; i < s_.size(); ++i) inner *= s_[i];
            auto z = full({outer, s_[dim_], inner}, 0.0f);
            g = mul(add(z, gy->reshape({outer, 1, inner})), scale_)->reshape(s_);
        }
        e.set_enabled(was);
        return {g};
    }
private:
    std::vector<int64_t> s_;
    int dim_;
    float scale_;
};

TensorPtr reduce_sum(const TensorPtr& t_in, int dim, bool keepdim, bool is_mean) {
    require_f32(t_in, "sum");
    TensorPtr t = contiguous(t_in);
    const int nd = int(t->ndim());
    TensorPtr out;
    float scale = 1.0f;
    if (dim == -1) {                                  // sentinel: all elements (core/ops.cpp:1787,1880-1893)
        out = Tensor::make({1}, kFloat32);
        size_t wsb = mft_colsum_workspace_bytes(t->numel(), 1);
        mft_rt::check(mft_colsum(t->dev_f32(), 1, out->dev_f32_out(), t->numel(), 1, mft_rt::workspace(wsb), wsb, st()), "mft_colsum");
        if (is_mean) scale = 1.0f / float(std::max<int64_t>(t->numel(), 1));
    } else {
        if (dim < 0) dim += nd;
        if (dim < 0 || dim >= nd) throw TensorError("sum: dimension out of range");
        const auto& s = t->shape();
        int64_t outer = 1, inner = 1;
        for (int i = 0; i < dim; ++i) outer *= s[i];
        for (int i = dim + 1; i < nd; ++i) inner *= s[i];
        std::vector<int64_t> so;
        for (int i = 0; i < nd; ++i) { if (i == dim) { if (keepdim) so.push_back(1); } else so.push_back(s[i]); }
        if (so.empty()) so.push_back(1);
        out = Tensor::make(so, kFloat32);
        size_t wsb = mft_colsum_workspace_bytes(s[dim], inner);
        for (int64_t o = 0; o < outer; ++o)
            mft_rt::check(mft_colsum(t->dev_f32() + o * s[dim] * inner, inner, out->dev_f32_out() + o * inner, s[dim], inner,
                                     mft_rt::workspace(wsb), wsb, st()), "mft_colsum");
        if (is_mean) scale = 1.0f / float(std::max<int64_t>(s[dim], 1));
    }
    if (scale != 1.0f) mft_rt::check(mft_affine(out->dev_f32(), out->dev_f32_mut(), out->numel(), scale, 0.f, st()), "mft_affine");
    track(out, {t_in}, std::make_shared<SumBackward>(t->shape(), dim == -1 ? -1 : dim, scale));
    return out;
}
}  // namespace

TensorPtr sum(const TensorPtr& t, int dim, bool keepdim) { return reduce_sum(t, dim, keepdim, false); }
TensorPtr mean(const TensorPtr& t, int dim, bool keepdim) { return reduce_sum(t, dim, keepdim, true); }

}  // namespace ops
