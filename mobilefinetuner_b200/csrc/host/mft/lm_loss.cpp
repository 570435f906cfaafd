#include "lm_loss.h"

#include <cstdlib>

#include <stdexcept>

#include "autograd_engine.h"
#include "mft_b200.h"
#include "ops.h"
#include "ops_internal.h"
#include "runtime.h"

namespace ops {

namespace {
// LMCrossEntropyBackward (core/lm_loss.cpp:19-104): recomputes softmax from the saved per-row logsumexp; the
// upstream scalar gradient and the valid-row count are read on the device (no host round trip).
class LmCeBackward : public BackwardFunction {
public:
    LmCeBackward(TensorPtr logits, TensorPtr labels, TensorPtr lse, TensorPtr valid, int ignore, int reduction)
        : logits_(std::move(logits)), labels_(std::move(labels)), lse_(std::move(lse)), valid_(std::move(valid)),
          ignore_(ignore), reduction_(reduction) {}
    std::vector<TensorPtr> apply(const TensorPtr& gout) override {
        const auto& s = logits_->shape();
        auto g = Tensor::make(s, kFloat32, false, logits_->ld());
        float* gp = nullptr;
        void* g16 = nullptr;
        int64_t ld16 = 0;
        if (internal::bf16_only_results() && logits_->grad_may_be_bf16() && logits_->ld() % 4 == 0) {
            auto e = g->bf16_only_out();          // frozen head: dlogits only feeds the bf16 dX product (an fp32 reader gets an upcast)
            g16 = e.ptr; ld16 = e.ld;
        } else {
            gp = g->dev_f32_out();
            if (internal::emit_bf16_operands()) { auto e = g->bf16_emit(); g16 = e.ptr; ld16 = e.ld; }   // read by the lm_head dX / dW products
        }
        mft_rt::check(mft_lm_ce_bwd_ex(logits_->dev_f32(), logits_->ld(), static_cast<const int32_t*>(labels_->dev()), lse_->dev_f32(),
                                       static_cast<const int32_t*>(valid_->dev()), gout->dev_f32(), int32_t(s[0]), int32_t(s[1]),
                                       int32_t(s[2]), ignore_, reduction_, gp, g->ld(), g16, ld16, mft_rt::stream()), "mft_lm_ce_bwd");
        return {g};
    }
private:
    TensorPtr logits_, labels_, lse_, valid_;
    int ignore_, reduction_;
};
}  // namespace

// core/lm_loss.cpp:106-210
TensorPtr lm_cross_entropy(const TensorPtr& logits, const TensorPtr& labels, int ignore_index, const std::string& reduction) {
    if (!logits || logits->ndim() != 3) throw std::runtime_error("lm_cross_entropy: logits must be [B,S,V]");
    if (!labels || labels->ndim() != 2) throw std::runtime_error("lm_cross_entropy: labels must be [B,S]");
    if (labels->dtype() != kInt32) throw std::runtime_error("lm_cross_entropy: labels must be int32");
    const auto& s = logits->shape();
    if (labels->shape()[0] != s[0] || labels->shape()[1] != s[1]) throw std::runtime_error("lm_cross_entropy: labels shape mismatch");
    int red;
    if (reduction == "mean") red = 0;
    else if (reduction == "sum" || reduction == "sum_debug") red = 1;
    else throw std::runtime_error("lm_cross_entropy: reduction '" + reduction + "' is not supported on the device path (use mean or sum)");
    const int64_t rows = s[0] * s[1];
    auto loss = Tensor::make({1}, kFloat32);
    auto valid = Tensor::make({1}, kInt32);
    auto lse = Tensor::make({rows}, kFloat32);
    size_t wsb = mft_lm_ce_workspace_bytes(rows);
    mft_rt::check(mft_lm_ce_fwd(logits->dev_f32(), logits->ld(), static_cast<const int32_t*>(labels->dev()), int32_t(s[0]), int32_t(s[1]),
                                int32_t(s[2]), ignore_index, red, loss->dev_f32_out(), static_cast<int32_t*>(valid->dev_out()),
                                lse->dev_f32_out(), mft_rt::workspace(wsb), wsb, mft_rt::stream()), "mft_lm_ce_fwd");
    internal::track(loss, {logits}, std::make_shared<LmCeBackward>(logits, labels, lse, valid, ignore_index, red));
    return loss;
}

}  // namespace ops
