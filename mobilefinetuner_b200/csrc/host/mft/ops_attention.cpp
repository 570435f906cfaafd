// Fused attention nodes of the host API: per-head RMSNorm + RoPE for q/k, and grouped-query attention with a real
// backward. One autograd node each where the reference builds a chain (graph/gemma_model.cpp:383-514): permutes, rms_norm,
// apply_rope, repeat_kv_heads, matmul, mul, add(mask), softmax, matmul. Same values; no [B,H,S,S] tensor, no repeated K/V.
#include <cmath>
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

class NormRopeBackward : public BackwardFunction {
public:
    NormRopeBackward(TensorPtr x, TensorPtr w, int heads, int S, int Hd, float eps, float theta, bool do_norm)
        : x_(std::move(x)), w_(std::move(w)), heads_(heads), S_(S), Hd_(Hd), eps_(eps), theta_(theta), do_norm_(do_norm) {}
    std::vector<TensorPtr> apply(const TensorPtr& gy_in) override {
        TensorPtr gy = contiguous(gy_in);
        auto gx = new_like(x_);
        const int64_t tokens = x_->numel() / (int64_t(heads_) * Hd_);
        mft_rt::check(mft_qk_norm_rope_bwd(x_->dev_f32(), do_norm_ ? w_->dev_f32() : nullptr, gy->dev_f32(), gx->dev_f32_out(), tokens, heads_, S_,
                                           Hd_, eps_, theta_, do_norm_ ? 1 : 0, st()), "mft_qk_norm_rope_bwd");
        return {gx, nullptr};
    }
private:
    TensorPtr x_, w_;
    int heads_, S_, Hd_;
    float eps_, theta_;
    bool do_norm_;
};

class GqaAttentionBackward : public BackwardFunction {
public:
    GqaAttentionBackward(TensorPtr q, TensorPtr k, TensorPtr v, TensorPtr bias, TensorPtr ctx, TensorPtr lse, int H, int KV, int Hd, float scale,
                         int window, GqaLayout lay)
        : q_(std::move(q)), k_(std::move(k)), v_(std::move(v)), bias_(std::move(bias)), ctx_(std::move(ctx)), lse_(std::move(lse)), H_(H), KV_(KV),
          Hd_(Hd), scale_(scale), window_(window), lay_(lay) {}
    std::vector<TensorPtr> apply(const TensorPtr& gy_in) override {
        TensorPtr gy = contiguous(gy_in);
        const int64_t B = ctx_->shape()[0], S = ctx_->shape()[1];
        const size_t wsb = mft_gqa_attention_bwd_workspace_bytes(int32_t(B), H_, KV_, int32_t(S), Hd_);
        if (lay_.fused_qkv) {
            // q, k, v are column blocks of one [B,S,(H+2KV)*Hd] projection: one gradient tensor in the same layout
            auto g = new_like(q_);
            float* base = g->dev_f32_out();
            const float* src = q_->dev_f32();
            const int64_t W = q_->shape().back();
            mft_rt::check(mft_gqa_attention_bwd(src + lay_.q_off, src + lay_.k_off, src + lay_.v_off, bias_ ? bias_->dev_f32() : nullptr,
                                                ctx_->dev_f32(), lse_->dev_f32(), gy->dev_f32(), base + lay_.q_off, base + lay_.k_off, base + lay_.v_off,
                                                int32_t(B), H_, KV_, int32_t(S), Hd_, S * W, W, Hd_, S * W, W, Hd_, S * H_ * Hd_, int64_t(H_) * Hd_, Hd_,
                                                S * W, W, Hd_, S * W, W, Hd_, scale_, window_, mft_rt::workspace(wsb), wsb, st()), "mft_gqa_attention_bwd");
            return {g};
        }
        auto gq = new_like(q_), gk = new_like(k_), gv = new_like(v_);
        const int64_t QW = int64_t(H_) * Hd_, KW = int64_t(KV_) * Hd_;
        mft_rt::check(mft_gqa_attention_bwd(q_->dev_f32(), k_->dev_f32(), v_->dev_f32(), bias_ ? bias_->dev_f32() : nullptr, ctx_->dev_f32(),
                                            lse_->dev_f32(), gy->dev_f32(), gq->dev_f32_out(), gk->dev_f32_out(), gv->dev_f32_out(),
                                            int32_t(B), H_, KV_, int32_t(S), Hd_, S * QW, QW, Hd_, S * KW, KW, Hd_, S * QW, QW, Hd_,
                                            S * QW, QW, Hd_, S * KW, KW, Hd_, scale_, window_, mft_rt::workspace(wsb), wsb, st()), "mft_gqa_attention_bwd");
        return {gq, gk, gv};
    }
private:
    TensorPtr q_, k_, v_, bias_, ctx_, lse_;
    int H_, KV_, Hd_;
    float scale_;
    int window_;
    GqaLayout lay_;
};

}  // namespace

bool gqa_head_dim_supported(int64_t Hd) { return Hd == 16 || Hd == 32 || Hd == 64 || Hd == 128 || Hd == 256; }

std::pair<TensorPtr, TensorPtr> qk_norm_rope(const TensorPtr& q_in, const TensorPtr& k_in, const TensorPtr& wq, const TensorPtr& wk, int H, int KV,
                                             int Hd, float eps, float theta, bool do_norm) {
    require_f32(q_in, "qk_norm_rope");
    require_f32(k_in, "qk_norm_rope");
    TensorPtr q = contiguous(q_in), k = contiguous(k_in);
    if (q->ndim() != 3 || k->ndim() != 3 || q->shape()[2] != int64_t(H) * Hd || k->shape()[2] != int64_t(KV) * Hd)
        throw TensorError("qk_norm_rope: q must be [B,S,H*Hd] and k [B,S,KV*Hd]");
    const int64_t B = q->shape()[0], S = q->shape()[1];
    auto yq = new_like(q), yk = new_like(k);
    mft_rt::check(mft_qk_norm_rope_fwd(q->dev_f32(), do_norm ? wq->dev_f32() : nullptr, yq->dev_f32_out(), B * S, H,
                                       k->dev_f32(), do_norm ? wk->dev_f32() : nullptr, yk->dev_f32_out(), B * S, KV, int32_t(S), Hd, eps, theta,
                                       do_norm ? 1 : 0, st()), "mft_qk_norm_rope_fwd");
    track(yq, {q_in, wq}, std::make_shared<NormRopeBackward>(q, wq, H, int(S), Hd, eps, theta, do_norm));
    track(yk, {k_in, wk}, std::make_shared<NormRopeBackward>(k, wk, KV, int(S), Hd, eps, theta, do_norm));
    return {yq, yk};
}

TensorPtr gqa_attention(const TensorPtr& q_in, const TensorPtr& k_in, const TensorPtr& v_in, const TensorPtr& key_bias, int H, int KV, int Hd,
                        float scale, int window) {
    require_f32(q_in, "gqa_attention");
    TensorPtr q = contiguous(q_in), k = contiguous(k_in), v = contiguous(v_in);
    if (q->ndim() != 3 || q->shape()[2] != int64_t(H) * Hd || k->shape()[2] != int64_t(KV) * Hd || v->shape() != k->shape())
        throw TensorError("gqa_attention: q must be [B,S,H*Hd], k and v [B,S,KV*Hd]");
    if (!gqa_head_dim_supported(Hd)) throw TensorError("gqa_attention: head_dim " + std::to_string(Hd) + " not supported (16/32/64/128/256)");
    const int64_t B = q->shape()[0], S = q->shape()[1], QW = int64_t(H) * Hd, KW = int64_t(KV) * Hd;
    TensorPtr bias = key_bias ? contiguous(key_bias) : nullptr;
    if (bias && bias->numel() != B * S) throw TensorError("gqa_attention: key bias must hold B*S values");
    auto ctx = Tensor::make({B, S, QW}, kFloat32);
    const bool need_bwd = autograd::Engine::instance().is_enabled() &&
                          (autograd::Engine::instance().needs_grad(q_in) || autograd::Engine::instance().needs_grad(k_in) ||
                           autograd::Engine::instance().needs_grad(v_in));
    TensorPtr lse = need_bwd ? Tensor::make({B, int64_t(H), S}, kFloat32) : nullptr;
    float* cp = ctx->dev_f32_out();
    void* c16 = nullptr;
    if (emit_bf16_operands() && QW % 8 == 0) c16 = ctx->bf16_emit().ptr;        // ctx feeds the output projection's GEMM
    mft_rt::check(mft_gqa_attention_fwd(q->dev_f32(), k->dev_f32(), v->dev_f32(), bias ? bias->dev_f32() : nullptr, cp,
                                        lse ? lse->dev_f32_out() : nullptr, c16, int32_t(B), H, KV, int32_t(S), Hd, S * QW, QW, Hd, S * KW, KW, Hd,
                                        S * QW, QW, Hd, scale, window, st()), "mft_gqa_attention_fwd");
    if (need_bwd)
        track(ctx, {q_in, k_in, v_in}, std::make_shared<GqaAttentionBackward>(q, k, v, bias, ctx, lse, H, KV, Hd, scale, window, GqaLayout{}));
    return ctx;
}

// GPT-2 true-gradient mode: attention over the fused [B,S,3C] projection WITH a backward (the reference's default path
// detaches q/k/v, graph/gpt2_model.cpp:616-637; its MFT_TRUE_ATTN_GRAD intent is what pytorch_alignment/gpt2_lora_finetune.py
// trains with). One node: input qkv, output ctx [B,S,C].
TensorPtr gpt2_causal_attention_with_grad(const TensorPtr& qkv_in, int n_head) {
    require_f32(qkv_in, "gpt2_causal_attention");
    TensorPtr qkv = contiguous(qkv_in);
    const auto& s = qkv->shape();
    if (s.size() != 3 || s[2] % 3) throw TensorError("gpt2_causal_attention: qkv must be [B, S, 3C]");
    const int64_t B = s[0], S = s[1], C = s[2] / 3, Hd = C / n_head;
    if (!gqa_head_dim_supported(Hd)) throw TensorError("gpt2_causal_attention (true gradient): head_dim " + std::to_string(Hd) + " not supported");
    auto ctx = Tensor::make({B, S, C}, kFloat32);
    auto lse = Tensor::make({B, int64_t(n_head), S}, kFloat32);
    const float* base = qkv->dev_f32();
    float* cp = ctx->dev_f32_out();
    void* c16 = nullptr;
    if (emit_bf16_operands() && C % 8 == 0) c16 = ctx->bf16_emit().ptr;
    const float scale = 1.0f / std::sqrt(float(Hd));
    mft_rt::check(mft_gqa_attention_fwd(base, base + C, base + 2 * C, nullptr, cp, lse->dev_f32_out(), c16, int32_t(B), n_head, n_head, int32_t(S),
                                        int32_t(Hd), S * 3 * C, 3 * C, Hd, S * 3 * C, 3 * C, Hd, S * C, C, Hd, scale, 0, st()), "mft_gqa_attention_fwd");
    GqaLayout lay;
    lay.fused_qkv = true; lay.q_off = 0; lay.k_off = C; lay.v_off = 2 * C;
    track(ctx, {qkv_in}, std::make_shared<GqaAttentionBackward>(qkv, nullptr, nullptr, nullptr, ctx, lse, n_head, n_head, int(Hd), scale, 0, lay));
    return ctx;
}

}  // namespace ops
