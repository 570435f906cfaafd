#include "memory_efficient_attention.h"

#include <cmath>
#include <stdexcept>

#include "mft_b200.h"
#include "ops.h"
#include "runtime.h"

namespace ops {

// core/memory_efficient_attention.cpp:40-185. The additive causal mask argument is redundant with
// use_causal_mask (the reference applies both, identical softmax — Appendix B12) and is ignored.
TensorPtr memory_efficient_attention(const TensorPtr& q_in, const TensorPtr& k_in, const TensorPtr& v_in,
                                     const TensorPtr& causal_mask, const MemoryEfficientAttentionConfig& cfg) {
    (void)causal_mask;
    if (!q_in || !k_in || !v_in) throw std::runtime_error("memory_efficient_attention: null input");
    if (q_in->ndim() != 4 || k_in->shape() != q_in->shape() || v_in->shape() != q_in->shape())
        throw std::runtime_error("memory_efficient_attention: q,k,v must be [B,H,S,Hd] with equal shapes");
    if (!cfg.use_causal_mask) throw std::runtime_error("memory_efficient_attention: only the causal form is implemented");
    TensorPtr q = contiguous(q_in), k = contiguous(k_in), v = contiguous(v_in);
    const auto& s = q->shape();
    const int64_t B = s[0], H = s[1], S = s[2], Hd = s[3];
    auto ctx = Tensor::make(s, kFloat32);
    const float scale = cfg.scale > 0 ? cfg.scale : 1.0f / std::sqrt(float(Hd));
    mft_rt::check(mft_causal_attention_fwd(q->dev_f32(), k->dev_f32(), v->dev_f32(), ctx->dev_f32_out(), int32_t(B), int32_t(H),
                                           int32_t(S), int32_t(Hd), H * S * Hd, Hd, S * Hd, H * S * Hd, Hd, S * Hd, H * S * Hd, Hd,
                                           S * Hd, scale, 0, mft_rt::stream()), "mft_causal_attention_fwd");
    return ctx;
}

}  // namespace ops
