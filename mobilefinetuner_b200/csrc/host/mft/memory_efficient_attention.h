// Causal SDPA forward — API mirror of the reference's core/memory_efficient_attention.h.
#pragma once
#include <vector>

#include "tensor.h"

namespace ops {

struct MemoryEfficientAttentionConfig {
    bool use_causal_mask = true;
    float scale = -1.0f;      // -1: 1/sqrt(head_dim)
    int chunk_size = 512;     // ignored: the CUDA kernel always streams K/V in tiles
    bool save_probs = false;  // ignored: probabilities are never materialised
};

// q,k,v [B,H,S,Hd] -> ctx [B,H,S,Hd]. Forward only (the reference's backward is a stub).
TensorPtr memory_efficient_attention(const TensorPtr& q, const TensorPtr& k, const TensorPtr& v,
                                     const TensorPtr& causal_mask = nullptr,
                                     const MemoryEfficientAttentionConfig& config = {});

}  // namespace ops
