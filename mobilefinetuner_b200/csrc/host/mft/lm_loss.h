// Shifted LM cross-entropy — API mirror of the reference's core/lm_loss.h:31-41.
#pragma once
#include <cmath>
#include <string>

#include "tensor.h"

namespace ops {

// logits [B,S,V] vs labels [B,S] (int32, NOT pre-shifted): position s is scored against labels[s+1]; labels equal to
// ignore_index (or outside [0,V)) are skipped; reduction "mean" | "sum" | "none".
TensorPtr lm_cross_entropy(const TensorPtr& logits, const TensorPtr& labels, int ignore_index = -100,
                           const std::string& reduction = "mean");

inline float perplexity_from_loss(float mean_nll) { return std::exp(mean_nll); }

}  // namespace ops
