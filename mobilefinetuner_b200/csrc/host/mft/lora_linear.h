// LoRALinear — API mirror of the reference's nn/lora_linear.h:17-97. Holds POINTERS to the model's base
// weight/bias TensorPtrs (so assign_weight after init_lora_modules is seen) plus a list of LoRA slices.
// forward() issues one fused tcgen05 GEMM (base + rank-r update + bias) and registers ONE autograd node
// where the reference builds matmul/add/matmul/matmul/mul/add (nn/lora_linear.cpp:47-108).
#pragma once
#include <memory>
#include <string>
#include <utility>
#include <vector>

#include "tensor.h"

namespace ops {

struct LinearFusion;   // ops.h

struct LoRASlice {
    TensorPtr A;   // [in, r]  (or PEFT [r, in])
    TensorPtr B;   // [r, out_slice] (or PEFT [out_slice, r])
    float scale;   // alpha / r
    int col0;      // first output column covered by this slice
    int cols;      // number of output columns covered
    LoRASlice(const TensorPtr& a, const TensorPtr& b, float s, int c0 = 0, int c = -1) : A(a), B(b), scale(s), col0(c0), cols(c) {}
};

class LoRALinear {
public:
    LoRALinear(TensorPtr* W_base_ref, TensorPtr* b_base_ref = nullptr) : W_ref_(W_base_ref), b_ref_(b_base_ref), merged_(false) {}

    void attach_lora(const TensorPtr& A, const TensorPtr& B, float scale, int col0 = 0, int cols = -1);
    void clear_lora();
    void merge_to_base();
    void unmerge_from_base();
    TensorPtr forward(const TensorPtr& x) const;
    // B200 extension: gelu(forward(x)) with the activation produced by the GEMM epilogue (returns the activation;
    // the pre-activation stays in the graph for GeluBackward).
    TensorPtr forward_gelu(const TensorPtr& x, TensorPtr* pre = nullptr) const;
    // B200 extension: forward with the neighbouring residual add / GELU backward folded into the GEMM epilogues
    // (see ops::LinearFusion).
    TensorPtr forward_fused(const TensorPtr& x, const LinearFusion& fusion) const;

    void set_debug_name(const std::string& name) { debug_name_ = name; }
    std::vector<std::pair<std::string, TensorPtr>> debug_params() const;
    std::vector<TensorPtr> trainable_parameters() const;
    const TensorPtr& W() const { return base_W(); }
    const TensorPtr& b() const { return base_b(); }
    const std::vector<LoRASlice>& slices() const { return slices_; }
    bool is_merged() const { return merged_; }
    // B200 extension: true when the forward adds a rank-r update (adapters attached and not merged into the base weight)
    bool has_adapter() const { return !merged_ && !slices_.empty(); }

private:
    TensorPtr* W_ref_ = nullptr;
    TensorPtr* b_ref_ = nullptr;
    TensorPtr b_cache_;
    std::vector<LoRASlice> slices_;
    bool merged_;
    std::string debug_name_;
    const TensorPtr& base_W() const;
    const TensorPtr& base_b() const;
    void add_delta_to_base(float sign);
    TensorPtr forward_impl(const TensorPtr& x, const LinearFusion& fusion) const;
};

}  // namespace ops
