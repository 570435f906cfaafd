// GPT-2 graph — API mirror of the reference's graph/gpt2_model.h:21-186 (same config/block structs, same
// public methods). The forward pass is rebuilt on fused device ops: embedding gather (+wpe), LayerNorm,
// LoRA-fused tcgen05 linears, flash-style causal attention that reads the [B,S,3C] qkv buffer in place,
// tanh-GELU, residual adds, and a tied lm_head that never materialises wte^T.
#pragma once
#include <memory>
#include <string>
#include <unordered_map>
#include <utility>
#include <vector>

#include "lora_linear.h"
#include "parameter_sharder.h"
#include "tensor.h"

namespace ops {

struct GPT2BlockWeights {
    TensorPtr ln_1_weight, ln_1_bias;
    TensorPtr ln_2_weight, ln_2_bias;
    TensorPtr attn_qkv_weight;    // [C, 3C]
    TensorPtr attn_qkv_bias;      // [3C]
    TensorPtr attn_proj_weight;   // [C, C]
    TensorPtr attn_proj_bias;     // [C]
    TensorPtr mlp_fc_in_weight;   // [C, 4C]
    TensorPtr mlp_fc_in_bias;     // [4C]
    TensorPtr mlp_fc_out_weight;  // [4C, C]
    TensorPtr mlp_fc_out_bias;    // [C]
    std::unique_ptr<LoRALinear> qkv_lin, proj_lin, fc_in_lin, fc_out_lin;
    bool lora_initialized = false;
};

struct GPT2Config {
    int vocab_size = 50257;
    int n_positions = 1024;
    int n_embd = 768;
    int n_layer = 12;
    int n_head = 12;
    float layernorm_eps = 1e-5f;
    bool tie_word_embeddings = true;
    bool use_cache = false;
    bool use_memory_efficient_attention = true;   // always true here: S x S scores are never materialised
    bool use_bf16_activations = false;            // unused: activations are fp32, GEMM inputs TF32
    static GPT2Config from_pretrained(const std::string& config_path);
};

class GPT2Model {
public:
    explicit GPT2Model(const GPT2Config& config);
    ~GPT2Model() = default;

    TensorPtr forward(const TensorPtr& input_ids, const TensorPtr& attention_mask = nullptr);
    void tie_weights();
    void assign_weight(const std::string& key, const TensorPtr& tensor);
    std::vector<TensorPtr> parameters();
    std::vector<TensorPtr> trainable_parameters();
    void freeze_base_parameters();
    void print_model_info() const;
    const GPT2Config& config() const { return config_; }

    std::pair<TensorPtr*, TensorPtr*> attn_qkv_params(int i);
    std::pair<TensorPtr*, TensorPtr*> attn_proj_params(int i);
    std::pair<TensorPtr*, TensorPtr*> mlp_fc_in_params(int i);
    std::pair<TensorPtr*, TensorPtr*> mlp_fc_out_params(int i);
    GPT2BlockWeights& get_block(int i) { return blocks_[i]; }
    const GPT2BlockWeights& get_block(int i) const { return blocks_[i]; }
    std::vector<TensorPtr> get_lora_parameters();
    void init_lora_modules();
    void set_parameter_sharder(sharding::ParameterSharder* sharder) { sharder_ = sharder; }

    TensorPtr& wte_weight_ref() { return wte_weight_; }
    TensorPtr& wpe_weight_ref() { return wpe_weight_; }
    TensorPtr& ln_f_weight_ref() { return ln_f_weight_; }
    TensorPtr& ln_f_bias_ref() { return ln_f_bias_; }

private:
    GPT2Config config_;
    TensorPtr wte_weight_, wpe_weight_;
    std::vector<GPT2BlockWeights> blocks_;
    TensorPtr ln_f_weight_, ln_f_bias_;
    bool weights_tied_ = false;
    sharding::ParameterSharder* sharder_ = nullptr;

    GPT2BlockWeights& checked_block(int i, const char* who);
    TensorPtr run_linear(const TensorPtr& x, const std::unique_ptr<LoRALinear>& lin, bool lora_ready,
                         const TensorPtr& W, const TensorPtr& b) const;
    TensorPtr run_linear_fused(const TensorPtr& x, const std::unique_ptr<LoRALinear>& lin, bool lora_ready,
                               const TensorPtr& W, const TensorPtr& b, const LinearFusion& fusion) const;
    TensorPtr forward_block(const TensorPtr& x, int block_idx, const TensorPtr& carry = nullptr, TensorPtr* carry_out = nullptr);
};

}  // namespace ops
