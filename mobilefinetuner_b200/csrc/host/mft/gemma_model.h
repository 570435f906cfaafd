// Gemma-3 text graph — API mirror of the reference's graph/gemma_model.h:17-178. Same config / block structs and
// public methods; the forward pass runs on the device op library (RMSNorm(1+w), LoRA-fused linears, interleaved
// RoPE, GQA repeat, QK^T/softmax/PV, GeGLU). Fully differentiable, like the reference's.
#pragma once
#include <memory>
#include <string>
#include <unordered_map>
#include <unordered_set>
#include <vector>

#include "lora_linear.h"
#include "parameter_sharder.h"
#include "tensor.h"

namespace ops {

struct GemmaTextConfig {
    int vocab_size = 262144;
    int hidden_size = 640;
    int intermediate_size = 2048;
    int num_hidden_layers = 18;
    int num_attention_heads = 4;
    int num_key_value_heads = 1;
    int head_dim = 256;
    int max_position_embeddings = 32768;
    int sliding_window = 512;
    bool attention_bias = false;
    bool use_bidirectional_attention = false;
    bool use_cache = true;
    float attention_dropout = 0.0f;
    float rms_norm_eps = 1e-6f;
    float query_pre_attn_scalar = 256.0f;
    float attn_logit_softcapping = 0.0f;
    float final_logit_softcapping = 0.0f;
    float rope_theta = 1000000.0f;
    float rope_local_base_freq = 10000.0f;
    std::string hidden_activation = "gelu_pytorch_tanh";
    std::vector<std::string> layer_types;
    static GemmaTextConfig from_pretrained(const std::string& model_dir);
};

struct GemmaBlockWeights {
    TensorPtr input_layernorm_weight, post_attention_layernorm_weight;
    TensorPtr pre_feedforward_layernorm_weight, post_feedforward_layernorm_weight;
    TensorPtr q_proj_weight, k_proj_weight, v_proj_weight, o_proj_weight;
    TensorPtr q_norm_weight, k_norm_weight;
    TensorPtr gate_proj_weight, up_proj_weight, down_proj_weight;
    std::unique_ptr<LoRALinear> q_proj_lora, k_proj_lora, v_proj_lora, o_proj_lora;
    std::unique_ptr<LoRALinear> gate_proj_lora, up_proj_lora, down_proj_lora;
    bool lora_initialized = false;
};

struct EmbeddingDumpRequest {
    bool active = false;
    int target_step = 0;
    std::string output_dir;
    bool fulfilled = false;
};

class GemmaModel {
public:
    explicit GemmaModel(const GemmaTextConfig& config);
    ~GemmaModel() = default;

    TensorPtr forward(const TensorPtr& input_ids, const TensorPtr& attention_mask = nullptr);
    void assign_weight(const std::string& key, const TensorPtr& tensor);
    const GemmaTextConfig& config() const { return config_; }
    void init_lora_modules();
    GemmaBlockWeights& get_block(int i);
    const GemmaBlockWeights& get_block(int i) const;
    std::vector<TensorPtr> get_lora_parameters() const;
    void merge_lora();
    void unmerge_lora();

    // Alignment/debug hooks of the reference (graph/gemma_model.cpp:736-944). Tensor dumps are host-side tooling
    // outside the hot path: the switches are accepted and recorded, nothing is written.
    void request_embedding_dump(int step, const std::string& output_dir);
    void enable_debug_dump(const std::string& dir, const std::vector<int>& layers);
    void disable_debug_dump();
    bool debug_enabled() const { return debug_enabled_; }
    const std::unordered_map<std::string, TensorPtr>& debug_tensors() const { return debug_tensors_; }
    void set_debug_retain_grads(bool) {}
    void set_numeric_perturb(bool, const std::string&, int64_t, float) {}
    void dump_layer_norm_weights(int, const std::string&) const {}

    void set_parameter_sharder(sharding::ParameterSharder* sharder) { sharder_ = sharder; }
    TensorPtr& embed_weight_ref() { return embed_weight_; }
    TensorPtr& norm_weight_ref() { return norm_weight_; }
    TensorPtr& lm_head_weight_ref() { return lm_head_weight_; }

private:
    GemmaTextConfig config_;
    TensorPtr embed_weight_, norm_weight_, lm_head_weight_;
    std::vector<GemmaBlockWeights> blocks_;
    bool lm_head_initialized_ = false;
    EmbeddingDumpRequest dump_request_;
    bool debug_enabled_ = false;
    std::unordered_map<std::string, TensorPtr> debug_tensors_;
    sharding::ParameterSharder* sharder_ = nullptr;
    std::unordered_map<int64_t, TensorPtr> causal_masks_, sliding_masks_;   // cached per sequence length

    TensorPtr mask_for(int64_t seq_len, bool sliding);
    TensorPtr padding_mask(const TensorPtr& attention_mask) const;
    TensorPtr attention(const TensorPtr& x, GemmaBlockWeights& b, const TensorPtr& pad_mask, const TensorPtr& base_mask, float theta) const;
    TensorPtr mlp(const TensorPtr& x, GemmaBlockWeights& b) const;
};

}  // namespace ops
