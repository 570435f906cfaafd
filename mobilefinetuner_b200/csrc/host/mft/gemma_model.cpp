#include "gemma_model.h"

#include <cmath>
#include <cstdlib>
#include <filesystem>
#include <fstream>
#include <iostream>
#include <regex>
#include <sstream>
#include <stdexcept>

#include "ops.h"
#include "ops_internal.h"
#include "runtime.h"

namespace ops {

namespace {
constexpr float kMaskValue = -1e10f;   // graph/gemma_model.cpp (finite mask value, Appendix B12)

std::string slurp(const std::string& path) {
    std::ifstream f(path);
    if (!f.is_open()) throw std::runtime_error("Failed to open config file: " + path);
    std::stringstream ss;
    ss << f.rdbuf();
    return ss.str();
}
bool scrape(const std::string& t, const std::string& key, const std::string& value_re, std::string& out) {
    std::regex re("\"" + key + "\"\\s*:\\s*" + value_re, std::regex::icase);
    std::smatch m;
    if (!std::regex_search(t, m, re)) return false;
    out = m[1].str();
    return true;
}
const char* kNum = "(-?[0-9]+(?:\\.[0-9]+)?(?:[eE][-+]?[0-9]+)?)";
int jint(const std::string& t, const std::string& k, int d) { std::string v; return scrape(t, k, "(-?[0-9]+)", v) ? std::stoi(v) : d; }
float jfloat(const std::string& t, const std::string& k, float d) { std::string v; return scrape(t, k, kNum, v) ? std::stof(v) : d; }
bool jbool(const std::string& t, const std::string& k, bool d) { std::string v; return scrape(t, k, "(true|false)", v) ? (v[0] == 't' || v[0] == 'T') : d; }

TensorPtr param(std::vector<int64_t> shape) {
    mft_rt::PersistentScope keep;
    return std::make_shared<Tensor>(shape, kFloat32, kCPU);
}
TensorPtr project(const TensorPtr& x, const std::unique_ptr<LoRALinear>& lin, const TensorPtr& W) {
    return lin ? lin->forward(x) : matmul(x, W);
}
}  // namespace

// graph/gemma_model.cpp:138-171
GemmaTextConfig GemmaTextConfig::from_pretrained(const std::string& model_dir) {
    GemmaTextConfig c;
    std::filesystem::path p = model_dir;
    if (std::filesystem::is_directory(p)) p /= "config.json";
    const std::string t = slurp(p.string());
    c.vocab_size = jint(t, "vocab_size", c.vocab_size);
    c.hidden_size = jint(t, "hidden_size", c.hidden_size);
    c.intermediate_size = jint(t, "intermediate_size", c.intermediate_size);
    c.num_hidden_layers = jint(t, "num_hidden_layers", c.num_hidden_layers);
    c.num_attention_heads = jint(t, "num_attention_heads", c.num_attention_heads);
    c.num_key_value_heads = jint(t, "num_key_value_heads", c.num_key_value_heads);
    c.head_dim = jint(t, "head_dim", c.head_dim);
    c.max_position_embeddings = jint(t, "max_position_embeddings", c.max_position_embeddings);
    c.sliding_window = jint(t, "sliding_window", c.sliding_window);
    c.attention_bias = jbool(t, "attention_bias", c.attention_bias);
    c.use_bidirectional_attention = jbool(t, "use_bidirectional_attention", c.use_bidirectional_attention);
    c.use_cache = jbool(t, "use_cache", c.use_cache);
    c.attention_dropout = jfloat(t, "attention_dropout", c.attention_dropout);
    c.rms_norm_eps = jfloat(t, "rms_norm_eps", c.rms_norm_eps);
    c.query_pre_attn_scalar = jfloat(t, "query_pre_attn_scalar", c.query_pre_attn_scalar);
    c.rope_theta = jfloat(t, "rope_theta", c.rope_theta);
    c.rope_local_base_freq = jfloat(t, "rope_local_base_freq", c.rope_local_base_freq);
    std::string v;
    if (scrape(t, "hidden_activation", "\"([^\"]+)\"", v)) c.hidden_activation = v;
    std::regex arr("\"layer_types\"\\s*:\\s*\\[([^\\]]*)\\]");
    std::smatch m;
    if (std::regex_search(t, m, arr)) {
        const std::string body = m[1].str();
        std::regex item("\"([^\"]+)\"");
        for (auto it = std::sregex_iterator(body.begin(), body.end(), item); it != std::sregex_iterator(); ++it)
            c.layer_types.push_back((*it)[1].str());
    }
    if (c.layer_types.empty())   // default when absent: every layer sliding (graph/gemma_model.cpp:166-169)
        c.layer_types.assign(c.num_hidden_layers, "sliding_attention");
    return c;
}

// graph/gemma_model.cpp:173-220
GemmaModel::GemmaModel(const GemmaTextConfig& config) : config_(config) {
    const int64_t C = config_.hidden_size, I = config_.intermediate_size, Hd = config_.head_dim;
    const int64_t QD = config_.num_attention_heads * Hd, KD = config_.num_key_value_heads * Hd;
    if (config_.num_key_value_heads <= 0 || config_.num_attention_heads % config_.num_key_value_heads)
        throw std::runtime_error("GemmaModel: num_attention_heads must be a multiple of num_key_value_heads");
    embed_weight_ = param({config_.vocab_size, C});
    norm_weight_ = param({C});
    blocks_.resize(config_.num_hidden_layers);
    for (auto& b : blocks_) {
        b.input_layernorm_weight = param({C});           b.post_attention_layernorm_weight = param({C});
        b.pre_feedforward_layernorm_weight = param({C}); b.post_feedforward_layernorm_weight = param({C});
        b.q_proj_weight = param({C, QD}); b.k_proj_weight = param({C, KD}); b.v_proj_weight = param({C, KD});
        b.o_proj_weight = param({QD, C});
        b.q_norm_weight = param({Hd}); b.k_norm_weight = param({Hd});
        b.gate_proj_weight = param({C, I}); b.up_proj_weight = param({C, I}); b.down_proj_weight = param({I, C});
    }
}

// graph/gemma_model.cpp:687-734
void GemmaModel::assign_weight(const std::string& key, const TensorPtr& tensor) {
    if (key == "embed_tokens.weight") {
        embed_weight_ = tensor;
        if (!lm_head_initialized_) {
            mft_rt::PersistentScope keep;
            lm_head_weight_ = transpose(tensor, 0, 1);   // untied copy [C,V], created once (:689-693)
            lm_head_initialized_ = true;
        }
        return;
    }
    if (key == "norm.weight") { norm_weight_ = tensor; return; }
    if (key == "lm_head.weight") { lm_head_weight_ = tensor; lm_head_initialized_ = true; return; }
    static const std::regex re(R"(layers\.(\d+)\.(.+))");
    std::smatch m;
    if (!std::regex_match(key, m, re)) throw std::runtime_error("Unknown Gemma weight key: " + key);
    const int layer = std::stoi(m[1].str());
    if (layer < 0 || layer >= config_.num_hidden_layers) throw std::runtime_error("Invalid Gemma layer index: " + std::to_string(layer));
    GemmaBlockWeights& b = blocks_[layer];
    const std::unordered_map<std::string, TensorPtr*> slots = {
        {"input_layernorm.weight", &b.input_layernorm_weight}, {"post_attention_layernorm.weight", &b.post_attention_layernorm_weight},
        {"pre_feedforward_layernorm.weight", &b.pre_feedforward_layernorm_weight},
        {"post_feedforward_layernorm.weight", &b.post_feedforward_layernorm_weight},
        {"self_attn.q_proj.weight", &b.q_proj_weight}, {"self_attn.k_proj.weight", &b.k_proj_weight},
        {"self_attn.v_proj.weight", &b.v_proj_weight}, {"self_attn.o_proj.weight", &b.o_proj_weight},
        {"self_attn.q_norm.weight", &b.q_norm_weight}, {"self_attn.k_norm.weight", &b.k_norm_weight},
        {"mlp.gate_proj.weight", &b.gate_proj_weight}, {"mlp.up_proj.weight", &b.up_proj_weight},
        {"mlp.down_proj.weight", &b.down_proj_weight}};
    auto it = slots.find(m[2].str());
    if (it == slots.end()) throw std::runtime_error("Unknown Gemma block weight: " + m[2].str());
    *it->second = tensor;
}

GemmaBlockWeights& GemmaModel::get_block(int i) {
    if (i < 0 || i >= int(blocks_.size())) throw std::out_of_range("Gemma block index");
    return blocks_[i];
}
const GemmaBlockWeights& GemmaModel::get_block(int i) const {
    if (i < 0 || i >= int(blocks_.size())) throw std::out_of_range("Gemma block index");
    return blocks_[i];
}

void GemmaModel::init_lora_modules() {
    for (auto& b : blocks_) {
        if (b.lora_initialized) continue;
        b.q_proj_lora = std::make_unique<LoRALinear>(&b.q_proj_weight, nullptr);
        b.k_proj_lora = std::make_unique<LoRALinear>(&b.k_proj_weight, nullptr);
        b.v_proj_lora = std::make_unique<LoRALinear>(&b.v_proj_weight, nullptr);
        b.o_proj_lora = std::make_unique<LoRALinear>(&b.o_proj_weight, nullptr);
        b.gate_proj_lora = std::make_unique<LoRALinear>(&b.gate_proj_weight, nullptr);
        b.up_proj_lora = std::make_unique<LoRALinear>(&b.up_proj_weight, nullptr);
        b.down_proj_lora = std::make_unique<LoRALinear>(&b.down_proj_weight, nullptr);
        b.lora_initialized = true;
    }
}
static std::vector<LoRALinear*> lora_modules(const GemmaBlockWeights& b) {
    return {b.q_proj_lora.get(), b.k_proj_lora.get(), b.v_proj_lora.get(), b.o_proj_lora.get(),
            b.gate_proj_lora.get(), b.up_proj_lora.get(), b.down_proj_lora.get()};
}
std::vector<TensorPtr> GemmaModel::get_lora_parameters() const {
    std::vector<TensorPtr> out;
    for (const auto& b : blocks_)
        for (LoRALinear* l : lora_modules(b)) {
            if (!l) continue;
            auto ps = l->trainable_parameters();
            out.insert(out.end(), ps.begin(), ps.end());
        }
    return out;
}
void GemmaModel::merge_lora() { for (auto& b : blocks_) if (b.lora_initialized) for (LoRALinear* l : lora_modules(b)) if (l) l->merge_to_base(); }
void GemmaModel::unmerge_lora() { for (auto& b : blocks_) if (b.lora_initialized) for (LoRALinear* l : lora_modules(b)) if (l) l->unmerge_from_base(); }

void GemmaModel::request_embedding_dump(int step, const std::string& output_dir) {
    dump_request_.active = true; dump_request_.target_step = step; dump_request_.output_dir = output_dir; dump_request_.fulfilled = false;
}
void GemmaModel::enable_debug_dump(const std::string&, const std::vector<int>&) { debug_enabled_ = true; }
void GemmaModel::disable_debug_dump() { debug_enabled_ = false; debug_tensors_.clear(); }

// graph/gemma_model.cpp:261-285 — additive masks: causal, or causal restricted to the last `window` keys
TensorPtr GemmaModel::mask_for(int64_t S, bool sliding) {
    auto& cache = sliding ? sliding_masks_ : causal_masks_;
    auto it = cache.find(S);
    if (it != cache.end()) return it->second;
    mft_rt::PersistentScope keep;
    auto m = std::make_shared<Tensor>(std::vector<int64_t>{S, S}, kFloat32, kCPU);
    float* d = m->data<float>();
    const int64_t w = config_.sliding_window;
    for (int64_t i = 0; i < S; ++i)
        for (int64_t j = 0; j < S; ++j) {
            const bool allow = sliding ? (j <= i && i - j < w) : (j <= i);
            d[i * S + j] = allow ? 0.0f : kMaskValue;
        }
    (void)m->dev();   // upload now, while the persistent scope is active
    cache[S] = m;
    return m;
}

// graph/gemma_model.cpp:287-322
TensorPtr GemmaModel::padding_mask(const TensorPtr& am) const {
    if (!am) return nullptr;
    if (am->ndim() != 2) throw std::runtime_error("Gemma attention_mask must be [batch, seq_len]");
    const int64_t B = am->shape()[0], S = am->shape()[1];
    auto m = std::make_shared<Tensor>(std::vector<int64_t>{B, 1, 1, S}, kFloat32, kCPU);
    float* d = m->data<float>();
    if (am->dtype() == kFloat32) { const float* s = am->data<float>(); for (int64_t i = 0; i < B * S; ++i) d[i] = s[i] <= 0.5f ? kMaskValue : 0.0f; }
    else if (am->dtype() == kInt32) { const int32_t* s = am->data<int32_t>(); for (int64_t i = 0; i < B * S; ++i) d[i] = s[i] == 0 ? kMaskValue : 0.0f; }
    else throw std::runtime_error("Gemma attention_mask must be int32 or float32");
    return m;
}

// graph/gemma_model.cpp:358-520
TensorPtr GemmaModel::attention(const TensorPtr& x, GemmaBlockWeights& b, const TensorPtr& pad_mask, const TensorPtr& base_mask, float theta) const {
    const int64_t B = x->shape()[0], S = x->shape()[1];
    const int64_t H = config_.num_attention_heads, KV = config_.num_key_value_heads, Hd = config_.head_dim;
    auto q = project(x, b.q_proj_lora, b.q_proj_weight);
    auto k = project(x, b.k_proj_lora, b.k_proj_weight);
    auto v = project(x, b.v_proj_lora, b.v_proj_weight);
    const char* no_norm_env = std::getenv("DISABLE_QK_NORM");
    const bool do_norm = !(no_norm_env && std::string(no_norm_env) == "1");
    // Fused path (default): q/k RMSNorm + RoPE in one launch on the projections' own layout, then one grouped-query
    // attention node with its own backward — no permutes, no repeated K/V, no [B,H,S,S] scores. Same values as the chain
    // below, which stays for the cases the fused kernels do not cover (trained q/k norm weights, odd head sizes) and as
    // MFT_GEMMA_FUSED_ATTN=0.
    static const bool fused_on = [] { const char* e = std::getenv("MFT_GEMMA_FUSED_ATTN"); return !(e && e[0] == '0'); }();
    if (fused_on && gqa_head_dim_supported(Hd) && H % KV == 0 && !internal::wants_grad(b.q_norm_weight) && !internal::wants_grad(b.k_norm_weight)) {
        auto qk = qk_norm_rope(q, k, b.q_norm_weight, b.k_norm_weight, int(H), int(KV), int(Hd), config_.rms_norm_eps, theta, do_norm);
        bool sliding = false;                         // which of the two cached masks the caller passed
        { auto it = sliding_masks_.find(S); sliding = it != sliding_masks_.end() && it->second == base_mask; }
        TensorPtr key_bias = pad_mask ? reshape(pad_mask, {B, S}) : nullptr;
        auto fctx = gqa_attention(qk.first, qk.second, v, key_bias, int(H), int(KV), int(Hd), std::pow(config_.query_pre_attn_scalar, -0.5f),
                                  sliding ? int(config_.sliding_window) : 0);
        return project(fctx, b.o_proj_lora, b.o_proj_weight);
    }
    q = permute(reshape(q, {B, S, H, Hd}), {0, 2, 1, 3});
    k = permute(reshape(k, {B, S, KV, Hd}), {0, 2, 1, 3});
    v = permute(reshape(v, {B, S, KV, Hd}), {0, 2, 1, 3});
    const char* no_qk_norm = std::getenv("DISABLE_QK_NORM");
    if (!(no_qk_norm && std::string(no_qk_norm) == "1")) {
        q = rms_norm(q, b.q_norm_weight, config_.rms_norm_eps);
        k = rms_norm(k, b.k_norm_weight, config_.rms_norm_eps);
    }
    q = apply_rope(q, int(S), int(Hd), theta);
    k = apply_rope(k, int(S), int(Hd), theta);
    auto k_full = repeat_kv_heads(k, int(H / KV));
    auto v_full = repeat_kv_heads(v, int(H / KV));
    auto scores = matmul_rhs_T(q, k_full);                       // Q K^T without materialising K^T
    scores = mul(scores, std::pow(config_.query_pre_attn_scalar, -0.5f));
    if (base_mask) scores = add(scores, base_mask);
    if (pad_mask) scores = add(scores, pad_mask);
    auto probs = softmax(scores, -1);
    auto ctx = matmul(probs, v_full);
    ctx = reshape(permute(ctx, {0, 2, 1, 3}), {B, S, b.q_proj_weight->shape()[1]});
    return project(ctx, b.o_proj_lora, b.o_proj_weight);
}

// graph/gemma_model.cpp:522-555 — GeGLU: down( gelu(gate(x)) * up(x) )
TensorPtr GemmaModel::mlp(const TensorPtr& x, GemmaBlockWeights& b) const {
    auto gate = project(x, b.gate_proj_lora, b.gate_proj_weight);
    auto up = project(x, b.up_proj_lora, b.up_proj_weight);
    static const bool fused = [] { const char* e = std::getenv("MFT_GEMMA_FUSED_MLP"); return !(e && e[0] == '0'); }();
    return project(fused ? geglu(gate, up) : mul(gelu(gate), up), b.down_proj_lora, b.down_proj_weight);
}

// graph/gemma_model.cpp:579-685
TensorPtr GemmaModel::forward(const TensorPtr& input_ids, const TensorPtr& attention_mask) {
    if (!input_ids || input_ids->ndim() != 2) throw std::runtime_error("GemmaModel::embedding_lookup expects [batch, seq_len]");
    if (sharder_) embed_weight_ = sharder_->require("embed_tokens.weight");
    const int64_t S = input_ids->shape()[1];
    auto h = embedding_lookup(input_ids, embed_weight_, nullptr, std::sqrt(float(config_.hidden_size)));
    TensorPtr causal = mask_for(S, false), sliding = mask_for(S, true);
    TensorPtr pad = padding_mask(attention_mask);
    const float eps = config_.rms_norm_eps;
    for (int i = 0; i < config_.num_hidden_layers; ++i) {
        GemmaBlockWeights& b = blocks_[i];
        const bool is_sliding = i < int(config_.layer_types.size()) && config_.layer_types[i] == "sliding_attention";
        const float theta = is_sliding ? config_.rope_local_base_freq : config_.rope_theta;
        auto normed = rms_norm(h, b.input_layernorm_weight, eps);
        auto attn = attention(normed, b, pad, is_sliding ? sliding : causal, theta);
        attn = rms_norm_post(attn, b.post_attention_layernorm_weight, eps);
        h = add(h, attn);
        auto mlp_in = rms_norm(h, b.pre_feedforward_layernorm_weight, eps);
        auto m = rms_norm_post(mlp(mlp_in, b), b.post_feedforward_layernorm_weight, eps);
        h = add(h, m);
    }
    h = rms_norm(h, norm_weight_, eps);
    return matmul(h, lm_head_weight_);
}

}  // namespace ops
