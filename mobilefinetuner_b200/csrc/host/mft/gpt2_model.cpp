#include <cstdlib>
#include "gpt2_model.h"

#include <filesystem>
#include <fstream>
#include <iostream>
#include <regex>
#include <sstream>
#include <stdexcept>

#include "ops.h"
#include "mft_b200.h"
#include "ops_internal.h"
#include "runtime.h"

namespace ops {

namespace {
std::string slurp(const std::string& path) {
    std::ifstream f(path);
    if (!f.is_open()) throw std::runtime_error("Failed to open config file: " + path);
    std::stringstream ss;
    ss << f.rdbuf();
    return ss.str();
}
// config.json values are scraped with regexes like the reference does (graph/gpt2_model.cpp:41-71)
bool find_number(const std::string& text, const std::string& key, std::string& out) {
    std::regex re("\"" + key + "\"\\s*:\\s*(-?[0-9]+(\\.[0-9]+)?([eE][-+]?[0-9]+)?)");
    std::smatch m;
    if (!std::regex_search(text, m, re)) return false;
    out = m[1].str();
    return true;
}
int json_int(const std::string& t, const std::string& k, int d) { std::string v; return find_number(t, k, v) ? std::stoi(v) : d; }
float json_float(const std::string& t, const std::string& k, float d) { std::string v; return find_number(t, k, v) ? std::stof(v) : d; }
bool json_bool(const std::string& t, const std::string& k, bool d) {
    std::regex re("\"" + k + "\"\\s*:\\s*(true|false)", std::regex::icase);
    std::smatch m;
    if (!std::regex_search(t, m, re)) return d;
    return m[1].str()[0] == 't' || m[1].str()[0] == 'T';
}
TensorPtr param(std::vector<int64_t> shape) {
    mft_rt::PersistentScope keep;   // parameters outlive the step arena
    return std::make_shared<Tensor>(shape, kFloat32, kCPU);
}
}  // namespace

// graph/gpt2_model.cpp:121-140
GPT2Config GPT2Config::from_pretrained(const std::string& config_path) {
    GPT2Config c;
    std::filesystem::path p = config_path;
    if (std::filesystem::is_directory(p)) p /= "config.json";
    const std::string t = slurp(p.string());
    c.vocab_size = json_int(t, "vocab_size", c.vocab_size);
    c.n_positions = json_int(t, "n_positions", c.n_positions);
    c.n_positions = json_int(t, "n_ctx", c.n_positions);
    c.n_embd = json_int(t, "n_embd", c.n_embd);
    c.n_layer = json_int(t, "n_layer", c.n_layer);
    c.n_head = json_int(t, "n_head", c.n_head);
    c.layernorm_eps = json_float(t, "layer_norm_epsilon", c.layernorm_eps);
    c.tie_word_embeddings = json_bool(t, "tie_word_embeddings", c.tie_word_embeddings);
    c.use_cache = json_bool(t, "use_cache", c.use_cache);
    return c;
}

// graph/gpt2_model.cpp:142-191 — every parameter exists (zero-filled) right after construction
GPT2Model::GPT2Model(const GPT2Config& config) : config_(config) {
    const int64_t C = config_.n_embd;
    if (config_.n_head <= 0 || C % config_.n_head) throw std::runtime_error("GPT2Model: n_embd must be divisible by n_head");
    wte_weight_ = param({config_.vocab_size, C});
    wpe_weight_ = param({config_.n_positions, C});
    blocks_.resize(config_.n_layer);
    for (auto& b : blocks_) {
        b.ln_1_weight = param({C});            b.ln_1_bias = param({C});
        b.attn_qkv_weight = param({C, 3 * C}); b.attn_qkv_bias = param({3 * C});
        b.attn_proj_weight = param({C, C});    b.attn_proj_bias = param({C});
        b.ln_2_weight = param({C});            b.ln_2_bias = param({C});
        b.mlp_fc_in_weight = param({C, 4 * C}); b.mlp_fc_in_bias = param({4 * C});
        b.mlp_fc_out_weight = param({4 * C, C}); b.mlp_fc_out_bias = param({C});
    }
    ln_f_weight_ = param({C});
    ln_f_bias_ = param({C});
    std::cout << "[GPT2Model] Initialized with " << config_.n_layer << " layers, " << config_.n_embd << " hidden, "
              << config_.n_head << " heads" << std::endl;
}

void GPT2Model::tie_weights() {
    weights_tied_ = true;
    std::cout << "[GPT2Model] Tied lm_head.weight <-> wte.weight" << std::endl;
}

void GPT2Model::init_lora_modules() {
    for (auto& b : blocks_) {
        if (b.lora_initialized) continue;
        b.qkv_lin = std::make_unique<LoRALinear>(&b.attn_qkv_weight, &b.attn_qkv_bias);
        b.proj_lin = std::make_unique<LoRALinear>(&b.attn_proj_weight, &b.attn_proj_bias);
        b.fc_in_lin = std::make_unique<LoRALinear>(&b.mlp_fc_in_weight, &b.mlp_fc_in_bias);
        b.fc_out_lin = std::make_unique<LoRALinear>(&b.mlp_fc_out_weight, &b.mlp_fc_out_bias);
        b.lora_initialized = true;
    }
}

std::vector<TensorPtr> GPT2Model::get_lora_parameters() {
    std::vector<TensorPtr> out;
    for (const auto& b : blocks_) {
        if (!b.lora_initialized) continue;
        for (const LoRALinear* lin : {b.qkv_lin.get(), b.proj_lin.get(), b.fc_in_lin.get(), b.fc_out_lin.get()}) {
            if (!lin) continue;
            auto ps = lin->trainable_parameters();
            out.insert(out.end(), ps.begin(), ps.end());
        }
    }
    return out;
}

// graph/gpt2_model.cpp:246-305 (internal key names "blocks.N.<sub>")
void GPT2Model::assign_weight(const std::string& key, const TensorPtr& tensor) {
    if (key == "wte.weight") { wte_weight_ = tensor; return; }
    if (key == "wpe.weight") { wpe_weight_ = tensor; return; }
    if (key == "ln_f.weight") { ln_f_weight_ = tensor; return; }
    if (key == "ln_f.bias") { ln_f_bias_ = tensor; return; }
    static const std::regex re(R"(blocks\.(\d+)\.(.+))");
    std::smatch m;
    if (!std::regex_match(key, m, re)) { std::cerr << "[WARN] Unknown key: " << key << std::endl; return; }
    const int idx = std::stoi(m[1].str());
    if (idx < 0 || idx >= config_.n_layer) throw std::runtime_error("Invalid block index: " + std::to_string(idx));
    GPT2BlockWeights& b = blocks_[idx];
    const std::unordered_map<std::string, TensorPtr*> slots = {
        {"ln_1.weight", &b.ln_1_weight}, {"ln_1.bias", &b.ln_1_bias},
        {"attn.qkv.weight", &b.attn_qkv_weight}, {"attn.qkv.bias", &b.attn_qkv_bias},
        {"attn.proj.weight", &b.attn_proj_weight}, {"attn.proj.bias", &b.attn_proj_bias},
        {"ln_2.weight", &b.ln_2_weight}, {"ln_2.bias", &b.ln_2_bias},
        {"mlp.fc_in.weight", &b.mlp_fc_in_weight}, {"mlp.fc_in.bias", &b.mlp_fc_in_bias},
        {"mlp.fc_out.weight", &b.mlp_fc_out_weight}, {"mlp.fc_out.bias", &b.mlp_fc_out_bias}};
    auto it = slots.find(m[2].str());
    if (it == slots.end()) { std::cerr << "[WARN] Unknown sub_key: " << m[2].str() << std::endl; return; }
    *it->second = tensor;
}

GPT2BlockWeights& GPT2Model::checked_block(int i, const char* who) {
    if (i < 0 || i >= config_.n_layer) throw std::out_of_range(std::string(who) + ": layer index");
    return blocks_[i];
}
std::pair<TensorPtr*, TensorPtr*> GPT2Model::attn_qkv_params(int i) { auto& b = checked_block(i, "attn_qkv_params"); return {&b.attn_qkv_weight, &b.attn_qkv_bias}; }
std::pair<TensorPtr*, TensorPtr*> GPT2Model::attn_proj_params(int i) { auto& b = checked_block(i, "attn_proj_params"); return {&b.attn_proj_weight, &b.attn_proj_bias}; }
std::pair<TensorPtr*, TensorPtr*> GPT2Model::mlp_fc_in_params(int i) { auto& b = checked_block(i, "mlp_fc_in_params"); return {&b.mlp_fc_in_weight, &b.mlp_fc_in_bias}; }
std::pair<TensorPtr*, TensorPtr*> GPT2Model::mlp_fc_out_params(int i) { auto& b = checked_block(i, "mlp_fc_out_params"); return {&b.mlp_fc_out_weight, &b.mlp_fc_out_bias}; }

// graph/gpt2_model.cpp:837-861 — order matters to callers that fill parameters positionally
std::vector<TensorPtr> GPT2Model::parameters() {
    std::vector<TensorPtr> p{wte_weight_, wpe_weight_};
    for (auto& b : blocks_)
        for (const TensorPtr& t : {b.ln_1_weight, b.ln_1_bias, b.attn_qkv_weight, b.attn_qkv_bias, b.attn_proj_weight, b.attn_proj_bias,
                                   b.ln_2_weight, b.ln_2_bias, b.mlp_fc_in_weight, b.mlp_fc_in_bias, b.mlp_fc_out_weight, b.mlp_fc_out_bias})
            p.push_back(t);
    p.push_back(ln_f_weight_);
    p.push_back(ln_f_bias_);
    return p;
}
std::vector<TensorPtr> GPT2Model::trainable_parameters() {
    std::vector<TensorPtr> out;
    for (auto& t : parameters()) if (t && t->requires_grad()) out.push_back(t);
    for (auto& t : get_lora_parameters()) if (t && t->requires_grad()) out.push_back(t);
    return out;
}
void GPT2Model::freeze_base_parameters() { for (auto& t : parameters()) if (t) t->set_requires_grad(false); }

void GPT2Model::print_model_info() const {
    std::cout << "\n[GPT2Model Info]\n  vocab_size: " << config_.vocab_size << "\n  n_positions: " << config_.n_positions
              << "\n  n_embd: " << config_.n_embd << "\n  n_layer: " << config_.n_layer << "\n  n_head: " << config_.n_head
              << "\n  layernorm_eps: " << config_.layernorm_eps << "\n  weights_tied: " << (weights_tied_ ? "true" : "false")
              << "\n  backend: mft_b200 (sm_100a, tcgen05 TF32 GEMM, flash attention)" << std::endl;
}

TensorPtr GPT2Model::run_linear(const TensorPtr& x, const std::unique_ptr<LoRALinear>& lin, bool lora_ready,
                                const TensorPtr& W, const TensorPtr& b) const {
    if (lora_ready && lin) return lin->forward(x);
    return linear(x, W, b);            // reference fallback: matmul + add (graph/gpt2_model.cpp:592-600)
}

TensorPtr GPT2Model::run_linear_fused(const TensorPtr& x, const std::unique_ptr<LoRALinear>& lin, bool lora_ready,
                                      const TensorPtr& W, const TensorPtr& b, const LinearFusion& fusion) const {
    if (lora_ready && lin) return lin->forward_fused(x, fusion);
    return fused_lora_linear(x, W, b, {}, fusion);
}

// graph/gpt2_model.cpp:530-817. Autograd-visible structure preserved: the attention output carries no history
// (q/k/v are cut from the graph by the reference's zeros+memcpy split, :616-637), residual adds and
// LayerNorms are tracked ops.
// `carry` (optional): the MLP output of the previous block, not yet added to the residual stream `x_in`. The add then happens
// inside this block's ln_1 kernel (ops::add_layer_norm). On return *carry_out holds this block's own MLP output when the
// caller asked for it (carry_out != nullptr), and the returned tensor is the stream WITHOUT it.
TensorPtr GPT2Model::forward_block(const TensorPtr& x_in, int i, const TensorPtr& carry, TensorPtr* carry_out) {
    GPT2BlockWeights& w = blocks_[i];
    if (sharder_) {
        const std::string p = "blocks." + std::to_string(i) + ".";
        for (const char* k : {"ln_1.weight", "ln_1.bias", "attn.qkv.weight", "attn.qkv.bias", "attn.proj.weight", "attn.proj.bias",
                              "ln_2.weight", "ln_2.bias", "mlp.fc_in.weight", "mlp.fc_in.bias", "mlp.fc_out.weight", "mlp.fc_out.bias"})
            sharder_->require(p + k);
    }
    const float eps = config_.layernorm_eps;
    // MFT_FUSE_ADD_LN=0 keeps residual adds and LayerNorms as separate kernels (default: fused, bit-identical values)
    static const bool fuse_ln = [] { const char* e = std::getenv("MFT_FUSE_ADD_LN"); return !(e && e[0] == '0'); }();
    TensorPtr x = x_in, x1;
    if (carry && fuse_ln) x1 = add_layer_norm(x_in, carry, w.ln_1_weight, w.ln_1_bias, eps, &x);
    else {
        if (carry) x = add(x_in, carry);
        x1 = layer_norm(x, w.ln_1_weight, w.ln_1_bias, eps);
    }
    // bf16 operand mode: results that are only ever read as bf16 operands skip their fp32 copy (the values stay reachable:
    // an fp32 reader triggers one upcast). qkv feeds the tcgen05 attention only; the GELU activation feeds fc_out's GEMM
    // only when fc_out has no adapter and a frozen weight; ln_2's rows feed fc_in's GEMM only under the same condition.
    const bool lowp = internal::gemm_precision() == 1 && internal::bf16_only_results();
    const int64_t head_dim = config_.n_embd / config_.n_head;
    LinearFusion qkv_f;
    qkv_f.out_bf16_only = lowp && head_dim == 64 && !(mft_attention_mode() & 1) && !internal::true_gradients();
    auto qkv = run_linear_fused(x1, w.qkv_lin, w.lora_initialized, w.attn_qkv_weight, w.attn_qkv_bias, qkv_f);
    const bool fc_in_plain = !(w.lora_initialized && w.fc_in_lin && w.fc_in_lin->has_adapter()) && !internal::wants_grad(w.mlp_fc_in_weight) &&
                             !internal::wants_grad(w.mlp_fc_in_bias);
    const bool fc_out_plain = !(w.lora_initialized && w.fc_out_lin && w.fc_out_lin->has_adapter()) && !internal::wants_grad(w.mlp_fc_out_weight);
    // default: the reference's stop-gradient attention; true-gradient mode: the same forward with a backward
    auto ctx = internal::true_gradients() ? gpt2_causal_attention_with_grad(qkv, config_.n_head) : gpt2_causal_attention(qkv, config_.n_head);
    // MFT_FUSE_EPILOGUE bit 0: residual adds ride in the projections' epilogues; bit 1: GeluBackward rides in the epilogue of
    // fc_out's dX product. Off by default: with four epilogue warps per CTA the extra operand traffic and the tanh cost more
    // inside the GEMM than the separate streaming kernels do (profiles/README.md).
    static const int fuse = [] { const char* e = std::getenv("MFT_FUSE_EPILOGUE"); return e ? std::atoi(e) : 0; }();
    TensorPtr x2, x3;
    if (fuse & 1) {
        LinearFusion attn_out;
        attn_out.residual = x;
        x2 = run_linear_fused(ctx, w.proj_lin, w.lora_initialized, w.attn_proj_weight, w.attn_proj_bias, attn_out);
        x3 = layer_norm(x2, w.ln_2_weight, w.ln_2_bias, eps);
    } else if (fuse_ln) {
        x3 = add_layer_norm(x, run_linear(ctx, w.proj_lin, w.lora_initialized, w.attn_proj_weight, w.attn_proj_bias), w.ln_2_weight, w.ln_2_bias, eps, &x2,
                            lowp && fc_in_plain);
    } else {
        x2 = add(x, run_linear(ctx, w.proj_lin, w.lora_initialized, w.attn_proj_weight, w.attn_proj_bias));
        x3 = layer_norm(x2, w.ln_2_weight, w.ln_2_bias, eps);
    }
    TensorPtr h, pre;   // fc_in + tanh-GELU: the activation is written by the GEMM epilogue (graph/gpt2_model.cpp:794-802)
    {
        LinearFusion f;
        f.gelu_act = &h;
        f.gelu_bf16_only = lowp && fc_out_plain && !fuse;
        // the pre-activation is read by GeluBackward only (which takes bf16): MFT_BF16_PREACT=0 keeps it in fp32
        static const bool preact16 = [] { const char* e = std::getenv("MFT_BF16_PREACT"); return !(e && e[0] == '0'); }();
        f.out_bf16_only = f.gelu_bf16_only && preact16;
        if (w.lora_initialized && w.fc_in_lin) pre = w.fc_in_lin->forward_fused(x3, f);
        else pre = fused_lora_linear(x3, w.mlp_fc_in_weight, w.mlp_fc_in_bias, {}, f);
    }
    // x2 + fc_out(gelu(pre)): residual in the forward epilogue, GeluBackward in the epilogue of the backward dX product
    if (!fuse) {
        auto m = run_linear(h, w.fc_out_lin, w.lora_initialized, w.mlp_fc_out_weight, w.mlp_fc_out_bias);
        if (carry_out) { *carry_out = m; return x2; }      // the next LayerNorm adds it
        return add(x2, m);
    }
    if (carry_out) *carry_out = nullptr;
    LinearFusion mlp_out;
    if (fuse & 1) mlp_out.residual = x2;
    if (fuse & 2) mlp_out.gelu_input_pre = pre;
    auto m = run_linear_fused(h, w.fc_out_lin, w.lora_initialized, w.mlp_fc_out_weight, w.mlp_fc_out_bias, mlp_out);
    return (fuse & 1) ? m : add(x2, m);
}

// graph/gpt2_model.cpp:323-442
TensorPtr GPT2Model::forward(const TensorPtr& input_ids, const TensorPtr& attention_mask) {
    (void)attention_mask;   // the reference's default attention path ignores the padding mask too (Appendix B11)
    if (!input_ids || input_ids->ndim() != 2) throw std::runtime_error("GPT2Model::forward expects input_ids [batch, seq_len]");
    const int64_t S = input_ids->shape()[1];
    if (S > config_.n_positions) throw std::runtime_error("seq_len exceeds n_positions");
    if (sharder_) { sharder_->require("wte.weight"); sharder_->require("wpe.weight"); }
    TensorPtr x = embedding_lookup(input_ids, wte_weight_, wpe_weight_);     // not differentiable (:346-381)
    TensorPtr carry;     // MLP output of the previous block, added inside the next LayerNorm kernel
    for (int i = 0; i < config_.n_layer; ++i) {
        TensorPtr next;
        x = forward_block(x, i, carry, &next);
        carry = next;
    }
    if (carry) {
        static const bool fuse_ln = [] { const char* e = std::getenv("MFT_FUSE_ADD_LN"); return !(e && e[0] == '0'); }();
        TensorPtr sum;
        if (fuse_ln) x = add_layer_norm(x, carry, ln_f_weight_, ln_f_bias_, config_.layernorm_eps, &sum);
        else x = layer_norm(add(x, carry), ln_f_weight_, ln_f_bias_, config_.layernorm_eps);
    } else x = layer_norm(x, ln_f_weight_, ln_f_bias_, config_.layernorm_eps);
    return lm_head_tied(x, wte_weight_);                                     // wte only gets a gradient through here
}

}  // namespace ops
