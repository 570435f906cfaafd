#include "gemma_lora_injector.h"

#include <cmath>
#include <iostream>
#include <random>
#include <stdexcept>

#include "ops.h"
#include "runtime.h"

namespace ops {

namespace {
struct Target { LoRALinear* lin; TensorPtr W; bool hidden_is_input; };

bool lookup(GemmaBlockWeights& b, const std::string& name, Target& t) {
    if (name == "q_proj") t = {b.q_proj_lora.get(), b.q_proj_weight, true};
    else if (name == "k_proj") t = {b.k_proj_lora.get(), b.k_proj_weight, true};
    else if (name == "v_proj") t = {b.v_proj_lora.get(), b.v_proj_weight, true};
    else if (name == "o_proj") t = {b.o_proj_lora.get(), b.o_proj_weight, false};
    else if (name == "gate_proj") t = {b.gate_proj_lora.get(), b.gate_proj_weight, true};
    else if (name == "up_proj") t = {b.up_proj_lora.get(), b.up_proj_weight, true};
    else if (name == "down_proj") t = {b.down_proj_lora.get(), b.down_proj_weight, false};
    else return false;
    return t.lin && t.W;
}

// graph/gemma_lora_injector.cpp:30-46 — the generator is shared by every call (static), so the init depends on
// injection order: layer-major, module order as listed in the spec.
std::mt19937& init_rng() {
    static std::mt19937 rng(42);
    return rng;
}
std::pair<TensorPtr, TensorPtr> make_adapter(int64_t in_dim, int64_t out_dim, int rank) {
    std::mt19937& rng = init_rng();
    mft_rt::PersistentScope keep;
    const float bound = 1.0f / std::sqrt(static_cast<float>(in_dim));
    std::uniform_real_distribution<float> dist(-bound, bound);
    auto A = zeros({rank, in_dim}, kFloat32, kCPU);
    float* a = A->data<float>();
    for (int64_t i = 0; i < A->numel(); ++i) a[i] = dist(rng);
    auto B = zeros({out_dim, rank}, kFloat32, kCPU);
    return {A, B};
}
}  // namespace

void GemmaLoraInjector::reseed_init(unsigned seed) { init_rng().seed(seed); }

// graph/gemma_lora_injector.cpp:50-109
void GemmaLoraInjector::inject(GemmaModel& model, const GemmaLoraSpec& spec) {
    model_ = &model;
    spec_ = spec;
    num_layers_ = model.config().num_hidden_layers;
    model.init_lora_modules();
    attached_modules_ = 0;
    std::vector<int> layers = spec.layers;
    if (layers.empty()) for (int i = 0; i < num_layers_; ++i) layers.push_back(i);
    const float scale = spec.alpha / static_cast<float>(spec.rank);
    const int64_t hidden = model.config().hidden_size;
    for (int layer : layers) {
        GemmaBlockWeights& b = model.get_block(layer);
        for (const auto& name : spec.target_modules) {
            Target t;
            if (!lookup(b, name, t) || t.W->ndim() != 2) {
                std::cerr << "[GemmaLoraInjector] Skip module " << name << " in layer " << layer << std::endl;
                continue;
            }
            const int64_t d0 = t.W->shape()[0], d1 = t.W->shape()[1];
            int64_t in_dim = d1, out_dim = d0;
            if (t.hidden_is_input) { if (d0 == hidden) { in_dim = d0; out_dim = d1; } else if (d1 == hidden) { in_dim = d1; out_dim = d0; } }
            else { if (d0 == hidden) { out_dim = d0; in_dim = d1; } else if (d1 == hidden) { out_dim = d1; in_dim = d0; } }
            auto ab = make_adapter(in_dim, out_dim, spec.rank);
            t.lin->attach_lora(ab.first, ab.second, scale, 0, int(out_dim));
            t.lin->set_debug_name("layers_" + std::to_string(layer) + "_self_attn_" + name);
            ++attached_modules_;
        }
    }
    std::cout << "[GemmaLoraInjector] Injected " << attached_modules_ << " LoRA modules across " << layers.size() << " layers." << std::endl;
}

std::vector<TensorPtr> GemmaLoraInjector::get_trainable_params() const {
    return model_ ? model_->get_lora_parameters() : std::vector<TensorPtr>{};
}
void GemmaLoraInjector::print_info() const {
    std::cout << "[GemmaLoRA]\n  Rank: " << spec_.rank << "\n  Alpha: " << spec_.alpha << "\n  Dropout: " << spec_.dropout << "\n  Targets: ";
    for (size_t i = 0; i < spec_.target_modules.size(); ++i) std::cout << (i ? ", " : "") << spec_.target_modules[i];
    std::cout << std::endl;
}
void GemmaLoraInjector::merge_all(GemmaModel& model) { model.merge_lora(); }
void GemmaLoraInjector::unmerge_all(GemmaModel& model) { model.unmerge_lora(); }
void GemmaLoraInjector::save_lora_safetensors(const std::string&) const { throw std::runtime_error("GemmaLoraInjector::save_lora_safetensors: checkpoint I/O is not part of this build"); }
void GemmaLoraInjector::load_lora_safetensors(const std::string&) { throw std::runtime_error("GemmaLoraInjector::load_lora_safetensors: checkpoint I/O is not part of this build"); }

}  // namespace ops
