#include "safetensors_loader.h"
#include <sstream>
#include <map>
#include <cstring>
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
// graph/gemma_lora_injector.cpp:129-212: attention adapters only (q, k, v, o), keys layer.{i}.attn.{q|k|v|proj}[.{slice}].lora_{A|B},
// tensors in the layout they are held in (PEFT [r,in] / [out,r]); MLP adapters are not written by the reference either.
void GemmaLoraInjector::save_lora_safetensors(const std::string& path) const {
    if (!model_) { std::cerr << "[GemmaLoraInjector] save_lora_safetensors: model not set" << std::endl; return; }
    std::map<std::string, TensorPtr> state;
    auto add = [&](int layer, const char* target, const LoRALinear* lin) {
        if (!lin) return;
        const auto& sl = lin->slices();
        for (size_t k = 0; k < sl.size(); ++k) {
            if (!sl[k].A || !sl[k].B) continue;
            std::string base = "layer." + std::to_string(layer) + "." + target;
            if (sl.size() > 1) base += "." + std::to_string(k);
            state[base + ".lora_A"] = sl[k].A;
            state[base + ".lora_B"] = sl[k].B;
        }
    };
    for (int i = 0; i < model_->config().num_hidden_layers; ++i) {
        const GemmaBlockWeights& b = model_->get_block(i);
        add(i, "attn.q", b.q_proj_lora.get());
        add(i, "attn.k", b.k_proj_lora.get());
        add(i, "attn.v", b.v_proj_lora.get());
        add(i, "attn.proj", b.o_proj_lora.get());
    }
    std::ostringstream alpha, dropout;
    alpha << spec_.alpha;
    dropout << spec_.dropout;
    write_safetensors(path, state, {{"rank", std::to_string(spec_.rank)}, {"alpha", alpha.str()}, {"dropout", dropout.str()},
                                    {"targets", "attn.q,attn.k,attn.v,attn.proj"}});
    std::cout << "[GemmaLoraInjector] Saved " << state.size() << " LoRA tensors to " << path << std::endl;
}

// Extension (the reference prints "not implemented yet", graph/gemma_lora_injector.cpp:214-216): restores the factors written
// by save_lora_safetensors into the adapters already injected on the model, so a run can resume.
void GemmaLoraInjector::load_lora_safetensors(const std::string& path) {
    if (!model_) throw std::runtime_error("GemmaLoraInjector::load_lora_safetensors: inject() first");
    SafeTensorsReader reader(path);
    reader.parse_header();
    int restored = 0;
    auto take = [&](int layer, const char* target, LoRALinear* lin) {
        if (!lin) return;
        const auto& sl = lin->slices();
        for (size_t k = 0; k < sl.size(); ++k) {
            std::string base = "layer." + std::to_string(layer) + "." + target;
            if (sl.size() > 1) base += "." + std::to_string(k);
            for (const char* ab : {"A", "B"}) {
                const std::string key = base + ".lora_" + ab;
                TensorPtr dst = ab[0] == 'A' ? sl[k].A : sl[k].B;
                TensorPtr src;
                try { src = reader.load_tensor(key, false); } catch (const std::exception&) { continue; }
                if (src->shape() != dst->shape()) throw std::runtime_error("GemmaLoraInjector::load_lora_safetensors: shape mismatch for " + key);
                std::memcpy(dst->data<float>(), static_cast<const Tensor&>(*src).data<float>(), size_t(dst->numel()) * sizeof(float));
                ++restored;
            }
        }
    };
    for (int i = 0; i < model_->config().num_hidden_layers; ++i) {
        GemmaBlockWeights& b = model_->get_block(i);
        take(i, "attn.q", b.q_proj_lora.get());
        take(i, "attn.k", b.k_proj_lora.get());
        take(i, "attn.v", b.v_proj_lora.get());
        take(i, "attn.proj", b.o_proj_lora.get());
    }
    std::cout << "[GemmaLoraInjector] Restored " << restored << " LoRA tensors from " << path << std::endl;
}

}  // namespace ops
