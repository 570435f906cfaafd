#include "lora_injector.h"

#include <cmath>
#include <iostream>
#include <random>
#include <stdexcept>

#include "ops.h"
#include "runtime.h"

namespace ops {

// graph/lora_injector.cpp:14-38 — stand-alone adapter state with a non-deterministic normal init
void LoraState::init(int64_t in_features, int64_t out_features, int rank, float alpha, float dropout) {
    mft_rt::PersistentScope keep;
    scale = alpha / static_cast<float>(rank);
    dropout_p = dropout;
    enabled = true;
    A = std::make_shared<Tensor>(std::vector<int64_t>{in_features, rank}, kFloat32, kCPU);
    std::random_device rd;
    std::mt19937 gen(rd());
    std::normal_distribution<float> dist(0.0f, 1.0f / std::sqrt(static_cast<float>(rank)));
    float* a = A->data<float>();
    for (int64_t i = 0; i < A->numel(); ++i) a[i] = dist(gen);
    B = zeros({rank, out_features}, kFloat32, kCPU);
    A->set_requires_grad(true);
    B->set_requires_grad(true);
}

namespace {
// graph/lora_injector.cpp:70-85
std::pair<TensorPtr, TensorPtr> make_adapter(int in_dim, int out_dim, int rank) {
    mft_rt::PersistentScope keep;   // adapters are parameters
    auto A = zeros({in_dim, rank}, kFloat32, kCPU);
    float* a = A->data<float>();
    const float bound = std::sqrt(6.0f / (in_dim + rank));
    std::mt19937 gen(42 + in_dim + out_dim);
    std::uniform_real_distribution<float> dist(-bound, bound);
    for (int64_t i = 0; i < A->numel(); ++i) a[i] = dist(gen);
    auto B = zeros({rank, out_dim}, kFloat32, kCPU);
    return {A, B};
}
}  // namespace

// graph/lora_injector.cpp:48-148
void LoraInjector::inject(GPT2Model& model, const LoraSpec& spec) {
    spec_ = spec;
    model_ = &model;
    num_layers_ = model.config().n_layer;
    const int C = model.config().n_embd, r = spec.rank;
    const float scale = spec.alpha / float(r);
    model.init_lora_modules();
    std::vector<int> layers = spec.layers;
    if (layers.empty()) for (int i = 0; i < num_layers_; ++i) layers.push_back(i);
    int count = 0;
    for (int i : layers) {
        if (i < 0 || i >= num_layers_) throw std::out_of_range("LoraInjector: layer index");
        GPT2BlockWeights& b = model.get_block(i);
        for (LoraTarget t : spec.targets) {
            switch (t) {
                case LoraTarget::AttnQKV:
                    if (spec.split_qkv) {
                        for (int part = 0; part < 3; ++part) {
                            auto ab = make_adapter(C, C, r);
                            b.qkv_lin->attach_lora(ab.first, ab.second, scale, part * C, C);
                            ++count;
                        }
                    } else {
                        auto ab = make_adapter(C, 3 * C, r);
                        b.qkv_lin->attach_lora(ab.first, ab.second, scale, 0, 3 * C);
                        ++count;
                    }
                    break;
                case LoraTarget::AttnProj: { auto ab = make_adapter(C, C, r); b.proj_lin->attach_lora(ab.first, ab.second, scale, 0, C); ++count; break; }
                case LoraTarget::MlpFcIn: { auto ab = make_adapter(C, 4 * C, r); b.fc_in_lin->attach_lora(ab.first, ab.second, scale, 0, 4 * C); ++count; break; }
                case LoraTarget::MlpFcOut: { auto ab = make_adapter(4 * C, C, r); b.fc_out_lin->attach_lora(ab.first, ab.second, scale, 0, C); ++count; break; }
            }
        }
    }
    std::cout << "[LoraInjector] Injected " << count << " LoRA modules to " << layers.size() << " layers" << std::endl;
    print_info();
}

void LoraInjector::print_info() const {
    if (!model_) return;
    auto ps = model_->get_lora_parameters();
    int64_t total = 0;
    for (const auto& p : ps) total += p->numel();
    std::cout << "[LoRA Info]\n  Rank: " << spec_.rank << "\n  Alpha: " << spec_.alpha << "\n  Scale: " << spec_.alpha / float(spec_.rank)
              << "\n  Total LoRA params: " << total << "\n  Trainable param count: " << ps.size() << std::endl;
}

std::vector<TensorPtr> LoraInjector::collect_lora_parameters() const {
    return model_ ? model_->get_lora_parameters() : std::vector<TensorPtr>{};
}
std::vector<TensorPtr> LoraInjector::get_trainable_params() { return collect_lora_parameters(); }

void LoraInjector::merge_all(GPT2Model& model) {
    for (int i = 0; i < model.config().n_layer; ++i) {
        auto& b = model.get_block(i);
        for (LoRALinear* l : {b.qkv_lin.get(), b.proj_lin.get(), b.fc_in_lin.get(), b.fc_out_lin.get()}) if (l) l->merge_to_base();
    }
    merged_ = true;
}
void LoraInjector::unmerge_all(GPT2Model& model) {
    for (int i = 0; i < model.config().n_layer; ++i) {
        auto& b = model.get_block(i);
        for (LoRALinear* l : {b.qkv_lin.get(), b.proj_lin.get(), b.fc_in_lin.get(), b.fc_out_lin.get()}) if (l) l->unmerge_from_base();
    }
    merged_ = false;
}
void LoraInjector::merge() { if (model_ && !merged_) merge_all(*model_); }
void LoraInjector::unmerge() { if (model_ && merged_) unmerge_all(*model_); }

// Adapter checkpoint I/O is the next row of the scope table, not part of the step.
void LoraInjector::save_lora_safetensors(const std::string&) const { throw std::runtime_error("LoraInjector::save_lora_safetensors: checkpoint I/O is not part of this build"); }
void LoraInjector::load_lora_safetensors(const std::string&) { throw std::runtime_error("LoraInjector::load_lora_safetensors: checkpoint I/O is not part of this build"); }

TensorPtr LoraInjector::lora_linear_forward(const TensorPtr& x, const TensorPtr& W, const TensorPtr& bias, const LoraState* lora, bool) {
    if (!lora || !lora->enabled) return linear(x, W, bias);
    return fused_lora_linear(x, W, bias, {LoraTerm{lora->A, lora->B, lora->scale}});
}

}  // namespace ops
