// Gemma LoRA injection — API mirror of the reference's graph/gemma_lora_injector.h:9-56 (PEFT layout:
// A [r,in] ~ U(-1/sqrt(in), 1/sqrt(in)) from ONE function-static mt19937(42) consumed in injection order, B [out,r] = 0).
#pragma once
#include <string>
#include <vector>

#include "gemma_model.h"

namespace ops {

struct GemmaLoraSpec {
    int rank = 8;
    float alpha = 32.0f;
    float dropout = 0.1f;
    std::vector<std::string> target_modules = {"q_proj", "k_proj", "v_proj", "o_proj", "gate_proj", "up_proj", "down_proj"};
    std::vector<int> layers;
    static GemmaLoraSpec full_attn_mlp() { return GemmaLoraSpec(); }
    static GemmaLoraSpec attention_only() { GemmaLoraSpec s; s.target_modules = {"q_proj", "k_proj", "v_proj", "o_proj"}; return s; }
    static GemmaLoraSpec attention_light() { GemmaLoraSpec s; s.target_modules = {"q_proj", "v_proj"}; return s; }
};

class GemmaLoraInjector {
public:
    GemmaLoraInjector() = default;
    ~GemmaLoraInjector() = default;
    void inject(GemmaModel& model, const GemmaLoraSpec& spec);
    // Extension: the reference's adapter init draws from one process-wide generator seeded with 42, so a second model built
    // in the same process starts from different factors. Reseeding restores the first-model state (used by the trainer C ABI).
    static void reseed_init(unsigned seed = 42);
    std::vector<TensorPtr> get_trainable_params() const;
    void print_info() const;
    void save_lora_safetensors(const std::string& path) const;
    void load_lora_safetensors(const std::string& path);
    void merge_all(GemmaModel& model);
    void unmerge_all(GemmaModel& model);

private:
    GemmaModel* model_ = nullptr;
    GemmaLoraSpec spec_;
    int num_layers_ = 0;
    int attached_modules_ = 0;
};

}  // namespace ops
