// GPT-2 LoRA injection — API mirror of the reference's graph/lora_injector.h:19-191. inject() reproduces the
// reference's deterministic initialisation bit for bit (A ~ U(-b,b), b = sqrt(6/(in+r)), mt19937(42+in+out)
// re-seeded per matrix; B = 0; graph/lora_injector.cpp:70-85) so both builds start from identical adapters.
#pragma once
#include <string>
#include <unordered_map>
#include <vector>

#include "gpt2_model.h"
#include "tensor.h"

namespace ops {

enum class LoraTarget { AttnQKV, AttnProj, MlpFcIn, MlpFcOut };

struct LoraSpec {
    int rank = 8;
    float alpha = 16.0f;
    float dropout = 0.05f;
    bool split_qkv = true;
    std::vector<LoraTarget> targets = {LoraTarget::AttnQKV, LoraTarget::AttnProj};
    std::vector<int> layers;   // empty = all layers
    LoraSpec() = default;
    static LoraSpec default_config() { return LoraSpec(); }
    static LoraSpec full_config() {
        LoraSpec s;
        s.targets = {LoraTarget::AttnQKV, LoraTarget::AttnProj, LoraTarget::MlpFcIn, LoraTarget::MlpFcOut};
        return s;
    }
};

struct LoraState {
    TensorPtr A;   // [in, r]
    TensorPtr B;   // [r, out]
    float scale = 1.0f;
    float dropout_p = 0.0f;
    bool enabled = true;
    void init(int64_t in_features, int64_t out_features, int rank, float alpha, float dropout);
};

class LoraInjector {
public:
    LoraInjector() = default;
    ~LoraInjector() = default;

    void inject(GPT2Model& model, const LoraSpec& spec);
    void merge();
    void unmerge();
    void merge_all(GPT2Model& model);
    void unmerge_all(GPT2Model& model);
    std::vector<TensorPtr> collect_lora_parameters() const;
    void save_lora_safetensors(const std::string& path) const;
    void load_lora_safetensors(const std::string& path);
    void print_info() const;
    std::vector<TensorPtr> get_trainable_params();
    static TensorPtr lora_linear_forward(const TensorPtr& x, const TensorPtr& W, const TensorPtr& bias,
                                         const LoraState* lora, bool training = false);

private:
    GPT2Model* model_ = nullptr;
    LoraSpec spec_;
    bool merged_ = false;
    int num_layers_ = 0;
};

}  // namespace ops
