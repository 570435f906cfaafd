// LoRA adapter save / resume for GPT-2 — API mirror of the reference's graph/lora_saver.h:17-48.
// Wire format: safetensors, keys `layer.{i}.{attn.q|attn.k|attn.v|attn.qkv|attn.proj|mlp.fc_in|mlp.fc_out}.lora_{A|B}`,
// fp32, sorted; `__metadata__` carries rank / alpha / dropout / split_qkv / targets as strings.
// Adapter tensors live in HBM while training; saving reads them through the tensors' host mirrors (one D2H each).
#pragma once
#include <string>
#include <unordered_map>
#include <vector>

#include "gpt2_model.h"
#include "lora_injector.h"
#include "tensor.h"

namespace ops {

struct LoRAState {
    int rank = 8;
    float alpha = 16.0f;
    float dropout = 0.0f;
    bool split_qkv = true;
    std::vector<LoraTarget> targets;
    std::unordered_map<std::string, TensorPtr> tensors;
    bool compatible_with(const GPT2Model& model) const;
};

class LoraSaver {
public:
    static void save_safetensors(const std::string& path, const GPT2Model& model, const LoraSpec& spec);
    static LoRAState load_safetensors(const std::string& path);
    static void attach_from_state(GPT2Model& model, const LoRAState& state);
    static std::string make_peft_key(int layer, const std::string& target, const std::string& ab);
    static bool parse_peft_key(const std::string& key, int& layer, std::string& target, std::string& ab);
};

}  // namespace ops
