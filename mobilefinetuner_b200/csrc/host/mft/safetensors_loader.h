// safetensors reading / writing — API mirror of the reference's graph/safetensors_loader.h:54-124 (SafeTensorsReader,
// SafeTensorsLoadOptions, GPT2KeyMapper, GemmaKeyMapper) plus the writer the savers share.
// File format: u64 little-endian header length, JSON header
//   { "<name>": {"dtype": "F32"|"F16"|"BF16"|"I32"|"I64", "shape": [...], "data_offsets": [begin, end]}, ...,
//     "__metadata__": {"k": "v", ...} }
// then the raw little-endian data. Tensors are returned as fp32 host tensors (F16/BF16 promoted); they move to HBM on
// first use by a kernel like every other ops::Tensor.
#pragma once
#include <cstdint>
#include <fstream>
#include <map>
#include <memory>
#include <string>
#include <unordered_map>
#include <vector>

#include "tensor.h"

namespace ops {

struct SafeTensorInfo {
    std::string dtype;
    std::vector<int64_t> shape;
    std::vector<size_t> data_offsets;   // [begin, end) relative to the start of the data section
};

struct SafeTensorsLoadOptions {
    bool transpose_linear = true;    // 2-D "*weight" tensors that are not norms / embeddings: [out,in] -> [in,out]
    bool auto_promote_fp16 = true;
    bool verbose = true;
    bool strict_shape_check = true;
};

class SafeTensorsReader {
public:
    explicit SafeTensorsReader(const std::string& filepath);
    ~SafeTensorsReader();
    void parse_header();
    std::vector<std::string> get_tensor_names() const;
    SafeTensorInfo get_tensor_info(const std::string& name) const;
    TensorPtr load_tensor(const std::string& name, bool transpose = false);
    // key_mapping: internal name -> name in the file ("transformer."-prefixed variant accepted); missing keys are skipped
    std::unordered_map<std::string, TensorPtr> load_tensors_mapped(const std::unordered_map<std::string, std::string>& key_mapping,
                                                                   const SafeTensorsLoadOptions& options = SafeTensorsLoadOptions());
    // extension: the "__metadata__" string map of the header
    const std::map<std::string, std::string>& metadata() const { return metadata_; }

private:
    TensorPtr read_tensor_data(const SafeTensorInfo& info, bool transpose);
    std::string filepath_;
    std::ifstream file_;
    size_t header_len_ = 0;
    size_t data_offset_ = 0;
    bool parsed_ = false;
    std::unordered_map<std::string, SafeTensorInfo> tensor_map_;
    std::vector<std::string> order_;
    std::map<std::string, std::string> metadata_;
};

class GPT2KeyMapper {
public:
    static std::unordered_map<std::string, std::string> generate_gpt2_mapping(int num_layers = 12);
    static void print_mapping(const std::unordered_map<std::string, std::string>& mapping);
};

class GemmaKeyMapper {
public:
    static std::unordered_map<std::string, std::string> generate_gemma_mapping(int num_layers = 18);
    static void print_mapping(const std::unordered_map<std::string, std::string>& mapping);
};

// ---- writer shared by LoraSaver, the Gemma injector and the checkpoint helpers (no counterpart class in the reference,
// which open-codes the same layout in graph/lora_saver.cpp:214-276 and gpt2_full_finetune/main.cpp:156-237).
// Tensors are written as F32 in sorted key order; device tensors are read back through their host mirror.
void write_safetensors(const std::string& path, const std::map<std::string, TensorPtr>& tensors,
                       const std::vector<std::pair<std::string, std::string>>& metadata = {});

}  // namespace ops
