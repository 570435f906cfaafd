// ParameterSharder — API mirror of the reference's opt_ops/sharding/parameter_sharder.h:36-93.
// OUT OF SCOPE for the B200 hot path (BASELINE.json north_star: weights stay HBM-resident; 180 GB per GPU):
// the interface is kept so the models and trainers compile, but nothing is ever written to disk — every
// registered tensor simply stays resident and require() hands it back.
#pragma once
#include <cstdint>
#include <mutex>
#include <string>
#include <unordered_map>
#include <vector>

#include "tensor.h"

namespace ops {
namespace sharding {

struct ShardConfig {
    std::string offload_dir;
    size_t max_resident_bytes = 256 * 1024 * 1024;
    bool quantize_fp16_on_disk = true;
};

enum class ShardState { InMemory, Offloaded };

class ParameterSharder {
public:
    explicit ParameterSharder(const ShardConfig& cfg) : cfg_(cfg) {}
    void register_parameter(const std::string& name, const TensorPtr& tensor, bool keep_in_memory = true,
                            TensorPtr* owner_ptr = nullptr) {
        (void)keep_in_memory; (void)owner_ptr;
        std::lock_guard<std::mutex> lk(mutex_);
        entries_[name] = tensor;
    }
    TensorPtr require(const std::string& name) {
        std::lock_guard<std::mutex> lk(mutex_);
        auto it = entries_.find(name);
        return it == entries_.end() ? nullptr : it->second;
    }
    void mark_dirty(const std::string&) {}
    void offload_all() {}
    std::string debug_string() const { return "ParameterSharder(HBM-resident, " + std::to_string(entries_.size()) + " tensors)"; }

private:
    ShardConfig cfg_;
    mutable std::mutex mutex_;
    std::unordered_map<std::string, TensorPtr> entries_;
};

}  // namespace sharding
}  // namespace ops
