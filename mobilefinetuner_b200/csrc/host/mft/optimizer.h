// Optimizer base — API mirror of the reference's optim/optimizer.h:25-106.
#pragma once
#include <memory>
#include <string>
#include <vector>

#include "tensor.h"

namespace ops {

struct OptimizerConfig {
    float learning_rate;
    float weight_decay;
    float clip_grad_norm;
    bool amsgrad;
    OptimizerConfig() : learning_rate(0.001f), weight_decay(0.0f), clip_grad_norm(1.0f), amsgrad(false) {}
    OptimizerConfig(float lr, float wd = 0.0f, float clip = 1.0f, bool ams = false)
        : learning_rate(lr), weight_decay(wd), clip_grad_norm(clip), amsgrad(ams) {}
    virtual ~OptimizerConfig() = default;
};

struct OptimizerState {
    virtual ~OptimizerState() = default;
    virtual void to_file(const std::string& path) const = 0;
    virtual void from_file(const std::string& path) = 0;
};

class Optimizer {
public:
    explicit Optimizer(const OptimizerConfig& config) : config_(config) {}
    virtual ~Optimizer() = default;

    virtual void step(const std::vector<TensorPtr>& parameters, const std::vector<TensorPtr>& gradients) = 0;
    virtual void zero_grad(const std::vector<TensorPtr>& parameters) {
        for (auto& p : parameters) p->zero_grad();
    }
    // Global L2 clip over the non-null gradients, entirely on the device (optim/optimizer.cpp:6-38; scale = max/norm).
    virtual void clip_grad_norm(const std::vector<TensorPtr>& parameters, float max_norm);
    // B200 extension: the trainers' own clip rule, scale = max/(norm + eps) when norm > max (gpt2_lora_finetune/main.cpp:491-516,
    // optim/gemma_trainer.cpp:84-102), on the device. Returns the norm.
    float clip_grad_norm_eps(const std::vector<TensorPtr>& parameters, float max_norm, float eps);
    virtual void save_state(const std::string& path) const = 0;
    virtual void load_state(const std::string& path) = 0;
    virtual float get_learning_rate() const { return config_.learning_rate; }
    virtual void set_learning_rate(float lr) { config_.learning_rate = lr; }
    const OptimizerConfig& get_config() const { return config_; }

protected:
    OptimizerConfig config_;
    virtual void update_param(TensorPtr param, const TensorPtr& grad, OptimizerState& state) = 0;
    virtual float compute_grad_norm(const std::vector<TensorPtr>& parameters);
    virtual void scale_gradients(const std::vector<TensorPtr>& parameters, float scale);
};

}  // namespace ops
