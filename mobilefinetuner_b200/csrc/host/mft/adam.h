// Adam — API mirror of the reference's optim/adam.h:23-104. Semantics preserved:
// L2-coupled weight decay, per-tensor step counters that only advance for tensors with a non-null gradient,
// bias correction through float pow. The update of ALL tensors is one multi-tensor CUDA launch.
#pragma once
#include <unordered_map>

#include "optimizer.h"

namespace ops {

struct AdamConfig : public OptimizerConfig {
    float beta1, beta2, epsilon;
    AdamConfig() : OptimizerConfig(), beta1(0.9f), beta2(0.999f), epsilon(1e-8f) {}
    AdamConfig(float lr, float b1 = 0.9f, float b2 = 0.999f, float eps = 1e-8f, float wd = 0.0f, float clip = 1.0f, bool ams = false)
        : OptimizerConfig(lr, wd, clip, ams), beta1(b1), beta2(b2), epsilon(eps) {}
};

struct AdamState : public OptimizerState {
    size_t step = 0;
    std::vector<TensorPtr> m, v, v_hat;
    void to_file(const std::string& path) const override;
    void from_file(const std::string& path) override;
};

class Adam : public Optimizer {
public:
    explicit Adam(const AdamConfig& config);
    ~Adam() override;

    void step(const std::vector<TensorPtr>& parameters, const std::vector<TensorPtr>& gradients) override;
    // Optimizer-state checkpoint. The reference writes an unordered dump and its load_state reads the
    // header only (optim/adam.cpp:92-164); here the state is a safetensors file keyed by the POSITION of each tensor in
    // the parameter list of the last step(): "adam.<i>.m", "adam.<i>.v" (F32) and metadata "adam.<i>.step", "num_params",
    // "lr", "beta1", "beta2", "eps", "weight_decay". load_state() keeps the loaded moments until the next step() (or
    // bind_state()) sees the parameter list and attaches entry i to parameters[i] (sizes must match).
    void save_state(const std::string& path) const override;
    void load_state(const std::string& path) override;
    void bind_state(const std::vector<TensorPtr>& parameters);
    void set_learning_rate(float new_lr) override { adam_config_.learning_rate = new_lr; }
    float get_learning_rate() const override { return adam_config_.learning_rate; }

    // B200 extension: the same update with the clip factor read from a device scalar (never visits the host).
    void step_scaled(const std::vector<TensorPtr>& parameters, const std::vector<TensorPtr>& gradients, const float* dev_grad_scale);
    // ... and with raw device gradient pointers (nullptr = no gradient), e.g. slices of an all-reduced bucket.
    void step_device(const std::vector<TensorPtr>& parameters, const std::vector<float*>& grad_ptrs, const float* dev_grad_scale);

protected:
    void update_param(TensorPtr param, const TensorPtr& grad, OptimizerState& state) override;

private:
    AdamConfig adam_config_;
    std::unordered_map<TensorPtr, AdamState> states_;
    std::vector<TensorPtr> order_;                 // parameter list of the last step (checkpoint order)
    std::vector<AdamState> loaded_;                // states read by load_state, not yet attached to tensors
    void* dev_slots_ = nullptr;
    size_t dev_slots_bytes_ = 0;
    void init_state(const TensorPtr& param);
    float compute_bias_correction1(size_t step) const;
    float compute_bias_correction2(size_t step) const;
};

// Device slot table shared by Adam and the clip helpers.
namespace optim_detail {
struct SlotTable { void* dev = nullptr; int count = 0; };
// Builds (and uploads) a table with one slot per parameter; grads[i] == nullptr marks "no gradient".
SlotTable upload_slots(const std::vector<TensorPtr>& params, const std::vector<float*>& grads,
                       const std::vector<TensorPtr>* m, const std::vector<TensorPtr>* v,
                       const std::vector<float>* bc1, const std::vector<float>* bc2, void** dev_buf, size_t* dev_bytes);
// sumsq of all non-null grads into a persistent device double; returns its device address.
double* device_grad_sumsq(const SlotTable& t);
}  // namespace optim_detail

}  // namespace ops
