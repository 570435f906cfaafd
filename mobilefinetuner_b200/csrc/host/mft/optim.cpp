// Optimizer::clip_grad_norm and Adam on the device (reference: optim/optimizer.cpp:6-38, optim/adam.cpp:25-228).
#include <cmath>
#include <cstring>
#include <stdexcept>

#include <map>
#include "adam.h"
#include "safetensors_loader.h"
#include "mft_b200.h"
#include "optimizer.h"
#include "runtime.h"

namespace ops {

namespace optim_detail {

SlotTable upload_slots(const std::vector<TensorPtr>& params, const std::vector<float*>& grads,
                       const std::vector<TensorPtr>* m, const std::vector<TensorPtr>* v,
                       const std::vector<float>* bc1, const std::vector<float>* bc2, void** dev_buf, size_t* dev_bytes) {
    std::vector<mft_tensor_slot> host(params.size());
    for (size_t i = 0; i < params.size(); ++i) {
        mft_tensor_slot& s = host[i];
        std::memset(&s, 0, sizeof(s));
        s.n = params[i]->numel();
        s.param = m ? params[i]->dev_f32_mut() : const_cast<float*>(params[i]->dev_f32());
        s.grad = i < grads.size() ? grads[i] : nullptr;
        s.m = (m && (*m)[i]) ? (*m)[i]->dev_f32_mut() : nullptr;
        s.v = (v && (*v)[i]) ? (*v)[i]->dev_f32_mut() : nullptr;
        s.bias_corr1 = bc1 ? (*bc1)[i] : 1.0f;
        s.bias_corr2 = bc2 ? (*bc2)[i] : 1.0f;
    }
    const size_t bytes = host.size() * sizeof(mft_tensor_slot);
    if (bytes > *dev_bytes) {
        mft_rt::PersistentScope keep;
        if (*dev_buf) mft_rt::dev_free(*dev_buf, *dev_bytes);
        *dev_bytes = bytes * 2;
        *dev_buf = mft_rt::dev_alloc(*dev_bytes);
    }
    mft_rt::h2d(*dev_buf, host.data(), bytes);   // pageable source: consumed before the call returns
    return SlotTable{*dev_buf, int(host.size())};
}

double* device_grad_sumsq(const SlotTable& t) {
    static double* out = nullptr;
    if (!out) { mft_rt::PersistentScope keep; out = static_cast<double*>(mft_rt::dev_alloc(64)); }
    const size_t wsb = mft_grad_sumsq_workspace_bytes();
    mft_rt::check(mft_grad_sumsq(static_cast<const mft_tensor_slot*>(t.dev), t.count, out, mft_rt::workspace(wsb), wsb, mft_rt::stream()),
                  "mft_grad_sumsq");
    return out;
}

}  // namespace optim_detail

using optim_detail::SlotTable;

namespace {
void* g_clip_slots = nullptr;
size_t g_clip_slots_bytes = 0;
std::vector<float*> grads_of(const std::vector<TensorPtr>& params) {
    std::vector<float*> g(params.size(), nullptr);
    for (size_t i = 0; i < params.size(); ++i)
        if (params[i] && params[i]->grad()) g[i] = params[i]->grad()->dev_f32_mut();
    return g;
}
}  // namespace

float Optimizer::compute_grad_norm(const std::vector<TensorPtr>& parameters) {
    SlotTable t = optim_detail::upload_slots(parameters, grads_of(parameters), nullptr, nullptr, nullptr, nullptr, &g_clip_slots, &g_clip_slots_bytes);
    double* ss = optim_detail::device_grad_sumsq(t);
    double host = 0;
    mft_rt::d2h(&host, ss, sizeof(double));
    return float(std::sqrt(host));
}

void Optimizer::scale_gradients(const std::vector<TensorPtr>& parameters, float scale) {
    for (auto& p : parameters) {
        auto g = p->grad();
        if (!g) continue;
        mft_rt::check(mft_affine(g->dev_f32(), g->dev_f32_mut(), g->numel(), scale, 0.f, mft_rt::stream()), "mft_affine");
    }
}

// optim/optimizer.cpp:6-12
void Optimizer::clip_grad_norm(const std::vector<TensorPtr>& parameters, float max_norm) {
    const float n = compute_grad_norm(parameters);
    if (n > max_norm) scale_gradients(parameters, max_norm / n);
}

float Optimizer::clip_grad_norm_eps(const std::vector<TensorPtr>& parameters, float max_norm, float eps) {
    const float n = compute_grad_norm(parameters);
    if (n > max_norm) scale_gradients(parameters, max_norm / (n + eps));
    return n;
}

// ---------------------------------------------------------------------------- Adam
Adam::Adam(const AdamConfig& config) : Optimizer(config), adam_config_(config) {}
Adam::~Adam() {
    if (dev_slots_) mft_rt::dev_free(dev_slots_, dev_slots_bytes_);
}

float Adam::compute_bias_correction1(size_t step) const { return 1.0f - std::pow(adam_config_.beta1, static_cast<float>(step)); }
float Adam::compute_bias_correction2(size_t step) const { return 1.0f - std::pow(adam_config_.beta2, static_cast<float>(step)); }

void Adam::init_state(const TensorPtr& param) {      // optim/adam.cpp:171-220
    mft_rt::PersistentScope keep;                    // moments live as long as the optimizer
    AdamState st;
    st.step = 0;
    st.m.push_back(zeros(param->shape(), param->dtype(), param->device()));
    st.v.push_back(zeros(param->shape(), param->dtype(), param->device()));
    states_[param] = std::move(st);
}

void Adam::step(const std::vector<TensorPtr>& parameters, const std::vector<TensorPtr>& gradients) {
    step_scaled(parameters, gradients, nullptr);
}

void Adam::step_scaled(const std::vector<TensorPtr>& parameters, const std::vector<TensorPtr>& gradients, const float* dev_grad_scale) {
    if (parameters.size() != gradients.size()) throw std::runtime_error("Parameters and gradients size mismatch");
    std::vector<float*> ptrs(gradients.size(), nullptr);
    for (size_t i = 0; i < gradients.size(); ++i) {
        if (!gradients[i]) continue;
        if (gradients[i]->numel() != parameters[i]->numel()) throw std::runtime_error("optimizer: gradient size mismatch");
        ptrs[i] = gradients[i]->dev_f32_mut();
    }
    step_device(parameters, ptrs, dev_grad_scale);
}

// optim/adam.cpp:25-90
void Adam::step_device(const std::vector<TensorPtr>& parameters, const std::vector<float*>& gradients, const float* dev_grad_scale) {
    if (parameters.size() != gradients.size()) throw std::runtime_error("Parameters and gradients size mismatch");
    if (adam_config_.amsgrad) throw std::runtime_error("Adam: amsgrad is not implemented on the B200 path");
    if (!loaded_.empty()) bind_state(parameters);
    order_ = parameters;
    std::vector<TensorPtr> m(parameters.size()), v(parameters.size());
    std::vector<float> bc1(parameters.size(), 1.f), bc2(parameters.size(), 1.f);
    bool any = false;
    for (size_t i = 0; i < parameters.size(); ++i) {
        if (!gradients[i]) continue;                  // tensors without a gradient keep their step count
        any = true;
        if (states_.find(parameters[i]) == states_.end()) init_state(parameters[i]);
        AdamState& st = states_[parameters[i]];
        ++st.step;
        m[i] = st.m[0]; v[i] = st.v[0];
        bc1[i] = compute_bias_correction1(st.step);
        bc2[i] = compute_bias_correction2(st.step);
    }
    if (!any) return;
    SlotTable t = optim_detail::upload_slots(parameters, gradients, &m, &v, &bc1, &bc2, &dev_slots_, &dev_slots_bytes_);
    mft_rt::check(mft_adam_step(static_cast<const mft_tensor_slot*>(t.dev), t.count, adam_config_.learning_rate, adam_config_.beta1,
                                adam_config_.beta2, adam_config_.epsilon, adam_config_.weight_decay, dev_grad_scale, mft_rt::stream()),
                  "mft_adam_step");
}

void Adam::update_param(TensorPtr param, const TensorPtr& grad, OptimizerState&) { step({param}, {grad}); }

// Optimizer-state checkpoint: see adam.h. Moments are read back through the tensors' host mirrors (one D2H each).
void Adam::save_state(const std::string& path) const {
    std::map<std::string, TensorPtr> tensors;
    std::vector<std::pair<std::string, std::string>> meta;
    meta.emplace_back("format", "mft_adam_state_v1");
    meta.emplace_back("num_params", std::to_string(order_.size()));
    meta.emplace_back("lr", std::to_string(adam_config_.learning_rate));
    meta.emplace_back("beta1", std::to_string(adam_config_.beta1));
    meta.emplace_back("beta2", std::to_string(adam_config_.beta2));
    meta.emplace_back("eps", std::to_string(adam_config_.epsilon));
    meta.emplace_back("weight_decay", std::to_string(adam_config_.weight_decay));
    for (size_t i = 0; i < order_.size(); ++i) {
        auto it = states_.find(order_[i]);
        if (it == states_.end()) continue;               // never received a gradient: no state, step 0
        const std::string k = "adam." + std::to_string(i);
        tensors[k + ".m"] = it->second.m[0];
        tensors[k + ".v"] = it->second.v[0];
        meta.emplace_back(k + ".step", std::to_string(it->second.step));
    }
    write_safetensors(path, tensors, meta);
}

void Adam::load_state(const std::string& path) {
    SafeTensorsReader rd(path);
    rd.parse_header();
    const auto& md = rd.metadata();
    auto fmt = md.find("format");
    if (fmt == md.end() || fmt->second != "mft_adam_state_v1") throw std::runtime_error("Adam::load_state: " + path + " is not an Adam state file");
    const size_t n = size_t(std::stoul(md.at("num_params")));
    loaded_.assign(n, AdamState{});
    mft_rt::PersistentScope keep;
    for (size_t i = 0; i < n; ++i) {
        const std::string k = "adam." + std::to_string(i);
        auto st = md.find(k + ".step");
        if (st == md.end()) continue;
        loaded_[i].step = size_t(std::stoul(st->second));
        loaded_[i].m.push_back(rd.load_tensor(k + ".m"));
        loaded_[i].v.push_back(rd.load_tensor(k + ".v"));
    }
    states_.clear();
    if (!order_.empty()) bind_state(order_);
}

void Adam::bind_state(const std::vector<TensorPtr>& parameters) {
    if (loaded_.empty()) return;
    if (loaded_.size() != parameters.size())
        throw std::runtime_error("Adam: the loaded state holds " + std::to_string(loaded_.size()) + " entries, the parameter list " + std::to_string(parameters.size()));
    for (size_t i = 0; i < parameters.size(); ++i) {
        if (loaded_[i].m.empty()) continue;
        if (loaded_[i].m[0]->numel() != parameters[i]->numel()) throw std::runtime_error("Adam: loaded state " + std::to_string(i) + " does not match its parameter's size");
        states_[parameters[i]] = std::move(loaded_[i]);
    }
    loaded_.clear();
    order_ = parameters;
}

void AdamState::to_file(const std::string&) const { throw std::runtime_error("AdamState::to_file: not implemented in this build"); }
void AdamState::from_file(const std::string&) { throw std::runtime_error("AdamState::from_file: not implemented in this build"); }

}  // namespace ops
