// optim/gemma_trainer.cpp:17-233 of the reference, restated over the device path.
#include "gemma_trainer.h"

#include <algorithm>
#include <chrono>
#include <cmath>
#include <iostream>
#include <thread>

#include "lm_loss.h"
#include "memory_manager.h"
#include "ops.h"
#include "wikitext2_dataset.h"

namespace ops {

GemmaLoRATrainer::GemmaLoRATrainer(GemmaModel& model, GemmaLoraInjector& injector, WikiText2Dataset& train_data, WikiText2Dataset& eval_data,
                                   const GemmaTrainerConfig& config)
    : model_(model), injector_(injector), train_data_(train_data), eval_data_(eval_data), config_(config) {
    AdamConfig ac;
    ac.learning_rate = config_.learning_rate;
    ac.beta1 = config_.adam_beta1;
    ac.beta2 = config_.adam_beta2;
    ac.epsilon = config_.adam_eps;
    ac.weight_decay = config_.weight_decay;
    ac.clip_grad_norm = config_.max_grad_norm;
    optimizer_ = std::make_unique<Adam>(ac);
    energy::PowerConfig pc;
    pc.check_interval_steps = config_.pm_interval;
    pc.battery_threshold = config_.pm_batt_thresh;
    pc.temp_threshold = config_.pm_temp_thresh;
    pc.freq_b_high = config_.pm_fb_high; pc.freq_b_low = config_.pm_fb_low;
    pc.freq_t_high = config_.pm_ft_high; pc.freq_t_low = config_.pm_ft_low;
    pc.enable_battery = config_.pm_enable_batt;
    pc.enable_temp = config_.pm_enable_temp;
    power_monitor_ = energy::PowerMonitor(pc);
    power_monitor_.set_manual_readings(config_.pm_manual_batt, config_.pm_manual_temp);
    if (!config_.pm_schedule.empty()) power_monitor_.set_step_schedule(energy::PowerMonitor::parse_schedule(config_.pm_schedule));
    std::cout << "[GemmaTrainer] LR=" << config_.learning_rate << ", epochs=" << config_.num_epochs << ", grad_accum=" << config_.grad_accum_steps << std::endl;
}

// warm-up over warmup_ratio of all optimiser updates, then linear (or cosine) decay to zero
float GemmaLoRATrainer::get_lr(int step) {
    const int64_t micro_per_epoch = (int64_t(train_data_.num_sequences()) + config_.micro_batch_size - 1) / config_.micro_batch_size;
    const int64_t updates = std::max<int64_t>(1, (micro_per_epoch * config_.num_epochs + config_.grad_accum_steps - 1) / config_.grad_accum_steps);
    const int warmup = int(float(updates) * config_.warmup_ratio);
    if (warmup > 0 && step <= warmup) return config_.learning_rate * (float(step) / float(warmup));
    float progress = float(step - warmup) / float(std::max<int64_t>(1, updates - warmup));
    progress = std::min(1.0f, std::max(0.0f, progress));
    if (config_.lr_scheduler == "cosine") return config_.learning_rate * 0.5f * (1.0f + std::cos(3.14159265f * progress));
    return config_.learning_rate * (1.0f - progress);
}

// global L2 norm over the adapters' gradients; scaled by max/(norm + 1e-6) when above max (reference: float accumulation on
// the host, optim/gemma_trainer.cpp:84-102; here a double reduction on the device — the factor agrees to ~1e-7 relative)
void GemmaLoRATrainer::clip_gradients() {
    optimizer_->clip_grad_norm_eps(injector_.get_trainable_params(), config_.max_grad_norm, 1e-6f);
}

float GemmaLoRATrainer::train_step(const Batch& batch) {
    ++micro_step_counter_;
    if (config_.dump_embedding && !embedding_dump_scheduled_ && micro_step_counter_ == config_.dump_embedding_step) {
        model_.request_embedding_dump(config_.dump_embedding_step, config_.dump_embedding_dir);
        embedding_dump_scheduled_ = true;
    }
    auto logits = model_.forward(batch.input_ids, batch.attention_mask);
    auto loss = lm_cross_entropy(logits, batch.labels, -100, "mean");
    const float loss_val = static_cast<const Tensor&>(*loss).data<float>()[0];
    auto scaled = mul(loss, 1.0f / float(config_.grad_accum_steps));
    scaled->backward();
    ++accum_counter_;
    accum_loss_ += loss_val;
    if (accum_counter_ < config_.grad_accum_steps) return -1.0f;
    clip_gradients();
    const std::vector<TensorPtr> params = injector_.get_trainable_params();
    std::vector<TensorPtr> grads;
    grads.reserve(params.size());
    for (const auto& p : params) grads.push_back(p->grad());
    ++global_step_;
    optimizer_->set_learning_rate(get_lr(global_step_));
    optimizer_->step(params, grads);
    for (const auto& p : params) if (p->grad()) p->zero_grad();
    accum_counter_ = 0;
    const float avg = accum_loss_ / float(config_.grad_accum_steps);
    accum_loss_ = 0.0f;
    return avg;
}

float GemmaLoRATrainer::evaluate() {
    std::cout << "[GemmaTrainer] Eval started..." << std::endl;
    eval_data_.reset_cursor();
    float total = 0.0f;
    int batches = 0;
    for (;;) {
        Batch b = eval_data_.next_batch(size_t(config_.micro_batch_size), false);
        if (!b.input_ids) break;
        auto logits = model_.forward(b.input_ids, b.attention_mask);
        auto loss = lm_cross_entropy(logits, b.labels, -100, "mean");
        total += static_cast<const Tensor&>(*loss).data<float>()[0];
        ++batches;
    }
    const float mean = batches > 0 ? total / float(batches) : 0.0f;
    std::cout << "  Eval Loss: " << mean << " PPL: " << perplexity_from_loss(mean) << std::endl;
    return mean;
}

void GemmaLoRATrainer::train() {
    for (int epoch = 0; epoch < config_.num_epochs; ++epoch) {
        std::cout << "\n=== Gemma Epoch " << (epoch + 1) << "/" << config_.num_epochs << " ===" << std::endl;
        train_data_.reset_cursor();
        float epoch_loss = 0.0f;
        int steps = 0;
        bool stop = false;
        for (;;) {
            Batch b = train_data_.next_batch(size_t(config_.micro_batch_size), false);
            if (!b.input_ids) break;
            const float loss = train_step(b);
            if (loss < 0.0f) continue;
            epoch_loss += loss;
            ++steps;
            if (global_step_ % config_.logging_steps == 0)
                std::cout << "[Step " << global_step_ << "] Loss=" << loss << " PPL=" << perplexity_from_loss(loss) << " LR=" << get_lr(global_step_) << std::endl;
            if (config_.eval_steps > 0 && global_step_ % config_.eval_steps == 0) evaluate();
            if (config_.max_steps > 0 && global_step_ >= config_.max_steps) { stop = true; break; }
            const int sleep_ms = power_monitor_.suggest_sleep_ms(global_step_);
            if (sleep_ms > 0) std::this_thread::sleep_for(std::chrono::milliseconds(sleep_ms));
        }
        std::cout << "Epoch " << (epoch + 1) << " avg loss: " << (steps > 0 ? epoch_loss / float(steps) : 0.0f) << std::endl;
        if (stop) {
            std::cout << "[GemmaTrainer] Reached max_steps=" << config_.max_steps << ", stopping early." << std::endl;
            break;
        }
    }
    MemoryManager::instance().force_cleanup();
}

void GemmaLoRATrainer::save_lora(const std::string& path) { injector_.save_lora_safetensors(path); }

}  // namespace ops
