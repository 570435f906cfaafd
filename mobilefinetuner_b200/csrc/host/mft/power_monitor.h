// Step-pacing policy — API mirror of the reference's opt_ops/energy/power_monitor.h:17-75. On a phone it throttles the
// training loop from battery level and CPU temperature; the B200 build keeps the class so the unchanged trainers link and
// their --pm_* flags keep their meaning (a schedule of sleeps per step range, else a rate derived from two thresholds).
// It is host-only bookkeeping and never touches the GPU path.
#pragma once
#include <cstdint>
#include <string>
#include <vector>

namespace ops {
namespace energy {

struct PowerConfig {
    int check_interval_steps = 0;      // 0: the policy is never re-evaluated
    float battery_threshold = 20.0f;   // percent
    float freq_b_high = 2.0f;          // steps per second above / below the battery threshold
    float freq_b_low = 0.5f;
    bool enable_battery = true;
    float temp_threshold = 42.0f;      // Celsius
    float freq_t_high = 2.0f;          // steps per second below / above the temperature threshold
    float freq_t_low = 0.5f;
    bool enable_temp = true;
};

struct StepSleep {
    int64_t start_step = 0;   // inclusive
    int64_t end_step = -1;    // inclusive; -1 = open ended
    int sleep_ms = 0;
};

class PowerMonitor {
public:
    explicit PowerMonitor(const PowerConfig& cfg = PowerConfig());
    void set_manual_readings(float battery_percent, float cpu_temp_c);
    void set_step_schedule(const std::vector<StepSleep>& schedule);
    static std::vector<StepSleep> parse_schedule(const std::string& spec);   // "0-99:0,100-:250"
    int suggest_sleep_ms(int64_t global_step);
    std::string debug_state() const;

private:
    PowerConfig cfg_;
    float battery_percent_ = 100.0f;
    float cpu_temp_c_ = 30.0f;
    std::vector<StepSleep> schedule_;
    int last_sleep_ms_ = 0;
};

}  // namespace energy
}  // namespace ops
