// opt_ops/energy/power_monitor.cpp:14-122 of the reference, restated.
#include "power_monitor.h"

#include <algorithm>
#include <cmath>
#include <sstream>

namespace ops {
namespace energy {

PowerMonitor::PowerMonitor(const PowerConfig& cfg) : cfg_(cfg) {}

void PowerMonitor::set_manual_readings(float battery_percent, float cpu_temp_c) {
    battery_percent_ = battery_percent;
    cpu_temp_c_ = cpu_temp_c;
}

void PowerMonitor::set_step_schedule(const std::vector<StepSleep>& schedule) { schedule_ = schedule; }

// comma-separated "<first>-<last>:<ms>" items, <last> may be empty; malformed items are skipped; result sorted by first step
std::vector<StepSleep> PowerMonitor::parse_schedule(const std::string& spec) {
    std::vector<StepSleep> out;
    std::stringstream items(spec);
    std::string item;
    while (std::getline(items, item, ',')) {
        const size_t colon = item.find(':');
        if (colon == std::string::npos) continue;
        const std::string range = item.substr(0, colon);
        const size_t dash = range.find('-');
        if (dash == std::string::npos) continue;
        StepSleep s;
        try {
            s.sleep_ms = std::max(0, std::stoi(item.substr(colon + 1)));
            s.start_step = std::stoll(range.substr(0, dash));
        } catch (...) { continue; }
        const std::string last = range.substr(dash + 1);
        try { s.end_step = last.empty() ? -1 : std::stoll(last); } catch (...) { s.end_step = -1; }
        out.push_back(s);
    }
    std::sort(out.begin(), out.end(), [](const StepSleep& a, const StepSleep& b) { return a.start_step < b.start_step; });
    return out;
}

int PowerMonitor::suggest_sleep_ms(int64_t global_step) {
    for (const StepSleep& s : schedule_) {
        if (global_step < s.start_step) break;
        if (s.end_step < 0 || global_step <= s.end_step) return s.sleep_ms;
    }
    if (cfg_.check_interval_steps > 0 && global_step % cfg_.check_interval_steps == 0) {
        const float fb = (cfg_.enable_battery && battery_percent_ < cfg_.battery_threshold) ? cfg_.freq_b_low : cfg_.freq_b_high;
        const float ft = (cfg_.enable_temp && cpu_temp_c_ > cfg_.temp_threshold) ? cfg_.freq_t_low : cfg_.freq_t_high;
        const float f = std::min(fb, ft);                      // steps per second the slower constraint allows
        last_sleep_ms_ = f > 0.0f ? int(std::round(std::min(1000.0f / f, 5000.0f))) : 0;
    }
    return last_sleep_ms_;
}

std::string PowerMonitor::debug_state() const {
    std::ostringstream o;
    o << "PowerMonitor{battery=" << battery_percent_ << "%, temp=" << cpu_temp_c_ << "C, last_sleep_ms=" << last_sleep_ms_ << "}";
    return o.str();
}

}  // namespace energy
}  // namespace ops
