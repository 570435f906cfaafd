#include "memory_manager.h"

#include <cstdio>
#include <cstdlib>
#include <fstream>
#include <iostream>
#include <sstream>

#include "runtime.h"

namespace ops {

MemoryManager& MemoryManager::instance() {
    static MemoryManager m;
    return m;
}
void* MemoryManager::allocate(size_t size) { return std::malloc(size ? size : 1); }
void MemoryManager::deallocate(void* ptr, size_t) { std::free(ptr); }
void MemoryManager::clear_unused_memory() { mft_rt::release_cached(); }
void MemoryManager::force_cleanup() {
    mft_rt::sync();
    mft_rt::release_cached();
}
size_t MemoryManager::get_memory_usage() const { return mft_rt::bytes_in_use(); }
size_t MemoryManager::get_peak_memory() const { return mft_rt::arena_high_water(); }
void MemoryManager::print_memory_stats() const {
    auto mb = [](size_t b) { return double(b) / (1024.0 * 1024.0); };
    std::printf("[MemoryManager] HBM in use %.1f MB, cached %.1f MB, step-arena high water %.1f MB\n", mb(mft_rt::bytes_in_use()),
                mb(mft_rt::bytes_cached()), mb(mft_rt::arena_high_water()));
}

double MemoryMonitor::memory_threshold_ = 0.9;

static bool meminfo(size_t& total_kb, size_t& avail_kb) {
    std::ifstream f("/proc/meminfo");
    if (!f) return false;
    std::string line;
    total_kb = avail_kb = 0;
    while (std::getline(f, line)) {
        std::istringstream ss(line);
        std::string key;
        size_t v = 0;
        ss >> key >> v;
        if (key == "MemTotal:") total_kb = v;
        else if (key == "MemAvailable:") avail_kb = v;
    }
    return total_kb > 0;
}
size_t MemoryMonitor::get_system_memory_usage() {
    size_t t, a;
    return meminfo(t, a) ? (t - a) * 1024 : 0;
}
size_t MemoryMonitor::get_system_available_memory() {
    size_t t, a;
    return meminfo(t, a) ? a * 1024 : 0;
}
bool MemoryMonitor::check_memory_threshold() {
    size_t t, a;
    return meminfo(t, a) && double(t - a) / double(t) > memory_threshold_;
}
bool MemoryMonitor::is_memory_pressure() { return check_memory_threshold(); }
void MemoryMonitor::set_memory_threshold(double fraction) { memory_threshold_ = fraction; }
void MemoryMonitor::log_memory_status(const std::string& context) {
    std::cout << "[MemoryMonitor] " << context << (context.empty() ? "" : ": ") << "host used " << get_system_memory_usage() / (1024 * 1024)
              << " MB, available " << get_system_available_memory() / (1024 * 1024) << " MB" << std::endl;
}

}  // namespace ops
