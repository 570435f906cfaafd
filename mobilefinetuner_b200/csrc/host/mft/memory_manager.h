// MemoryManager facade — API mirror of the reference's core/memory_manager.h:112-196 for the calls the trainers make
// (force_cleanup / clear_unused_memory / cleanup_dead_references / print_memory_stats, MemoryGuard, MemoryMonitor).
// In the reference this class owns a host malloc pool behind every Tensor; here tensors live in HBM and are served by the
// caching allocator / step arena of mft_rt (runtime.h), so the "pool" calls map onto that allocator: cleanup returns
// cached device blocks to the driver, the statistics are HBM bytes. allocate()/deallocate() remain plain host memory.
#pragma once
#include <cstddef>
#include <memory>
#include <string>

#include "tensor.h"

namespace ops {

class MemoryManager {
public:
    static MemoryManager& instance();
    void* allocate(size_t size);
    void deallocate(void* ptr, size_t size);
    std::shared_ptr<Tensor> get_cached_tensor(const std::string&) { return nullptr; }   // no tensor cache: nothing is ever found
    void cache_tensor(const std::string&, std::shared_ptr<Tensor>) {}
    void register_tensor(std::shared_ptr<Tensor>) {}          // the autograd engine owns graph lifetimes here
    void clear_computation_graph() {}
    void cleanup_dead_references() {}
    void clear_cache() {}
    void clear_unused_memory();                               // cached (free) HBM blocks -> driver
    void force_cleanup();                                     // sync the stream, then the same
    void print_memory_stats() const;
    size_t get_memory_usage() const;                          // HBM bytes held by live tensors
    size_t get_peak_memory() const;                           // high-water mark of the step arena (0 without one)
    void set_cache_size(size_t) {}
    MemoryManager(const MemoryManager&) = delete;
    MemoryManager& operator=(const MemoryManager&) = delete;

private:
    MemoryManager() = default;
};

class MemoryGuard {
public:
    MemoryGuard() = default;
    ~MemoryGuard() { if (active_) MemoryManager::instance().clear_unused_memory(); }
    void release() { active_ = false; }
    MemoryGuard(const MemoryGuard&) = delete;
    MemoryGuard& operator=(const MemoryGuard&) = delete;

private:
    bool active_ = true;
};

// host RAM figures from /proc/meminfo (the trainers only log them)
class MemoryMonitor {
public:
    static size_t get_system_memory_usage();
    static size_t get_system_available_memory();
    static bool is_memory_pressure();
    static void log_memory_status(const std::string& context = "");
    static void set_memory_threshold(double fraction);
    static bool check_memory_threshold();

private:
    static double memory_threshold_;
};

}  // namespace ops
