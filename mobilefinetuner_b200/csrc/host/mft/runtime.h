// Device runtime under the host API: the compute stream, a caching HBM allocator with an optional
// per-step bump arena (deterministic addresses => the whole step is CUDA-graph capturable), and
// status -> exception mapping for the C ABI. Nothing here computes on the CPU.
#pragma once
#include <cstddef>
#include <cstdint>
#include <string>

namespace mft_rt {

void* stream();                                  // the cudaStream_t every host-API op enqueues on
void set_stream(void* s);
void sync();                                     // block until the stream is idle

void* dev_alloc(size_t bytes);                   // HBM; from the step arena when one is active
void dev_free(void* p, size_t bytes);
void h2d(void* dst, const void* src, size_t bytes);                      // async on stream()
void d2h(void* dst, const void* src, size_t bytes);                      // copies, then syncs
void h2d_2d(void* dst, size_t dpitch, const void* src, size_t spitch, size_t width, size_t rows);
void d2h_2d(void* dst, size_t dpitch, const void* src, size_t spitch, size_t width, size_t rows);
void d2d(void* dst, const void* src, size_t bytes);
void d2d_2d(void* dst, size_t dpitch, const void* src, size_t spitch, size_t width, size_t rows);
void memset0(void* dst, size_t bytes);
void* pinned_alloc(size_t bytes);               // page-locked host memory (persistent staging buffers)
void pinned_free(void* p);

// Step arena: between arena_begin() and arena_end() every dev_alloc is a bump allocation from one slab that
// arena_reset() rewinds; persistent allocations (parameters, optimizer state, cached weight shadows) are made
// under a PersistentScope and bypass it.
void arena_enable(size_t bytes);
void arena_acquire(size_t bytes);                // ref-counted enable (one per trainer); refuses to move a slab that is in use
void arena_release();                            // last release switches the arena off
void arena_reset();
bool arena_active();
size_t arena_high_water();
struct PersistentScope {
    PersistentScope();
    ~PersistentScope();
private:
    int prev_;
};
void release_cached();                           // return cached blocks to the driver
size_t bytes_in_use();
size_t bytes_cached();

void check(int status, const char* what);        // non-zero C-ABI status -> throws ops::TensorError
[[noreturn]] void fail(const std::string& msg);

// Scratch for kernels that need workspace (norm/colsum/CE/sumsq partials). Grows on demand, persistent.
void* workspace(size_t bytes);

}  // namespace mft_rt
