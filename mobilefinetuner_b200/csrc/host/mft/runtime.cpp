#include "runtime.h"

#include <cuda_runtime.h>

#include <cstdio>
#include <cstdlib>
#include <unordered_map>
#include <vector>

#include "mft_b200.h"
#include "tensor.h"

namespace mft_rt {

namespace {
cudaStream_t g_stream = nullptr;
bool g_stream_made = false;

struct Pool {
    std::unordered_map<size_t, std::vector<void*>> free_lists;
    size_t in_use = 0, cached = 0;
    ~Pool() {}  // blocks are released with the context at process exit
} g_pool;

struct Arena {
    char* base = nullptr;
    size_t size = 0, off = 0, high = 0;
    bool enabled = false;
    int users = 0;             // live trainers sharing the slab
} g_arena;
int g_persistent_depth = 0;

void* g_ws = nullptr;
size_t g_ws_bytes = 0;

size_t round_up(size_t b) {
    if (b < 512) return 512;
    if (b < (1u << 20)) return (b + 511) & ~size_t(511);
    return (b + (size_t(2) << 20) - 1) & ~((size_t(2) << 20) - 1);   // large blocks: 2 MiB granularity
}

void cuda_ok(cudaError_t e, const char* what) {
    if (e != cudaSuccess) fail(std::string(what) + ": " + cudaGetErrorString(e));
}
}  // namespace

[[noreturn]] void fail(const std::string& msg) { throw ops::TensorError(msg); }

void check(int status, const char* what) {
    if (status == 0) return;
    const char* m = mft_last_error();
    fail(std::string(what) + " failed (status " + std::to_string(status) + "): " + (m ? m : ""));
}

void* stream() {
    if (!g_stream_made) {
        // Fail loudly when there is no B200: there is no CPU fallback anywhere in this library.
        if (mft_device_check() != 0) fail(std::string("mft_b200 requires an sm_100 GPU: ") + mft_last_error());
        cuda_ok(cudaStreamCreateWithFlags(&g_stream, cudaStreamNonBlocking), "cudaStreamCreate");
        g_stream_made = true;
    }
    return g_stream;
}
void set_stream(void* s) { g_stream = static_cast<cudaStream_t>(s); g_stream_made = true; }
void sync() { cuda_ok(cudaStreamSynchronize(static_cast<cudaStream_t>(stream())), "cudaStreamSynchronize"); }

void* dev_alloc(size_t bytes) {
    if (bytes == 0) bytes = 4;
    if (g_arena.enabled && g_persistent_depth == 0) {
        size_t need = (bytes + 1023) & ~size_t(1023);
        if (g_arena.off + need > g_arena.size)
            fail("mft step arena exhausted (" + std::to_string(g_arena.size >> 20) + " MiB); raise MFT_ARENA_MB");
        void* p = g_arena.base + g_arena.off;
        g_arena.off += need;
        if (g_arena.off > g_arena.high) g_arena.high = g_arena.off;
        return p;
    }
    size_t rb = round_up(bytes);
    auto it = g_pool.free_lists.find(rb);
    if (it != g_pool.free_lists.end() && !it->second.empty()) {
        void* p = it->second.back();
        it->second.pop_back();
        g_pool.cached -= rb;
        g_pool.in_use += rb;
        return p;
    }
    (void)stream();
    void* p = nullptr;
    cudaError_t e = cudaMalloc(&p, rb);
    if (e != cudaSuccess) {
        release_cached();
        e = cudaMalloc(&p, rb);
    }
    cuda_ok(e, "cudaMalloc");
    g_pool.in_use += rb;
    return p;
}

void dev_free(void* p, size_t bytes) {
    if (!p) return;
    if (g_arena.base && p >= g_arena.base && p < g_arena.base + g_arena.size) return;  // arena memory: rewound wholesale
    if (bytes == 0) bytes = 4;
    size_t rb = round_up(bytes);
    // Stream-ordered reuse: every op runs on the one compute stream, so a cached block can be handed out again
    // immediately (later kernels are ordered after the earlier readers).
    g_pool.free_lists[rb].push_back(p);
    g_pool.in_use -= rb;
    g_pool.cached += rb;
}

void release_cached() {
    cudaDeviceSynchronize();
    for (auto& kv : g_pool.free_lists) {
        for (void* p : kv.second) cudaFree(p);
        kv.second.clear();
    }
    g_pool.cached = 0;
}
size_t bytes_in_use() { return g_pool.in_use; }
size_t bytes_cached() { return g_pool.cached; }

void arena_enable(size_t bytes) {
    if (g_arena.base && g_arena.size >= bytes) { g_arena.enabled = true; return; }
    (void)stream();
    if (g_arena.base) { cudaDeviceSynchronize(); cudaFree(g_arena.base); g_arena.base = nullptr; }
    void* p = nullptr;
    cuda_ok(cudaMalloc(&p, bytes), "cudaMalloc(step arena)");
    g_arena.base = static_cast<char*>(p);
    g_arena.size = bytes;
    g_arena.off = 0;
    g_arena.enabled = true;
}
// Trainers share one slab: they all enqueue on the one compute stream and nothing in the arena outlives a step, so a
// reset by one trainer only recycles memory whose readers are already ordered before the next writer. The slab cannot move
// while a trainer is alive (its captured graphs and gradients point into it).
void arena_acquire(size_t bytes) {
    if (g_arena.users > 0 && g_arena.base && g_arena.size < bytes)
        fail("mft step arena: a live trainer uses a " + std::to_string(g_arena.size >> 20) + " MiB arena; a second trainer asked for " +
             std::to_string(bytes >> 20) + " MiB. Destroy the first trainer or create the larger one first.");
    arena_enable(bytes);
    ++g_arena.users;
}
void arena_release() {
    if (g_arena.users > 0) --g_arena.users;
    if (g_arena.users == 0) {              // ops outside a trainer go back to the caching allocator
        g_arena.enabled = false;
        g_arena.off = 0;
    }
}
void arena_reset() { g_arena.off = 0; }
bool arena_active() { return g_arena.enabled && g_persistent_depth == 0; }
size_t arena_high_water() { return g_arena.high; }
PersistentScope::PersistentScope() : prev_(g_persistent_depth) { ++g_persistent_depth; }
PersistentScope::~PersistentScope() { g_persistent_depth = prev_; }

void h2d(void* dst, const void* src, size_t bytes) {
    if (!bytes) return;
    cuda_ok(cudaMemcpyAsync(dst, src, bytes, cudaMemcpyHostToDevice, static_cast<cudaStream_t>(stream())), "H2D copy");
    // pageable sources are staged by the driver before the call returns; pinned ones must stay alive until the
    // stream reaches the copy — callers that reuse the buffer synchronise.
}
void d2h(void* dst, const void* src, size_t bytes) {
    if (!bytes) return;
    cuda_ok(cudaMemcpyAsync(dst, src, bytes, cudaMemcpyDeviceToHost, static_cast<cudaStream_t>(stream())), "D2H copy");
    sync();
}
void h2d_2d(void* dst, size_t dpitch, const void* src, size_t spitch, size_t width, size_t rows) {
    if (!width || !rows) return;
    cuda_ok(cudaMemcpy2DAsync(dst, dpitch, src, spitch, width, rows, cudaMemcpyHostToDevice,
                              static_cast<cudaStream_t>(stream())), "H2D 2-D copy");
}
void d2h_2d(void* dst, size_t dpitch, const void* src, size_t spitch, size_t width, size_t rows) {
    if (!width || !rows) return;
    cuda_ok(cudaMemcpy2DAsync(dst, dpitch, src, spitch, width, rows, cudaMemcpyDeviceToHost,
                              static_cast<cudaStream_t>(stream())), "D2H 2-D copy");
    sync();
}
void d2d(void* dst, const void* src, size_t bytes) {
    if (!bytes) return;
    cuda_ok(cudaMemcpyAsync(dst, src, bytes, cudaMemcpyDeviceToDevice, static_cast<cudaStream_t>(stream())), "D2D copy");
}
void d2d_2d(void* dst, size_t dpitch, const void* src, size_t spitch, size_t width, size_t rows) {
    if (!width || !rows) return;
    cuda_ok(cudaMemcpy2DAsync(dst, dpitch, src, spitch, width, rows, cudaMemcpyDeviceToDevice,
                              static_cast<cudaStream_t>(stream())), "D2D 2-D copy");
}
void memset0(void* dst, size_t bytes) {
    if (!bytes) return;
    cuda_ok(cudaMemsetAsync(dst, 0, bytes, static_cast<cudaStream_t>(stream())), "memset");
}

void* pinned_alloc(size_t bytes) {
    void* p = nullptr;
    cuda_ok(cudaMallocHost(&p, bytes ? bytes : 16), "cudaMallocHost");
    return p;
}
void pinned_free(void* p) { if (p) cudaFreeHost(p); }

void* workspace(size_t bytes) {
    if (bytes <= g_ws_bytes) return g_ws;
    // Grown rarely (first steps only). Older scratch may still be in flight: retire it to the pool, which is
    // stream-ordered like every other block.
    PersistentScope keep;
    if (g_ws) dev_free(g_ws, g_ws_bytes);
    g_ws_bytes = round_up(bytes * 2);
    g_ws = dev_alloc(g_ws_bytes);
    return g_ws;
}

}  // namespace mft_rt
