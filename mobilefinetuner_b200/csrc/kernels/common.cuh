// Shared device/host helpers for the mft_b200 kernels (sm_100a only).
#pragma once
#include <cuda_bf16.h>
#include <cuda_runtime.h>
#include <stdint.h>
#include <stdio.h>
#include <string.h>
#include "mft_b200.h"

namespace mft {

// Records a message for mft_last_error() and returns `code`.
int set_error(int code, const char* fmt, ...);

inline cudaStream_t as_stream(mft_stream_t s) { return reinterpret_cast<cudaStream_t>(s); }

// Kernel-launch accounting (mft_launch_count): every entry point reports how many kernels it enqueued.
void add_launches(int n);

// Post-launch check used by every entry point (counts one kernel; multi-kernel entry points add the rest).
#define MFT_CHECK_LAUNCH(name)                                                             \
    do {                                                                                   \
        ::mft::add_launches(1);                                                            \
        cudaError_t _e = cudaGetLastError();                                               \
        if (_e != cudaSuccess) return ::mft::set_error(2, "%s: %s", name, cudaGetErrorString(_e)); \
    } while (0)

#define MFT_CUDA_OK(expr, name)                                                            \
    do {                                                                                   \
        cudaError_t _e = (expr);                                                           \
        if (_e != cudaSuccess) return ::mft::set_error(2, "%s: %s", name, cudaGetErrorString(_e)); \
    } while (0)

constexpr int kNumSMs = 148;  // B200

// Function attributes (dynamic shared-memory limit) are per device: `flags` is a per-call-site table indexed by the current
// device, so a process that drives several GPUs sets the attribute on each of them once.
constexpr int kMaxDevices = 64;
inline bool first_use_on_device(bool (&flags)[kMaxDevices]) {
    int dev = 0;
    if (cudaGetDevice(&dev) != cudaSuccess || dev < 0 || dev >= kMaxDevices) return true;   // unknown: set it again, it is cheap
    if (flags[dev]) return false;
    flags[dev] = true;
    return true;
}

__device__ __forceinline__ float warp_sum(float v) {
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
    return v;
}
__device__ __forceinline__ double warp_sum(double v) {
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
    return v;
}
__device__ __forceinline__ float warp_max(float v) {
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) v = fmaxf(v, __shfl_xor_sync(0xffffffffu, v, o));
    return v;
}

// Block-wide reductions (blockDim.x multiple of 32, <= 1024). `smem` must hold 32 elements.
template <typename T>
__device__ __forceinline__ T block_sum(T v, T* smem) {
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5, nw = (blockDim.x + 31) >> 5;
    v = warp_sum(v);
    __syncthreads();
    if (lane == 0) smem[warp] = v;
    __syncthreads();
    T r = (lane < nw) ? smem[lane] : T(0);
    r = warp_sum(r);
    return r;  // valid in every thread
}
__device__ __forceinline__ float block_max(float v, float* smem) {
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5, nw = (blockDim.x + 31) >> 5;
    v = warp_max(v);
    __syncthreads();
    if (lane == 0) smem[warp] = v;
    __syncthreads();
    float r = (lane < nw) ? smem[lane] : -INFINITY;
    r = warp_max(r);
    return r;
}

// Streaming 128-bit accesses (data touched once: keep it out of L1).
__device__ __forceinline__ float4 ld_stream(const float4* p) {
    float4 r;
    asm volatile("ld.global.nc.L1::no_allocate.v4.f32 {%0,%1,%2,%3}, [%4];"
                 : "=f"(r.x), "=f"(r.y), "=f"(r.z), "=f"(r.w) : "l"(p));
    return r;
}
__device__ __forceinline__ float2 ld_stream(const float2* p) {
    float2 r;
    asm volatile("ld.global.nc.L1::no_allocate.v2.f32 {%0,%1}, [%2];" : "=f"(r.x), "=f"(r.y) : "l"(p));
    return r;
}
__device__ __forceinline__ float ld_stream(const float* p) {
    float r;
    asm volatile("ld.global.nc.L1::no_allocate.f32 %0, [%1];" : "=f"(r) : "l"(p));
    return r;
}
__device__ __forceinline__ void st_stream(float4* p, const float4& v) {
    asm volatile("st.global.L1::no_allocate.v4.f32 [%0], {%1,%2,%3,%4};"
                 :: "l"(p), "f"(v.x), "f"(v.y), "f"(v.z), "f"(v.w) : "memory");
}

// Four fp32 values -> four bf16 (round to nearest even), one 8-byte store. Producers of GEMM operands use it to leave a
// bf16 copy next to their fp32 result, so the bf16 tensor-core GEMM that consumes it needs no separate cast pass.
__device__ __forceinline__ void st_bf16x4(void* dst, const float4& v) {
    const __nv_bfloat162 lo = __floats2bfloat162_rn(v.x, v.y), hi = __floats2bfloat162_rn(v.z, v.w);
    uint2 p;
    p.x = *reinterpret_cast<const uint32_t*>(&lo);
    p.y = *reinterpret_cast<const uint32_t*>(&hi);
    *reinterpret_cast<uint2*>(dst) = p;
}

// ---- legacy-pipe tensor-core helpers (mma.sync m16n8k8 TF32), shared by attention and the LoRA factor kernels
__device__ __forceinline__ uint32_t f2tf32(float x) {
    uint32_t r;
    asm("cvt.rna.tf32.f32 %0, %1;" : "=r"(r) : "f"(x));
    return r;
}
__device__ __forceinline__ void split_tf32(float x, uint32_t& hi, uint32_t& lo) {
    hi = f2tf32(x);
    lo = f2tf32(x - __uint_as_float(hi));
}
__device__ __forceinline__ void mma_tf32(float (&c)[4], const uint32_t (&a)[4], uint32_t b0, uint32_t b1) {
    asm volatile(
        "mma.sync.aligned.m16n8k8.row.col.f32.tf32.tf32.f32 {%0,%1,%2,%3}, {%4,%5,%6,%7}, {%8,%9}, {%0,%1,%2,%3};"
        : "+f"(c[0]), "+f"(c[1]), "+f"(c[2]), "+f"(c[3])
        : "r"(a[0]), "r"(a[1]), "r"(a[2]), "r"(a[3]), "r"(b0), "r"(b1));
}
__device__ __forceinline__ float gelu_tanh_f(float x) {
    // reference constants: core/ops.cpp:1055-1059
    const float k0 = 0.7978845608f, k1 = 0.044715f;
    float inner = k0 * (x + k1 * x * x * x);
    return 0.5f * x * (1.0f + tanhf(inner));
}
__device__ __forceinline__ float gelu_tanh_grad_f(float x) {
    // core/backward_functions.cpp:155-177
    const float k0 = 0.7978845608f, k1 = 0.044715f;
    float x2 = x * x;
    float inner = k0 * (x + k1 * x * x2);
    float t = tanhf(inner);
    float sech2 = 1.0f - t * t;
    float dinner = k0 * (1.0f + 3.0f * k1 * x2);
    return 0.5f * (1.0f + t) + 0.5f * x * sech2 * dinner;
}


// ---- programmatic dependent launch (opt-in: MFT_PDL=1). Every kernel of the library starts with pdl_sync(): it lets the NEXT
// kernel of the stream begin launching (its CTAs become resident as this kernel's leave), then waits until the PREVIOUS kernel
// has completed and its writes are visible. Launched through pdl_launch with the attribute, a kernel overlaps only its launch
// latency with the tail of its predecessor, never its memory accesses. Without the attribute both instructions are no-ops.
// Measured on B200 with the captured steps (profiles/r2_pdl_on_off.txt): c1 285 k vs 304 k tokens/s, c3 80.5 k vs 83.4 k with
// the attribute on vs off — early-resident CTAs spinning in griddepcontrol.wait cost more than the launch gap they hide, so
// the attribute stays off; all 222 GPU tests pass either way.
__device__ __forceinline__ void pdl_sync() {
    asm volatile("griddepcontrol.launch_dependents;" ::: "memory");
    asm volatile("griddepcontrol.wait;" ::: "memory");
}
bool pdl_enabled();      // elementwise.cu (reads MFT_PDL once)

template <typename... KArgs>
struct PdlLaunch {
    void (*kernel)(KArgs...);
    dim3 grid, block;
    size_t smem;
    cudaStream_t stream;
    template <typename... Args>
    void operator()(Args&&... args) const {
        cudaLaunchConfig_t cfg;
        memset(&cfg, 0, sizeof(cfg));
        cfg.gridDim = grid; cfg.blockDim = block; cfg.dynamicSmemBytes = smem; cfg.stream = stream;
        cudaLaunchAttribute attr[1];
        attr[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
        attr[0].val.programmaticStreamSerializationAllowed = 1;
        cfg.attrs = attr;
        cfg.numAttrs = pdl_enabled() ? 1 : 0;
        cudaLaunchKernelEx(&cfg, kernel, static_cast<KArgs>(args)...);     // errors surface through MFT_CHECK_LAUNCH (cudaGetLastError)
    }
};
template <typename... KArgs>
inline PdlLaunch<KArgs...> pdl_launch(void (*kernel)(KArgs...), dim3 grid, dim3 block, size_t smem, cudaStream_t stream) {
    return PdlLaunch<KArgs...>{kernel, grid, block, smem, stream};
}

inline bool aligned16(const void* p) { return (reinterpret_cast<uintptr_t>(p) & 15u) == 0; }
inline int64_t ceil_div(int64_t a, int64_t b) { return (a + b - 1) / b; }

// Grid size for grid-stride streaming kernels: enough CTAs to fill the chip a few times over,
// in multiples of the SM count.
inline int stream_grid(int64_t work_items, int threads, int per_thread = 1) {
    int64_t blocks = ceil_div(work_items, int64_t(threads) * per_thread);
    int64_t cap = int64_t(kNumSMs) * 8;
    if (blocks > cap) blocks = cap;
    if (blocks < 1) blocks = 1;
    return static_cast<int>(blocks);
}

}  // namespace mft
