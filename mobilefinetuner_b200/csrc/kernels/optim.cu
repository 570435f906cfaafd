// Multi-tensor gradient-norm / clip / Adam. One launch covers every trainable tensor through a slot
// table in device memory. Arithmetic follows optim/adam.cpp:25-90 operation by operation (explicit
// round-to-nearest mul/add so nvcc cannot contract into FMAs the CPU reference does not use).
// HBM-bound: sumsq reads 4 B/param, Adam moves 28 B/param (read p,g,m,v; write p,m,v).
#include "common.cuh"

namespace mft {

constexpr int kOptCtas = kNumSMs * 4;
constexpr int kOptThreads = 256;

// Work decomposition shared by the multi-tensor kernels: every tensor is cut into chunks of kOptChunk elements and the CTAs
// grid-stride over the concatenated chunk list, so 96 small LoRA factors fill the chip as well as one large matrix does
// (a loop "for each tensor: grid-stride over its elements" leaves most CTAs idle on tensors smaller than the grid).
constexpr int kOptChunk = kOptThreads * 16;
constexpr int kOptMaxSlots = 12000;       // (slots + 1) ints of shared memory

// pre[s] = number of chunks before tensor s (tensors without a gradient contribute none when grads_only). Returns the total.
__device__ int chunk_prefix(const mft_tensor_slot* __restrict__ slots, int num_slots, int* pre, bool grads_only) {
    for (int s = threadIdx.x; s < num_slots; s += blockDim.x)
        pre[s + 1] = (grads_only && !slots[s].grad) ? 0 : int((slots[s].n + kOptChunk - 1) / kOptChunk);
    if (threadIdx.x == 0) pre[0] = 0;
    __syncthreads();
    if (threadIdx.x < 32) {
        int carry = 0;
        for (int base = 0; base < num_slots; base += 32) {
            const int i = base + int(threadIdx.x);
            int v = i < num_slots ? pre[i + 1] : 0;
#pragma unroll
            for (int o = 1; o < 32; o <<= 1) {
                const int u = __shfl_up_sync(0xffffffffu, v, o);
                if (int(threadIdx.x) >= o) v += u;
            }
            if (i < num_slots) pre[i + 1] = v + carry;
            carry += __shfl_sync(0xffffffffu, v, 31);
        }
    }
    __syncthreads();
    return pre[num_slots];
}
__device__ __forceinline__ int chunk_slot(const int* pre, int num_slots, int c) {   // s with pre[s] <= c < pre[s+1]
    int lo = 0, hi = num_slots - 1;
    while (lo < hi) {
        const int mid = (lo + hi + 1) >> 1;
        if (pre[mid] <= c) lo = mid; else hi = mid - 1;
    }
    return lo;
}

__global__ void __launch_bounds__(kOptThreads)
grad_sumsq_kernel(const mft_tensor_slot* __restrict__ slots, int num_slots, double* __restrict__ partial) {
    pdl_sync();
    extern __shared__ int pre[];
    __shared__ double dred[32];
    const int total = chunk_prefix(slots, num_slots, pre, true);
    double acc = 0.0;
    for (int c = blockIdx.x; c < total; c += gridDim.x) {
        const int s = chunk_slot(pre, num_slots, c);
        const float* g = slots[s].grad;
        const int64_t n = slots[s].n, e0 = int64_t(c - pre[s]) * kOptChunk;
        if ((reinterpret_cast<uintptr_t>(g) & 15u) == 0) {
#pragma unroll
            for (int j = 0; j < 4; ++j) {
                const int64_t i = e0 + (j * kOptThreads + threadIdx.x) * 4;
                if (i + 3 < n) {
                    const float4 a = *reinterpret_cast<const float4*>(g + i);
                    acc += double(a.x) * a.x + double(a.y) * a.y + double(a.z) * a.z + double(a.w) * a.w;
                } else {
                    for (int64_t k = i; k < n; ++k) acc += double(g[k]) * g[k];
                }
            }
        } else {
            for (int j = 0; j < 16; ++j) {
                const int64_t i = e0 + j * kOptThreads + threadIdx.x;
                if (i < n) acc += double(g[i]) * g[i];
            }
        }
    }
    acc = block_sum<double>(acc, dred);
    if (threadIdx.x == 0) partial[blockIdx.x] = acc;
}

__global__ void __launch_bounds__(1024) sumsq_final_kernel(const double* __restrict__ partial, int parts, double* __restrict__ out) {
    pdl_sync();
    __shared__ double dred[32];
    double s = 0.0;
    for (int i = threadIdx.x; i < parts; i += blockDim.x) s += partial[i];
    s = block_sum<double>(s, dred);
    if (threadIdx.x == 0) out[0] = s;
}

__global__ void clip_scale_kernel(const double* __restrict__ sumsq, float max_norm, float pre_scale,
                                  float* __restrict__ scale_out, float* __restrict__ norm_out) {
    pdl_sync();
    // gpt2_lora_finetune/main.cpp:491-516: norm = sqrt(sum g^2) (double); if norm > max: scale = max/(norm+1e-6)
    double norm = sqrt(sumsq[0]) * double(pre_scale);
    float sc = 1.0f;
    if (max_norm > 0.f && norm > double(max_norm)) sc = float(double(max_norm) / (norm + 1e-6));
    if (scale_out) scale_out[0] = sc * pre_scale;
    if (norm_out) norm_out[0] = float(norm);
}

__global__ void __launch_bounds__(kOptThreads)
scale_grads_kernel(const mft_tensor_slot* __restrict__ slots, int num_slots, const float* __restrict__ scale) {
    pdl_sync();
    const float sc = scale[0];
    if (sc == 1.0f) return;
    const int64_t tid = blockIdx.x * int64_t(kOptThreads) + threadIdx.x;
    const int64_t stride = int64_t(gridDim.x) * kOptThreads;
    for (int s = 0; s < num_slots; ++s) {
        float* g = slots[s].grad;
        if (!g) continue;
        const int64_t n = slots[s].n;
        for (int64_t i = tid; i < n; i += stride) g[i] *= sc;
    }
}

__device__ __forceinline__ void adam_one(float& p, float g, float& m, float& v, float lr, float b1, float b2,
                                         float omb1, float omb2, float eps, float wd, float bc1, float bc2, float gs) {
    if (gs != 1.0f) g = __fmul_rn(g, gs);
    if (wd > 0.0f) g = __fadd_rn(g, __fmul_rn(wd, p));
    m = __fadd_rn(__fmul_rn(b1, m), __fmul_rn(omb1, g));
    v = __fadd_rn(__fmul_rn(b2, v), __fmul_rn(__fmul_rn(omb2, g), g));
    const float vc = __fdiv_rn(v, bc2);
    const float mc = __fdiv_rn(m, bc1);
    p = __fsub_rn(p, __fdiv_rn(__fmul_rn(lr, mc), __fadd_rn(__fsqrt_rn(vc), eps)));
}

// steps[s] += 1 for every slot that has a gradient (optim/adam.cpp:35-43: the per-tensor step only advances then)
__global__ void adam_advance_kernel(const mft_tensor_slot* __restrict__ slots, int num_slots, int32_t* __restrict__ steps) {
    pdl_sync();
    for (int s = blockIdx.x * blockDim.x + threadIdx.x; s < num_slots; s += gridDim.x * blockDim.x)
        if (slots[s].grad) steps[s] += 1;
}

__global__ void __launch_bounds__(kOptThreads)
adam_kernel(const mft_tensor_slot* __restrict__ slots, int num_slots, float lr, const float* __restrict__ lr_dev, float b1, float b2,
            float eps, float wd, const float* __restrict__ grad_scale, const int32_t* __restrict__ steps) {
    pdl_sync();
    extern __shared__ int pre[];
    const int total = chunk_prefix(slots, num_slots, pre, true);
    const float gs = grad_scale ? grad_scale[0] : 1.0f;
    if (lr_dev) lr = lr_dev[0];
    const float omb1 = 1.0f - b1, omb2 = 1.0f - b2;
    for (int c = blockIdx.x; c < total; c += gridDim.x) {
        const int s = chunk_slot(pre, num_slots, c);
        mft_tensor_slot sl = slots[s];
        if (steps) {   // device-resident step counters: bias corrections computed here (optim/adam.cpp:222-228)
            sl.bias_corr1 = 1.0f - powf(b1, float(steps[s]));
            sl.bias_corr2 = 1.0f - powf(b2, float(steps[s]));
        }
        const int64_t e0 = int64_t(c - pre[s]) * kOptChunk;
        const bool vec = ((reinterpret_cast<uintptr_t>(sl.param) | reinterpret_cast<uintptr_t>(sl.grad) |
                           reinterpret_cast<uintptr_t>(sl.m) | reinterpret_cast<uintptr_t>(sl.v)) & 15u) == 0;
        if (vec) {
            float4 p[4], g[4], m[4], v[4];
            bool full[4];
#pragma unroll
            for (int j = 0; j < 4; ++j) {      // all loads of the chunk first
                const int64_t i = e0 + (j * kOptThreads + threadIdx.x) * 4;
                full[j] = i + 3 < sl.n;
                if (full[j]) {
                    p[j] = *reinterpret_cast<const float4*>(sl.param + i); g[j] = *reinterpret_cast<const float4*>(sl.grad + i);
                    m[j] = *reinterpret_cast<const float4*>(sl.m + i); v[j] = *reinterpret_cast<const float4*>(sl.v + i);
                }
            }
#pragma unroll
            for (int j = 0; j < 4; ++j) {
                const int64_t i = e0 + (j * kOptThreads + threadIdx.x) * 4;
                if (full[j]) {
                    adam_one(p[j].x, g[j].x, m[j].x, v[j].x, lr, b1, b2, omb1, omb2, eps, wd, sl.bias_corr1, sl.bias_corr2, gs);
                    adam_one(p[j].y, g[j].y, m[j].y, v[j].y, lr, b1, b2, omb1, omb2, eps, wd, sl.bias_corr1, sl.bias_corr2, gs);
                    adam_one(p[j].z, g[j].z, m[j].z, v[j].z, lr, b1, b2, omb1, omb2, eps, wd, sl.bias_corr1, sl.bias_corr2, gs);
                    adam_one(p[j].w, g[j].w, m[j].w, v[j].w, lr, b1, b2, omb1, omb2, eps, wd, sl.bias_corr1, sl.bias_corr2, gs);
                    *reinterpret_cast<float4*>(sl.param + i) = p[j]; *reinterpret_cast<float4*>(sl.m + i) = m[j];
                    *reinterpret_cast<float4*>(sl.v + i) = v[j];
                } else {
                    for (int64_t k = i; k < sl.n; ++k)
                        adam_one(sl.param[k], sl.grad[k], sl.m[k], sl.v[k], lr, b1, b2, omb1, omb2, eps, wd, sl.bias_corr1, sl.bias_corr2, gs);
                }
            }
        } else {
            for (int j = 0; j < 16; ++j) {
                const int64_t i = e0 + j * kOptThreads + threadIdx.x;
                if (i < sl.n) adam_one(sl.param[i], sl.grad[i], sl.m[i], sl.v[i], lr, b1, b2, omb1, omb2, eps, wd, sl.bias_corr1, sl.bias_corr2, gs);
            }
        }
    }
}

}  // namespace mft

using namespace mft;

namespace mft {
// bucket[offset + i] = grad[i] (0 where a tensor has no gradient): one launch packs every gradient for the all-reduce
__global__ void __launch_bounds__(kOptThreads)
pack_grads_kernel(const mft_tensor_slot* __restrict__ slots, int num_slots, float* __restrict__ bucket) {
    pdl_sync();
    extern __shared__ int pre[];
    const int total = chunk_prefix(slots, num_slots, pre, false);
    for (int c = blockIdx.x; c < total; c += gridDim.x) {
        const int s = chunk_slot(pre, num_slots, c);
        const float* g = slots[s].grad;
        float* dst = bucket + slots[s].bucket_offset;
        const int64_t n = slots[s].n, e0 = int64_t(c - pre[s]) * kOptChunk;
#pragma unroll 4
        for (int j = 0; j < 16; ++j) {
            const int64_t i = e0 + j * kOptThreads + threadIdx.x;
            if (i < n) dst[i] = g ? g[i] : 0.f;
        }
    }
}
}  // namespace mft

extern "C" int mft_pack_grads(const mft_tensor_slot* slots, int32_t num_slots, float* bucket, mft_stream_t stream) {
    if (num_slots == 0) return 0;
    if (!slots || !bucket) return mft::set_error(1, "mft_pack_grads: null pointer");
    if (num_slots > mft::kOptMaxSlots) return mft::set_error(1, "mft_pack_grads: more than %d tensors", mft::kOptMaxSlots);
    pdl_launch(mft::pack_grads_kernel, mft::kOptCtas, mft::kOptThreads, size_t(num_slots + 1) * sizeof(int), mft::as_stream(stream))(slots, num_slots, bucket);
    MFT_CHECK_LAUNCH("mft_pack_grads");
    return 0;
}

extern "C" size_t mft_grad_sumsq_workspace_bytes(void) { return size_t(kOptCtas) * sizeof(double); }

extern "C" int mft_grad_sumsq(const mft_tensor_slot* slots, int32_t num_slots, double* out, void* workspace,
                              size_t workspace_bytes, mft_stream_t stream) {
    if (!out) return set_error(1, "mft_grad_sumsq: null output");
    if (num_slots > 0 && !slots) return set_error(1, "mft_grad_sumsq: null slot table");
    if (!workspace || workspace_bytes < mft_grad_sumsq_workspace_bytes())
        return set_error(1, "mft_grad_sumsq: workspace too small");
    cudaStream_t st = as_stream(stream);
    double* partial = static_cast<double*>(workspace);
    if (num_slots > kOptMaxSlots) return set_error(1, "mft_grad_sumsq: more than %d tensors", kOptMaxSlots);
    pdl_launch(grad_sumsq_kernel, kOptCtas, kOptThreads, size_t(num_slots + 1) * sizeof(int), st)(slots, num_slots, partial);
    pdl_launch(sumsq_final_kernel, 1, 1024, 0, st)(partial, kOptCtas, out);
    add_launches(1);
    MFT_CHECK_LAUNCH("mft_grad_sumsq");
    return 0;
}

extern "C" int mft_clip_scale(const double* sumsq, float max_norm, float pre_scale, float* scale_out, float* norm_out,
                              mft_stream_t stream) {
    if (!sumsq) return set_error(1, "mft_clip_scale: null pointer");
    pdl_launch(clip_scale_kernel, 1, 1, 0, as_stream(stream))(sumsq, max_norm, pre_scale, scale_out, norm_out);
    MFT_CHECK_LAUNCH("mft_clip_scale");
    return 0;
}

extern "C" int mft_scale_grads(const mft_tensor_slot* slots, int32_t num_slots, const float* scale, mft_stream_t stream) {
    if (num_slots == 0) return 0;
    if (!slots || !scale) return set_error(1, "mft_scale_grads: null pointer");
    pdl_launch(scale_grads_kernel, kOptCtas, kOptThreads, 0, as_stream(stream))(slots, num_slots, scale);
    MFT_CHECK_LAUNCH("mft_scale_grads");
    return 0;
}

extern "C" int mft_adam_step(const mft_tensor_slot* slots, int32_t num_slots, float lr, float beta1, float beta2,
                             float eps, float weight_decay, const float* grad_scale, mft_stream_t stream) {
    if (num_slots == 0) return 0;
    if (!slots) return set_error(1, "mft_adam_step: null slot table");
    if (num_slots > kOptMaxSlots) return set_error(1, "mft_adam_step: more than %d tensors", kOptMaxSlots);
    pdl_launch(adam_kernel, kOptCtas, kOptThreads, size_t(num_slots + 1) * sizeof(int), as_stream(stream))(slots, num_slots, lr, nullptr, beta1, beta2, eps, weight_decay, grad_scale, nullptr);
    MFT_CHECK_LAUNCH("mft_adam_step");
    return 0;
}

extern "C" int mft_adam_step_dev(const mft_tensor_slot* slots, int32_t num_slots, const float* lr, float beta1, float beta2,
                                 float eps, float weight_decay, const float* grad_scale, int32_t* steps, mft_stream_t stream) {
    if (num_slots == 0) return 0;
    if (!slots || !steps || !lr) return set_error(1, "mft_adam_step_dev: null pointer");
    if (num_slots > kOptMaxSlots) return set_error(1, "mft_adam_step_dev: more than %d tensors", kOptMaxSlots);
    pdl_launch(adam_advance_kernel, unsigned(ceil_div(num_slots, 256)), 256, 0, as_stream(stream))(slots, num_slots, steps);
    add_launches(1);
    pdl_launch(adam_kernel, kOptCtas, kOptThreads, size_t(num_slots + 1) * sizeof(int), as_stream(stream))(slots, num_slots, 0.0f, lr, beta1, beta2, eps, weight_decay, grad_scale, steps);
    MFT_CHECK_LAUNCH("mft_adam_step_dev");
    return 0;
}

// ---------------------------------------------------------------- bf16 shadows of trained matrices, all in one launch
// After the optimizer step every trained 2-D tensor that feeds a bf16 GEMM operand needs its bf16 copy (same layout) and /
// or its K-major bf16 transpose again. One launch over a slot table does all of them (blockIdx.y = tensor), instead of one
// cast / transpose-cast launch per tensor per step (96 adapter factors = 4 tiny launches each in the Gemma configuration).
namespace mft {
__global__ void refresh_shadows_kernel(const mft_shadow_slot* __restrict__ slots) {
    pdl_sync();
    const mft_shadow_slot s = slots[blockIdx.y];
    const int64_t n = s.rows * s.cols;
    for (int64_t i = blockIdx.x * int64_t(blockDim.x) + threadIdx.x; i < n; i += int64_t(gridDim.x) * blockDim.x) {
        const int64_t r = i / s.cols, c = i % s.cols;
        const __nv_bfloat16 v = __float2bfloat16_rn(s.src[r * s.ld_src + c]);
        if (s.plain) static_cast<__nv_bfloat16*>(s.plain)[r * s.ld_plain + c] = v;
        if (s.transposed) static_cast<__nv_bfloat16*>(s.transposed)[c * s.ld_transposed + r] = v;
    }
}
}  // namespace mft

extern "C" int mft_refresh_bf16_shadows(const mft_shadow_slot* slots, int32_t num_slots, int64_t max_elems, mft_stream_t stream) {
    if (num_slots == 0) return 0;
    if (!slots) return set_error(1, "mft_refresh_bf16_shadows: null slot table");
    if (num_slots > 65535) return set_error(1, "mft_refresh_bf16_shadows: more than 65535 tensors");
    int64_t bx = ceil_div(max_elems > 0 ? max_elems : 1, int64_t(256) * 8);
    if (bx < 1) bx = 1;
    if (bx > 64) bx = 64;
    pdl_launch(mft::refresh_shadows_kernel, dim3(unsigned(bx), unsigned(num_slots)), 256, 0, as_stream(stream))(slots);
    MFT_CHECK_LAUNCH("mft_refresh_bf16_shadows");
    return 0;
}
