// Multi-tensor gradient-norm / clip / Adam. One launch covers every trainable tensor through a slot
// table in device memory. Arithmetic follows optim/adam.cpp:25-90 operation by operation (explicit
// round-to-nearest mul/add so nvcc cannot contract into FMAs the CPU reference does not use).
// HBM-bound: sumsq reads 4 B/param, Adam moves 28 B/param (read p,g,m,v; write p,m,v).
#include "common.cuh"

namespace mft {

constexpr int kOptCtas = kNumSMs * 4;
constexpr int kOptThreads = 256;

__global__ void __launch_bounds__(kOptThreads)
grad_sumsq_kernel(const mft_tensor_slot* __restrict__ slots, int num_slots, double* __restrict__ partial) {
    __shared__ double dred[32];
    double acc = 0.0;
    const int64_t tid = blockIdx.x * int64_t(kOptThreads) + threadIdx.x;
    const int64_t stride = int64_t(gridDim.x) * kOptThreads;
    for (int s = 0; s < num_slots; ++s) {
        const float* g = slots[s].grad;
        const int64_t n = slots[s].n;
        if (!g) continue;
        if ((reinterpret_cast<uintptr_t>(g) & 15u) == 0) {
            const float4* g4 = reinterpret_cast<const float4*>(g);
            const int64_t n4 = n >> 2;
            for (int64_t i = tid; i < n4; i += stride) {
                float4 a = g4[i];
                acc += double(a.x) * a.x + double(a.y) * a.y + double(a.z) * a.z + double(a.w) * a.w;
            }
            for (int64_t i = (n4 << 2) + tid; i < n; i += stride) acc += double(g[i]) * g[i];
        } else {
            for (int64_t i = tid; i < n; i += stride) acc += double(g[i]) * g[i];
        }
    }
    acc = block_sum<double>(acc, dred);
    if (threadIdx.x == 0) partial[blockIdx.x] = acc;
}

__global__ void __launch_bounds__(1024) sumsq_final_kernel(const double* __restrict__ partial, int parts, double* __restrict__ out) {
    __shared__ double dred[32];
    double s = 0.0;
    for (int i = threadIdx.x; i < parts; i += blockDim.x) s += partial[i];
    s = block_sum<double>(s, dred);
    if (threadIdx.x == 0) out[0] = s;
}

__global__ void clip_scale_kernel(const double* __restrict__ sumsq, float max_norm, float pre_scale,
                                  float* __restrict__ scale_out, float* __restrict__ norm_out) {
    // gpt2_lora_finetune/main.cpp:491-516: norm = sqrt(sum g^2) (double); if norm > max: scale = max/(norm+1e-6)
    double norm = sqrt(sumsq[0]) * double(pre_scale);
    float sc = 1.0f;
    if (max_norm > 0.f && norm > double(max_norm)) sc = float(double(max_norm) / (norm + 1e-6));
    if (scale_out) scale_out[0] = sc * pre_scale;
    if (norm_out) norm_out[0] = float(norm);
}

__global__ void __launch_bounds__(kOptThreads)
scale_grads_kernel(const mft_tensor_slot* __restrict__ slots, int num_slots, const float* __restrict__ scale) {
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
    for (int s = blockIdx.x * blockDim.x + threadIdx.x; s < num_slots; s += gridDim.x * blockDim.x)
        if (slots[s].grad) steps[s] += 1;
}

__global__ void __launch_bounds__(kOptThreads)
adam_kernel(const mft_tensor_slot* __restrict__ slots, int num_slots, float lr, const float* __restrict__ lr_dev, float b1, float b2,
            float eps, float wd, const float* __restrict__ grad_scale, const int32_t* __restrict__ steps) {
    const float gs = grad_scale ? grad_scale[0] : 1.0f;
    if (lr_dev) lr = lr_dev[0];
    const float omb1 = 1.0f - b1, omb2 = 1.0f - b2;
    const int64_t tid = blockIdx.x * int64_t(kOptThreads) + threadIdx.x;
    const int64_t stride = int64_t(gridDim.x) * kOptThreads;
    for (int s = 0; s < num_slots; ++s) {
        mft_tensor_slot sl = slots[s];
        if (!sl.grad) continue;
        if (steps) {   // device-resident step counters: bias corrections computed here (optim/adam.cpp:222-228)
            sl.bias_corr1 = 1.0f - powf(b1, float(steps[s]));
            sl.bias_corr2 = 1.0f - powf(b2, float(steps[s]));
        }
        const bool vec = ((reinterpret_cast<uintptr_t>(sl.param) | reinterpret_cast<uintptr_t>(sl.grad) |
                           reinterpret_cast<uintptr_t>(sl.m) | reinterpret_cast<uintptr_t>(sl.v)) & 15u) == 0;
        if (vec) {
            float4* p4 = reinterpret_cast<float4*>(sl.param);
            const float4* g4 = reinterpret_cast<const float4*>(sl.grad);
            float4* m4 = reinterpret_cast<float4*>(sl.m);
            float4* v4 = reinterpret_cast<float4*>(sl.v);
            const int64_t n4 = sl.n >> 2;
            for (int64_t i = tid; i < n4; i += stride) {
                float4 p = p4[i], g = g4[i], m = m4[i], v = v4[i];
                adam_one(p.x, g.x, m.x, v.x, lr, b1, b2, omb1, omb2, eps, wd, sl.bias_corr1, sl.bias_corr2, gs);
                adam_one(p.y, g.y, m.y, v.y, lr, b1, b2, omb1, omb2, eps, wd, sl.bias_corr1, sl.bias_corr2, gs);
                adam_one(p.z, g.z, m.z, v.z, lr, b1, b2, omb1, omb2, eps, wd, sl.bias_corr1, sl.bias_corr2, gs);
                adam_one(p.w, g.w, m.w, v.w, lr, b1, b2, omb1, omb2, eps, wd, sl.bias_corr1, sl.bias_corr2, gs);
                p4[i] = p; m4[i] = m; v4[i] = v;
            }
            for (int64_t i = (n4 << 2) + tid; i < sl.n; i += stride)
                adam_one(sl.param[i], sl.grad[i], sl.m[i], sl.v[i], lr, b1, b2, omb1, omb2, eps, wd, sl.bias_corr1, sl.bias_corr2, gs);
        } else {
            for (int64_t i = tid; i < sl.n; i += stride)
                adam_one(sl.param[i], sl.grad[i], sl.m[i], sl.v[i], lr, b1, b2, omb1, omb2, eps, wd, sl.bias_corr1, sl.bias_corr2, gs);
        }
    }
}

}  // namespace mft

using namespace mft;

namespace mft {
// bucket[offset + i] = grad[i] (0 where a tensor has no gradient): one launch packs every gradient for the all-reduce
__global__ void __launch_bounds__(kOptThreads)
pack_grads_kernel(const mft_tensor_slot* __restrict__ slots, int num_slots, float* __restrict__ bucket) {
    const int64_t tid = blockIdx.x * int64_t(kOptThreads) + threadIdx.x;
    const int64_t stride = int64_t(gridDim.x) * kOptThreads;
    for (int s = 0; s < num_slots; ++s) {
        const float* g = slots[s].grad;
        float* dst = bucket + slots[s].bucket_offset;
        const int64_t n = slots[s].n;
        for (int64_t i = tid; i < n; i += stride) dst[i] = g ? g[i] : 0.f;
    }
}
}  // namespace mft

extern "C" int mft_pack_grads(const mft_tensor_slot* slots, int32_t num_slots, float* bucket, mft_stream_t stream) {
    if (num_slots == 0) return 0;
    if (!slots || !bucket) return mft::set_error(1, "mft_pack_grads: null pointer");
    mft::pack_grads_kernel<<<mft::kOptCtas, mft::kOptThreads, 0, mft::as_stream(stream)>>>(slots, num_slots, bucket);
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
    grad_sumsq_kernel<<<kOptCtas, kOptThreads, 0, st>>>(slots, num_slots, partial);
    sumsq_final_kernel<<<1, 1024, 0, st>>>(partial, kOptCtas, out);
    add_launches(1);
    MFT_CHECK_LAUNCH("mft_grad_sumsq");
    return 0;
}

extern "C" int mft_clip_scale(const double* sumsq, float max_norm, float pre_scale, float* scale_out, float* norm_out,
                              mft_stream_t stream) {
    if (!sumsq) return set_error(1, "mft_clip_scale: null pointer");
    clip_scale_kernel<<<1, 1, 0, as_stream(stream)>>>(sumsq, max_norm, pre_scale, scale_out, norm_out);
    MFT_CHECK_LAUNCH("mft_clip_scale");
    return 0;
}

extern "C" int mft_scale_grads(const mft_tensor_slot* slots, int32_t num_slots, const float* scale, mft_stream_t stream) {
    if (num_slots == 0) return 0;
    if (!slots || !scale) return set_error(1, "mft_scale_grads: null pointer");
    scale_grads_kernel<<<kOptCtas, kOptThreads, 0, as_stream(stream)>>>(slots, num_slots, scale);
    MFT_CHECK_LAUNCH("mft_scale_grads");
    return 0;
}

extern "C" int mft_adam_step(const mft_tensor_slot* slots, int32_t num_slots, float lr, float beta1, float beta2,
                             float eps, float weight_decay, const float* grad_scale, mft_stream_t stream) {
    if (num_slots == 0) return 0;
    if (!slots) return set_error(1, "mft_adam_step: null slot table");
    adam_kernel<<<kOptCtas, kOptThreads, 0, as_stream(stream)>>>(slots, num_slots, lr, nullptr, beta1, beta2, eps, weight_decay, grad_scale, nullptr);
    MFT_CHECK_LAUNCH("mft_adam_step");
    return 0;
}

extern "C" int mft_adam_step_dev(const mft_tensor_slot* slots, int32_t num_slots, const float* lr, float beta1, float beta2,
                                 float eps, float weight_decay, const float* grad_scale, int32_t* steps, mft_stream_t stream) {
    if (num_slots == 0) return 0;
    if (!slots || !steps || !lr) return set_error(1, "mft_adam_step_dev: null pointer");
    adam_advance_kernel<<<unsigned(ceil_div(num_slots, 256)), 256, 0, as_stream(stream)>>>(slots, num_slots, steps);
    add_launches(1);
    adam_kernel<<<kOptCtas, kOptThreads, 0, as_stream(stream)>>>(slots, num_slots, 0.0f, lr, beta1, beta2, eps, weight_decay, grad_scale, steps);
    MFT_CHECK_LAUNCH("mft_adam_step_dev");
    return 0;
}
