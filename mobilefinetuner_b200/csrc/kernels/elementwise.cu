#include <stdlib.h>
// HBM-streaming elementwise / layout kernels: 128-bit vectorised, grid-stride, coalesced.
// Each is bound by HBM bandwidth; algorithmic bytes per element are stated per kernel.
#include <cuda_bf16.h>
#include "common.cuh"

namespace mft {

bool pdl_enabled() {
    static const bool on = [] { const char* e = getenv("MFT_PDL"); return e && *e == '1'; }();      // off by default, see common.cuh
    return on;
}


// ---------------------------------------------------------------- binary with broadcast
template <int OP>
__device__ __forceinline__ float bin_apply(float a, float b) {
    if (OP == 0) return a + b;
    if (OP == 1) return a - b;
    if (OP == 2) return a * b;
    if (OP == 3) return a / b;
    if (OP == 4) return a == b ? 1.f : 0.f;
    if (OP == 5) return a != b ? 1.f : 0.f;
    if (OP == 6) return a > b ? 1.f : 0.f;
    if (OP == 7) return a < b ? 1.f : 0.f;
    if (OP == 8) return a >= b ? 1.f : 0.f;
    return a <= b ? 1.f : 0.f;
}

// b_mode 0: same shape (12 B/elem); vectorised when n % 4 == 0 and pointers aligned.
template <int OP>
__global__ void binary_same_vec4(const float4* __restrict__ a, const float4* __restrict__ b,
                                 float4* __restrict__ out, int64_t n4) {
    pdl_sync();
    for (int64_t i = blockIdx.x * int64_t(blockDim.x) + threadIdx.x; i < n4; i += int64_t(gridDim.x) * blockDim.x) {
        float4 x = ld_stream(a + i), y = ld_stream(b + i), r;
        r.x = bin_apply<OP>(x.x, y.x); r.y = bin_apply<OP>(x.y, y.y);
        r.z = bin_apply<OP>(x.z, y.z); r.w = bin_apply<OP>(x.w, y.w);
        st_stream(out + i, r);
    }
}

// General broadcast path (scalar): out[o,m,i] = a[o,m,i] op b[f(o,m,i)].
template <int OP>
__global__ void binary_bcast(const float* __restrict__ a, const float* __restrict__ b, float* __restrict__ out,
                             int64_t outer, int64_t mid, int64_t inner, int b_mode, int64_t b_div) {
    pdl_sync();
    const int64_t n = outer * mid * inner;
    for (int64_t idx = blockIdx.x * int64_t(blockDim.x) + threadIdx.x; idx < n; idx += int64_t(gridDim.x) * blockDim.x) {
        int64_t i = idx % inner;
        int64_t om = idx / inner;
        int64_t m = om % mid, o = om / mid;
        int64_t bi;
        if (b_mode == 0) bi = idx;
        else if (b_mode == 1) bi = i;
        else if (b_mode == 2) bi = m * inner + i;
        else bi = (o / b_div) * inner + i;
        out[idx] = bin_apply<OP>(a[idx], __ldg(b + bi));
    }
}

// Bias broadcast, vectorised: rows x inner with inner % 4 == 0.
template <int OP>
__global__ void binary_bias_vec4(const float4* __restrict__ a, const float4* __restrict__ b,
                                 float4* __restrict__ out, int64_t n4, int64_t inner4) {
    pdl_sync();
    for (int64_t i = blockIdx.x * int64_t(blockDim.x) + threadIdx.x; i < n4; i += int64_t(gridDim.x) * blockDim.x) {
        float4 x = ld_stream(a + i), y = __ldg(b + (i % inner4)), r;
        r.x = bin_apply<OP>(x.x, y.x); r.y = bin_apply<OP>(x.y, y.y);
        r.z = bin_apply<OP>(x.z, y.z); r.w = bin_apply<OP>(x.w, y.w);
        st_stream(out + i, r);
    }
}

template <int OP>
static int launch_binary(const float* a, const float* b, float* out, int64_t outer, int64_t mid, int64_t inner,
                         int b_mode, int64_t b_div, cudaStream_t st) {
    const int64_t n = outer * mid * inner;
    if (n == 0) return 0;
    const bool al = aligned16(a) && aligned16(b) && aligned16(out);
    if (b_mode == 0 && (n % 4 == 0) && al) {
        int64_t n4 = n / 4;
        pdl_launch(binary_same_vec4<OP>, stream_grid(n4, 256), 256, 0, st)(
            reinterpret_cast<const float4*>(a), reinterpret_cast<const float4*>(b), reinterpret_cast<float4*>(out), n4);
    } else if (b_mode == 1 && (inner % 4 == 0) && al) {
        int64_t n4 = n / 4;
        pdl_launch(binary_bias_vec4<OP>, stream_grid(n4, 256), 256, 0, st)(
            reinterpret_cast<const float4*>(a), reinterpret_cast<const float4*>(b), reinterpret_cast<float4*>(out), n4,
            inner / 4);
    } else {
        pdl_launch(binary_bcast<OP>, stream_grid(n, 256), 256, 0, st)(a, b, out, outer, mid, inner, b_mode, b_div < 1 ? 1 : b_div);
    }
    return 0;
}

// ---------------------------------------------------------------- affine / accumulate / fill
__global__ void affine_kernel(const float* __restrict__ a, float* __restrict__ out, int64_t n, float alpha, float beta) {
    pdl_sync();
    const int64_t n4 = n >> 2;
    const bool vec = ((reinterpret_cast<uintptr_t>(a) | reinterpret_cast<uintptr_t>(out)) & 15u) == 0;
    int64_t tid = blockIdx.x * int64_t(blockDim.x) + threadIdx.x, stride = int64_t(gridDim.x) * blockDim.x;
    if (vec) {
        const float4* a4 = reinterpret_cast<const float4*>(a);
        float4* o4 = reinterpret_cast<float4*>(out);
        for (int64_t i = tid; i < n4; i += stride) {
            float4 x = ld_stream(a4 + i);
            x.x = x.x * alpha + beta; x.y = x.y * alpha + beta; x.z = x.z * alpha + beta; x.w = x.w * alpha + beta;
            st_stream(o4 + i, x);
        }
        for (int64_t i = n4 * 4 + tid; i < n; i += stride) out[i] = a[i] * alpha + beta;
    } else {
        for (int64_t i = tid; i < n; i += stride) out[i] = a[i] * alpha + beta;
    }
}

__global__ void accumulate_kernel(float* __restrict__ acc, const float* __restrict__ g, int64_t n) {
    pdl_sync();
    const int64_t n4 = n >> 2;
    const bool vec = ((reinterpret_cast<uintptr_t>(acc) | reinterpret_cast<uintptr_t>(g)) & 15u) == 0;
    int64_t tid = blockIdx.x * int64_t(blockDim.x) + threadIdx.x, stride = int64_t(gridDim.x) * blockDim.x;
    if (vec) {
        float4* a4 = reinterpret_cast<float4*>(acc);
        const float4* g4 = reinterpret_cast<const float4*>(g);
        for (int64_t i = tid; i < n4; i += stride) {
            float4 x = a4[i], y = ld_stream(g4 + i);
            x.x += y.x; x.y += y.y; x.z += y.z; x.w += y.w;
            a4[i] = x;
        }
        for (int64_t i = n4 * 4 + tid; i < n; i += stride) acc[i] += g[i];
    } else {
        for (int64_t i = tid; i < n; i += stride) acc[i] += g[i];
    }
}

__global__ void fill_kernel(float* __restrict__ out, int64_t n, float v) {
    pdl_sync();
    for (int64_t i = blockIdx.x * int64_t(blockDim.x) + threadIdx.x; i < n; i += int64_t(gridDim.x) * blockDim.x) out[i] = v;
}

// ---------------------------------------------------------------- unary / gelu / mul backward
template <int KIND>
__device__ __forceinline__ float unary_apply(float x) {
    if (KIND == 0) return gelu_tanh_f(x);
    if (KIND == 1) return x > 0.f ? x : 0.f;
    if (KIND == 2) return 1.0f / (1.0f + expf(-x));
    if (KIND == 3) return x / (1.0f + expf(-x));
    if (KIND == 4) return tanhf(x);
    if (KIND == 5) return expf(x);
    if (KIND == 6) return logf(x);
    if (KIND == 7) return sqrtf(x);
    return fabsf(x);
}
template <int KIND>
__global__ void unary_kernel(const float* __restrict__ x, float* __restrict__ y, int64_t n) {
    pdl_sync();
    const int64_t n4 = n >> 2;
    const bool vec = ((reinterpret_cast<uintptr_t>(x) | reinterpret_cast<uintptr_t>(y)) & 15u) == 0;
    int64_t tid = blockIdx.x * int64_t(blockDim.x) + threadIdx.x, stride = int64_t(gridDim.x) * blockDim.x;
    if (vec) {
        const float4* x4 = reinterpret_cast<const float4*>(x);
        float4* y4 = reinterpret_cast<float4*>(y);
        for (int64_t i = tid; i < n4; i += stride) {
            float4 v = ld_stream(x4 + i);
            v.x = unary_apply<KIND>(v.x); v.y = unary_apply<KIND>(v.y);
            v.z = unary_apply<KIND>(v.z); v.w = unary_apply<KIND>(v.w);
            st_stream(y4 + i, v);
        }
        for (int64_t i = n4 * 4 + tid; i < n; i += stride) y[i] = unary_apply<KIND>(x[i]);
    } else {
        for (int64_t i = tid; i < n; i += stride) y[i] = unary_apply<KIND>(x[i]);
    }
}

// gx = gy * gelu'(x): reads 8 B, writes 4 B per element.
__global__ void gelu_bwd_kernel(const float* __restrict__ x, const float* __restrict__ gy, float* __restrict__ gx, int64_t n,
                                __nv_bfloat16* __restrict__ gx16) {
    pdl_sync();   // gx16 != NULL (dense, 8-byte aligned, n % 4 == 0): bf16 copy of gx
    const int64_t n4 = n >> 2;
    const bool vec = ((reinterpret_cast<uintptr_t>(x) | reinterpret_cast<uintptr_t>(gy) | reinterpret_cast<uintptr_t>(gx)) & 15u) == 0;
    int64_t tid = blockIdx.x * int64_t(blockDim.x) + threadIdx.x, stride = int64_t(gridDim.x) * blockDim.x;
    if (vec) {
        const float4* x4 = reinterpret_cast<const float4*>(x);
        const float4* g4 = reinterpret_cast<const float4*>(gy);
        float4* o4 = reinterpret_cast<float4*>(gx);
        for (int64_t i = tid; i < n4; i += stride) {
            float4 v = ld_stream(x4 + i), g = ld_stream(g4 + i), r;
            r.x = g.x * gelu_tanh_grad_f(v.x); r.y = g.y * gelu_tanh_grad_f(v.y);
            r.z = g.z * gelu_tanh_grad_f(v.z); r.w = g.w * gelu_tanh_grad_f(v.w);
            st_stream(o4 + i, r);
            if (gx16) st_bf16x4(gx16 + i * 4, r);
        }
        for (int64_t i = n4 * 4 + tid; i < n; i += stride) gx[i] = gy[i] * gelu_tanh_grad_f(x[i]);
    } else {
        for (int64_t i = tid; i < n; i += stride) gx[i] = gy[i] * gelu_tanh_grad_f(x[i]);
    }
}

// gx = gy * gelu'(x) with gy and gx as bf16 (the gradient arrives from a bf16-only GEMM result and feeds a bf16 GEMM
// operand): x is read as fp32 (XBF16 = false: 4 B) or as the bf16 pre-activation (XBF16 = true: 2 B), plus 2 B read and 2 B
// written per element. gx32 != NULL additionally writes the fp32 gradient. n % 8 == 0, 16-byte aligned.
// gelu'(x) with the hardware tanh (MUFU.TANH, relative error ~2^-11): only where the result is rounded to bf16 anyway
__device__ __forceinline__ float gelu_tanh_grad_fast(float x) {
    const float k0 = 0.7978845608f, k1 = 0.044715f;
    const float x2 = x * x;
    float t;
    asm("tanh.approx.f32 %0, %1;" : "=f"(t) : "f"(k0 * fmaf(k1 * x, x2, x)));
    const float dinner = k0 * fmaf(3.0f * k1, x2, 1.0f);
    return fmaf(0.5f * x * fmaf(-t, t, 1.0f), dinner, fmaf(0.5f, t, 0.5f));
}

template <bool XBF16>
__device__ __forceinline__ void gelu_bwd_lowp_load(const void* xv, const uint4* g8, int64_t i, float (&xs)[8], uint4& gr) {
    if (XBF16) {
        const uint4 xr = __ldcs(reinterpret_cast<const uint4*>(xv) + i);
        const __nv_bfloat162* xp = reinterpret_cast<const __nv_bfloat162*>(&xr);
#pragma unroll
        for (int e = 0; e < 4; ++e) { const float2 f = __bfloat1622float2(xp[e]); xs[2 * e] = f.x; xs[2 * e + 1] = f.y; }
    } else {
        const float4 a = ld_stream(reinterpret_cast<const float4*>(xv) + 2 * i), b = ld_stream(reinterpret_cast<const float4*>(xv) + 2 * i + 1);
        xs[0] = a.x; xs[1] = a.y; xs[2] = a.z; xs[3] = a.w; xs[4] = b.x; xs[5] = b.y; xs[6] = b.z; xs[7] = b.w;
    }
    gr = __ldcs(g8 + i);
}
template <bool FAST>      // FAST: bf16 is the only result (the fp32 copy, when asked for, keeps the exact tanhf)
__device__ __forceinline__ void gelu_bwd_lowp_store(const float (&xs)[8], const uint4& gr, float* gx32, uint4* o8, int64_t i) {
    const __nv_bfloat162* gp = reinterpret_cast<const __nv_bfloat162*>(&gr);
    float r[8];
#pragma unroll
    for (int e = 0; e < 4; ++e) {
        const float2 g = __bfloat1622float2(gp[e]);
        r[2 * e] = g.x * (FAST ? gelu_tanh_grad_fast(xs[2 * e]) : gelu_tanh_grad_f(xs[2 * e]));
        r[2 * e + 1] = g.y * (FAST ? gelu_tanh_grad_fast(xs[2 * e + 1]) : gelu_tanh_grad_f(xs[2 * e + 1]));
    }
    uint4 out;
    __nv_bfloat162* op = reinterpret_cast<__nv_bfloat162*>(&out);
#pragma unroll
    for (int e = 0; e < 4; ++e) op[e] = __floats2bfloat162_rn(r[2 * e], r[2 * e + 1]);
    __stcs(o8 + i, out);
    if (!FAST && gx32) {
        st_stream(reinterpret_cast<float4*>(gx32) + 2 * i, make_float4(r[0], r[1], r[2], r[3]));
        st_stream(reinterpret_cast<float4*>(gx32) + 2 * i + 1, make_float4(r[4], r[5], r[6], r[7]));
    }
}

// Two 8-element groups per thread and iteration: both groups' loads are issued before either is used (bytes in flight).
template <bool XBF16, bool FAST>
__global__ void __launch_bounds__(256)
gelu_bwd_lowp_kernel(const void* __restrict__ xv, const __nv_bfloat16* __restrict__ gy, float* __restrict__ gx32,
                     __nv_bfloat16* __restrict__ gx16, int64_t n) {
    pdl_sync();
    const int64_t n8 = n >> 3, stride = int64_t(gridDim.x) * blockDim.x;
    const uint4* g8 = reinterpret_cast<const uint4*>(gy);
    uint4* o8 = reinterpret_cast<uint4*>(gx16);
    int64_t i = blockIdx.x * int64_t(blockDim.x) + threadIdx.x;
    for (; i + stride < n8; i += 2 * stride) {
        float xa[8], xb[8];
        uint4 ga, gb;
        gelu_bwd_lowp_load<XBF16>(xv, g8, i, xa, ga);
        gelu_bwd_lowp_load<XBF16>(xv, g8, i + stride, xb, gb);
        gelu_bwd_lowp_store<FAST>(xa, ga, gx32, o8, i);
        gelu_bwd_lowp_store<FAST>(xb, gb, gx32, o8, i + stride);
    }
    if (i < n8) {
        float xa[8];
        uint4 ga;
        gelu_bwd_lowp_load<XBF16>(xv, g8, i, xa, ga);
        gelu_bwd_lowp_store<FAST>(xa, ga, gx32, o8, i);
    }
}

// bf16 -> fp32, row pitches in elements (fills the fp32 buffer of a tensor that was produced as bf16 only)
__global__ void cast_from_bf16_kernel(const __nv_bfloat16* __restrict__ in, int64_t ldi, float* __restrict__ out, int64_t ldo, int64_t rows, int64_t cols) {
    pdl_sync();
    const int64_t c4 = (cols + 3) >> 2, n = rows * c4;
    const bool vec = ((ldi & 3) == 0) && ((ldo & 3) == 0) && ((reinterpret_cast<uintptr_t>(in) & 7u) == 0) && ((reinterpret_cast<uintptr_t>(out) & 15u) == 0);
    for (int64_t idx = blockIdx.x * int64_t(blockDim.x) + threadIdx.x; idx < n; idx += int64_t(gridDim.x) * blockDim.x) {
        const int64_t r = idx / c4, c = (idx % c4) << 2;
        const __nv_bfloat16* src = in + r * ldi + c;
        float* dst = out + r * ldo + c;
        if (vec && c + 4 <= cols) {
            const uint2 v = *reinterpret_cast<const uint2*>(src);
            const float2 a = __bfloat1622float2(*reinterpret_cast<const __nv_bfloat162*>(&v.x)), b = __bfloat1622float2(*reinterpret_cast<const __nv_bfloat162*>(&v.y));
            *reinterpret_cast<float4*>(dst) = make_float4(a.x, a.y, b.x, b.y);
        } else {
            for (int e = 0; e < 4 && c + e < cols; ++e) dst[e] = __bfloat162float(src[e]);
        }
    }
}

// GeGLU: y = gelu_tanh(gate) * up (Gemma MLP, graph/gemma_model.cpp:534-551: ops::gelu then ops::mul) in one pass, with an
// optional bf16 copy of y for the down projection's GEMM; and its backward: d gate = gy * up * gelu'(gate), d up = gy * gelu(gate)
// (GeluBackward + MulBackward, core/backward_functions.cpp:155-177, 66-83), each with an optional bf16 copy.
__global__ void geglu_fwd_kernel(const float* __restrict__ gate, const float* __restrict__ up, float* __restrict__ y, __nv_bfloat16* __restrict__ y16, int64_t n) {
    pdl_sync();
    const int64_t n4 = n >> 2;
    for (int64_t i = blockIdx.x * int64_t(blockDim.x) + threadIdx.x; i < n4; i += int64_t(gridDim.x) * blockDim.x) {
        const float4 g = ld_stream(reinterpret_cast<const float4*>(gate) + i), u = ld_stream(reinterpret_cast<const float4*>(up) + i);
        const float4 r = make_float4(gelu_tanh_f(g.x) * u.x, gelu_tanh_f(g.y) * u.y, gelu_tanh_f(g.z) * u.z, gelu_tanh_f(g.w) * u.w);
        st_stream(reinterpret_cast<float4*>(y) + i, r);
        if (y16) st_bf16x4(y16 + i * 4, r);
    }
}
__global__ void geglu_bwd_kernel(const float* __restrict__ gate, const float* __restrict__ up, const float* __restrict__ gy, float* __restrict__ dgate,
                                 float* __restrict__ dup, __nv_bfloat16* __restrict__ dgate16, __nv_bfloat16* __restrict__ dup16, int64_t n) {
    pdl_sync();
    const int64_t n4 = n >> 2;
    for (int64_t i = blockIdx.x * int64_t(blockDim.x) + threadIdx.x; i < n4; i += int64_t(gridDim.x) * blockDim.x) {
        const float4 g = ld_stream(reinterpret_cast<const float4*>(gate) + i), u = ld_stream(reinterpret_cast<const float4*>(up) + i);
        const float4 d = ld_stream(reinterpret_cast<const float4*>(gy) + i);
        // same association as the two reference nodes: MulBackward gives gy * up, GeluBackward multiplies by gelu'
        const float4 dg = make_float4((d.x * u.x) * gelu_tanh_grad_f(g.x), (d.y * u.y) * gelu_tanh_grad_f(g.y), (d.z * u.z) * gelu_tanh_grad_f(g.z),
                                      (d.w * u.w) * gelu_tanh_grad_f(g.w));
        const float4 du = make_float4(d.x * gelu_tanh_f(g.x), d.y * gelu_tanh_f(g.y), d.z * gelu_tanh_f(g.z), d.w * gelu_tanh_f(g.w));
        st_stream(reinterpret_cast<float4*>(dgate) + i, dg);
        st_stream(reinterpret_cast<float4*>(dup) + i, du);
        if (dgate16) st_bf16x4(dgate16 + i * 4, dg);
        if (dup16) st_bf16x4(dup16 + i * 4, du);
    }
}

// y = pow(x, p0) (kind 0) or clamp(x, p0, p1) (kind 1)
__global__ void unary_param_kernel(int kind, const float* __restrict__ x, float* __restrict__ y, int64_t n, float p0, float p1) {
    pdl_sync();
    for (int64_t i = blockIdx.x * int64_t(blockDim.x) + threadIdx.x; i < n; i += int64_t(gridDim.x) * blockDim.x) {
        float v = x[i];
        y[i] = kind == 0 ? powf(v, p0) : fminf(fmaxf(v, p0), p1);
    }
}

__global__ void mul_bwd_kernel(const float* __restrict__ a, const float* __restrict__ b, const float* __restrict__ gy,
                               float* __restrict__ ga, float* __restrict__ gb, int64_t n) {
    pdl_sync();
    for (int64_t i = blockIdx.x * int64_t(blockDim.x) + threadIdx.x; i < n; i += int64_t(gridDim.x) * blockDim.x) {
        float g = gy[i], x = a[i], y = b[i];
        if (ga) ga[i] = g * y;
        if (gb) gb[i] = g * x;
    }
}

}  // namespace mft

using namespace mft;

extern "C" int mft_binary(int op, const float* a, const float* b, float* out, int64_t outer, int64_t mid,
                          int64_t inner, int b_mode, int64_t b_div, mft_stream_t stream) {
    if (op < 0 || op > 9 || b_mode < 0 || b_mode > 3) return set_error(1, "mft_binary: bad op/b_mode");
    if (outer * mid * inner == 0) return 0;
    if (!a || !b || !out) return set_error(1, "mft_binary: null pointer");
    cudaStream_t st = as_stream(stream);
    switch (op) {
        case 0: launch_binary<0>(a, b, out, outer, mid, inner, b_mode, b_div, st); break;
        case 1: launch_binary<1>(a, b, out, outer, mid, inner, b_mode, b_div, st); break;
        case 2: launch_binary<2>(a, b, out, outer, mid, inner, b_mode, b_div, st); break;
        case 3: launch_binary<3>(a, b, out, outer, mid, inner, b_mode, b_div, st); break;
        case 4: launch_binary<4>(a, b, out, outer, mid, inner, b_mode, b_div, st); break;
        case 5: launch_binary<5>(a, b, out, outer, mid, inner, b_mode, b_div, st); break;
        case 6: launch_binary<6>(a, b, out, outer, mid, inner, b_mode, b_div, st); break;
        case 7: launch_binary<7>(a, b, out, outer, mid, inner, b_mode, b_div, st); break;
        case 8: launch_binary<8>(a, b, out, outer, mid, inner, b_mode, b_div, st); break;
        default: launch_binary<9>(a, b, out, outer, mid, inner, b_mode, b_div, st); break;
    }
    MFT_CHECK_LAUNCH("mft_binary");
    return 0;
}

extern "C" int mft_affine(const float* a, float* out, int64_t n, float alpha, float beta, mft_stream_t stream) {
    if (n == 0) return 0;
    if (!a || !out) return set_error(1, "mft_affine: null pointer");
    pdl_launch(affine_kernel, stream_grid(n, 256, 4), 256, 0, as_stream(stream))(a, out, n, alpha, beta);
    MFT_CHECK_LAUNCH("mft_affine");
    return 0;
}

extern "C" int mft_accumulate(float* acc, const float* g, int64_t n, mft_stream_t stream) {
    if (n == 0) return 0;
    if (!acc || !g) return set_error(1, "mft_accumulate: null pointer");
    pdl_launch(accumulate_kernel, stream_grid(n, 256, 4), 256, 0, as_stream(stream))(acc, g, n);
    MFT_CHECK_LAUNCH("mft_accumulate");
    return 0;
}

extern "C" int mft_fill(float* out, int64_t n, float value, mft_stream_t stream) {
    if (n == 0) return 0;
    if (!out) return set_error(1, "mft_fill: null pointer");
    if (value == 0.0f) {
        MFT_CUDA_OK(cudaMemsetAsync(out, 0, size_t(n) * 4, as_stream(stream)), "mft_fill");
        return 0;
    }
    pdl_launch(fill_kernel, stream_grid(n, 256), 256, 0, as_stream(stream))(out, n, value);
    MFT_CHECK_LAUNCH("mft_fill");
    return 0;
}

extern "C" int mft_unary(int kind, const float* x, float* y, int64_t n, mft_stream_t stream) {
    if (n == 0) return 0;
    if (!x || !y) return set_error(1, "mft_unary: null pointer");
    cudaStream_t st = as_stream(stream);
    int g = stream_grid(n, 256, 4);
    switch (kind) {
        case 0: pdl_launch(unary_kernel<0>, g, 256, 0, st)(x, y, n); break;
        case 1: pdl_launch(unary_kernel<1>, g, 256, 0, st)(x, y, n); break;
        case 2: pdl_launch(unary_kernel<2>, g, 256, 0, st)(x, y, n); break;
        case 3: pdl_launch(unary_kernel<3>, g, 256, 0, st)(x, y, n); break;
        case 4: pdl_launch(unary_kernel<4>, g, 256, 0, st)(x, y, n); break;
        case 5: pdl_launch(unary_kernel<5>, g, 256, 0, st)(x, y, n); break;
        case 6: pdl_launch(unary_kernel<6>, g, 256, 0, st)(x, y, n); break;
        case 7: pdl_launch(unary_kernel<7>, g, 256, 0, st)(x, y, n); break;
        case 8: pdl_launch(unary_kernel<8>, g, 256, 0, st)(x, y, n); break;
        default: return set_error(1, "mft_unary: unknown kind %d", kind);
    }
    MFT_CHECK_LAUNCH("mft_unary");
    return 0;
}

extern "C" int mft_gelu_fwd(const float* x, float* y, int64_t n, mft_stream_t stream) {
    return mft_unary(0, x, y, n, stream);
}

extern "C" int mft_gelu_bwd(const float* x, const float* gy, float* gx, int64_t n, mft_stream_t stream) {
    if (n == 0) return 0;
    if (!x || !gy || !gx) return set_error(1, "mft_gelu_bwd: null pointer");
    pdl_launch(gelu_bwd_kernel, stream_grid(n, 256, 4), 256, 0, as_stream(stream))(x, gy, gx, n, nullptr);
    MFT_CHECK_LAUNCH("mft_gelu_bwd");
    return 0;
}

// same, plus a dense bf16 copy of gx for the bf16 tensor-core GEMM that consumes it (fc_in's dX product)
extern "C" int mft_gelu_bwd_ex(const float* x, const float* gy, float* gx, void* gx_bf16, int64_t n, mft_stream_t stream) {
    if (n == 0) return 0;
    if (!x || !gy || !gx) return set_error(1, "mft_gelu_bwd_ex: null pointer");
    const bool ok = gx_bf16 && (n & 3) == 0 && aligned16(x) && aligned16(gy) && aligned16(gx) && (reinterpret_cast<uintptr_t>(gx_bf16) & 7u) == 0;
    pdl_launch(gelu_bwd_kernel, stream_grid(n, 256, 4), 256, 0, as_stream(stream))(x, gy, gx, n, ok ? static_cast<__nv_bfloat16*>(gx_bf16) : nullptr);
    MFT_CHECK_LAUNCH("mft_gelu_bwd_ex");
    if (gx_bf16 && !ok) return mft_cast_bf16(gx, n, gx_bf16, n, 1, n, stream);
    return 0;
}

extern "C" int mft_geglu_fwd(const float* gate, const float* up, float* y, void* y_bf16, int64_t n, mft_stream_t stream) {
    if (n == 0) return 0;
    if (!gate || !up || !y) return set_error(1, "mft_geglu_fwd: null pointer");
    if ((n & 3) || !aligned16(gate) || !aligned16(up) || !aligned16(y) || (y_bf16 && (reinterpret_cast<uintptr_t>(y_bf16) & 7u)))
        return set_error(1, "mft_geglu_fwd: n must be a multiple of 4 and the tensors 16-byte aligned");
    pdl_launch(geglu_fwd_kernel, stream_grid(n, 256, 4), 256, 0, as_stream(stream))(gate, up, y, static_cast<__nv_bfloat16*>(y_bf16), n);
    MFT_CHECK_LAUNCH("mft_geglu_fwd");
    return 0;
}
extern "C" int mft_geglu_bwd(const float* gate, const float* up, const float* gy, float* dgate, float* dup, void* dgate_bf16, void* dup_bf16, int64_t n,
                             mft_stream_t stream) {
    if (n == 0) return 0;
    if (!gate || !up || !gy || !dgate || !dup) return set_error(1, "mft_geglu_bwd: null pointer");
    if ((n & 3) || !aligned16(gate) || !aligned16(up) || !aligned16(gy) || !aligned16(dgate) || !aligned16(dup) ||
        (dgate_bf16 && (reinterpret_cast<uintptr_t>(dgate_bf16) & 7u)) || (dup_bf16 && (reinterpret_cast<uintptr_t>(dup_bf16) & 7u)))
        return set_error(1, "mft_geglu_bwd: n must be a multiple of 4 and the tensors 16-byte aligned");
    pdl_launch(geglu_bwd_kernel, stream_grid(n, 256, 4), 256, 0, as_stream(stream))(gate, up, gy, dgate, dup, static_cast<__nv_bfloat16*>(dgate_bf16),
                                                                        static_cast<__nv_bfloat16*>(dup_bf16), n);
    MFT_CHECK_LAUNCH("mft_geglu_bwd");
    return 0;
}

// gy and gx in bf16 (x fp32, or bf16 when x_is_bf16); gx (fp32) may be NULL
extern "C" int mft_gelu_bwd_lowp(const void* x, int x_is_bf16, const void* gy_bf16, float* gx, void* gx_bf16, int64_t n, mft_stream_t stream) {
    if (n == 0) return 0;
    if (!x || !gy_bf16 || !gx_bf16) return set_error(1, "mft_gelu_bwd_lowp: null pointer");
    if ((n & 7) || !aligned16(x) || !aligned16(gy_bf16) || !aligned16(gx_bf16) || (gx && !aligned16(gx)))
        return set_error(1, "mft_gelu_bwd_lowp: n must be a multiple of 8 and every pointer 16-byte aligned");
    const int grid = stream_grid(n, 256, 16);
    const void* gyp = gy_bf16;
    auto* g16 = static_cast<const __nv_bfloat16*>(gyp);
    auto* o16 = static_cast<__nv_bfloat16*>(gx_bf16);
    cudaStream_t st = as_stream(stream);
    if (x_is_bf16) { if (gx) pdl_launch(gelu_bwd_lowp_kernel<true, false>, grid, 256, 0, st)(x, g16, gx, o16, n); else pdl_launch(gelu_bwd_lowp_kernel<true, true>, grid, 256, 0, st)(x, g16, gx, o16, n); }
    else { if (gx) pdl_launch(gelu_bwd_lowp_kernel<false, false>, grid, 256, 0, st)(x, g16, gx, o16, n); else pdl_launch(gelu_bwd_lowp_kernel<false, true>, grid, 256, 0, st)(x, g16, gx, o16, n); }
    MFT_CHECK_LAUNCH("mft_gelu_bwd_lowp");
    return 0;
}

extern "C" int mft_cast_from_bf16(const void* in, int64_t ldi, float* out, int64_t ldo, int64_t rows, int64_t cols, mft_stream_t stream) {
    if (rows == 0 || cols == 0) return 0;
    if (!in || !out) return set_error(1, "mft_cast_from_bf16: null pointer");
    pdl_launch(cast_from_bf16_kernel, stream_grid(rows * ((cols + 3) / 4), 256), 256, 0, as_stream(stream))(static_cast<const __nv_bfloat16*>(in), ldi, out, ldo, rows, cols);
    MFT_CHECK_LAUNCH("mft_cast_from_bf16");
    return 0;
}

extern "C" int mft_mul_bwd(const float* a, const float* b, const float* gy, float* ga, float* gb, int64_t n,
                           mft_stream_t stream) {
    if (n == 0) return 0;
    if (!a || !b || !gy) return set_error(1, "mft_mul_bwd: null pointer");
    pdl_launch(mul_bwd_kernel, stream_grid(n, 256), 256, 0, as_stream(stream))(a, b, gy, ga, gb, n);
    MFT_CHECK_LAUNCH("mft_mul_bwd");
    return 0;
}

extern "C" int mft_unary_param(int kind, const float* x, float* y, int64_t n, float p0, float p1, mft_stream_t stream) {
    if (n == 0) return 0;
    if (!x || !y) return set_error(1, "mft_unary_param: null pointer");
    if (kind < 0 || kind > 1) return set_error(1, "mft_unary_param: unknown kind %d", kind);
    pdl_launch(unary_param_kernel, stream_grid(n, 256), 256, 0, as_stream(stream))(kind, x, y, n, p0, p1);
    MFT_CHECK_LAUNCH("mft_unary_param");
    return 0;
}
