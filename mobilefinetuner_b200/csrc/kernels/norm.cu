// LayerNorm / RMSNorm(1+w) / softmax: one warp per row, row cached in registers (one HBM read),
// warp-shuffle reductions, 128-bit accesses. Weight-gradient column sums go through per-CTA partials
// in caller-provided scratch and a fixed-order second stage (deterministic, no atomics).
#include <algorithm>
#include "common.cuh"

namespace mft {

constexpr int kNormWarps = 8;               // warps (rows in flight) per CTA
constexpr int kNormMaxCtas = kNumSMs * 2;   // persistent-style grid for the backward kernels

// ---------------------------------------------------------------- LayerNorm forward
// NV = float4 per lane (dim <= 128*NV). Algorithmic bytes: 8*rows*dim.
template <int NV>
__global__ void __launch_bounds__(kNormWarps * 32)
layernorm_fwd_kernel(const float* __restrict__ x, const float* __restrict__ gamma, const float* __restrict__ beta,
                     float* __restrict__ y, int64_t rows, int dim, float eps,
                     const float* __restrict__ add, float* __restrict__ sum_out,     // add != NULL: normalise x + add, store the sum
                     __nv_bfloat16* __restrict__ y16, int64_t ld16) {
    pdl_sync();                 // y16 != NULL: bf16 copy of y (pitch ld16)
    const int lane = threadIdx.x & 31;
    const int dim4 = dim >> 2;
    const int64_t warps = (int64_t(gridDim.x) * blockDim.x) >> 5;
    for (int64_t r = (blockIdx.x * int64_t(blockDim.x) + threadIdx.x) >> 5; r < rows; r += warps) {
        const float4* xr = reinterpret_cast<const float4*>(x + r * dim);
        float4 v[NV];
        float s = 0.f;
#pragma unroll
        for (int j = 0; j < NV; ++j) {
            int c = lane + 32 * j;
            v[j] = (c < dim4) ? ld_stream(xr + c) : make_float4(0.f, 0.f, 0.f, 0.f);
            if (add && c < dim4) {
                const float4 a = ld_stream(reinterpret_cast<const float4*>(add + r * dim) + c);
                v[j].x += a.x; v[j].y += a.y; v[j].z += a.z; v[j].w += a.w;
                st_stream(reinterpret_cast<float4*>(sum_out + r * dim) + c, v[j]);
            }
            s += v[j].x + v[j].y + v[j].z + v[j].w;
        }
        const float mean = warp_sum(s) / float(dim);
        float q = 0.f;
#pragma unroll
        for (int j = 0; j < NV; ++j) {
            int c = lane + 32 * j;
            if (c < dim4) {
                float a = v[j].x - mean, b = v[j].y - mean, cc = v[j].z - mean, d = v[j].w - mean;
                q += a * a + b * b + cc * cc + d * d;
            }
        }
        const float rstd = 1.0f / sqrtf(warp_sum(q) / float(dim) + eps);
        float4* yr = reinterpret_cast<float4*>(y + r * dim);
#pragma unroll
        for (int j = 0; j < NV; ++j) {
            int c = lane + 32 * j;
            if (c < dim4) {
                float4 g = __ldg(reinterpret_cast<const float4*>(gamma) + c);
                float4 b = __ldg(reinterpret_cast<const float4*>(beta) + c);
                float4 o;
                o.x = (v[j].x - mean) * rstd * g.x + b.x;
                o.y = (v[j].y - mean) * rstd * g.y + b.y;
                o.z = (v[j].z - mean) * rstd * g.z + b.z;
                o.w = (v[j].w - mean) * rstd * g.w + b.w;
                if (y) st_stream(yr + c, o);                 // y == NULL: the bf16 copy is the only result
                if (y16) st_bf16x4(y16 + r * ld16 + c * 4, o);
            }
        }
    }
}

// Generic fallback (any dim / alignment): one warp per row, three passes through L1/L2.
__global__ void layernorm_fwd_generic(const float* __restrict__ x, const float* __restrict__ gamma,
                                      const float* __restrict__ beta, float* __restrict__ y, int64_t rows, int dim, float eps) {
    pdl_sync();
    const int lane = threadIdx.x & 31;
    const int64_t warps = (int64_t(gridDim.x) * blockDim.x) >> 5;
    for (int64_t r = (blockIdx.x * int64_t(blockDim.x) + threadIdx.x) >> 5; r < rows; r += warps) {
        const float* xr = x + r * dim;
        float s = 0.f;
        for (int i = lane; i < dim; i += 32) s += xr[i];
        const float mean = warp_sum(s) / float(dim);
        float q = 0.f;
        for (int i = lane; i < dim; i += 32) { float d = xr[i] - mean; q += d * d; }
        const float rstd = 1.0f / sqrtf(warp_sum(q) / float(dim) + eps);
        for (int i = lane; i < dim; i += 32) y[r * dim + i] = (xr[i] - mean) * rstd * gamma[i] + beta[i];
    }
}

// ---------------------------------------------------------------- LayerNorm backward
// dx per row; per-CTA partial dgamma/dbeta -> partial[cta][2][dim]. Algorithmic bytes: 12*rows*dim.
template <int NV, bool WG>
__global__ void __launch_bounds__(kNormWarps * 32)
layernorm_bwd_kernel(const float* __restrict__ x, const float* __restrict__ gamma, const float* __restrict__ gy,
                     float* __restrict__ gx, float* __restrict__ partial, int64_t rows, int dim, float eps,
                     const float* __restrict__ addend,     // addend != NULL: gx = dx + addend (gradient already waiting at x)
                     __nv_bfloat16* __restrict__ gx16, int64_t ld16) {
    pdl_sync();   // gx16 != NULL: bf16 copy of gx
    extern __shared__ float smem[];  // [kNormWarps][2][dim]
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    const int dim4 = dim >> 2;
    const int nw = WG ? kNormWarps : int(blockDim.x >> 5);     // the dx-only form runs with smaller CTAs (better SM balance)
    const int64_t warps = int64_t(gridDim.x) * nw;
    float4 dg[NV], db[NV], gam[NV];
#pragma unroll
    for (int j = 0; j < NV; ++j) {
        dg[j] = make_float4(0.f, 0.f, 0.f, 0.f);
        db[j] = dg[j];
        int c = lane + 32 * j;
        gam[j] = (c < dim4) ? __ldg(reinterpret_cast<const float4*>(gamma) + c) : dg[j];
    }
    for (int64_t r = int64_t(blockIdx.x) * nw + warp; r < rows; r += warps) {
        const float4* xr = reinterpret_cast<const float4*>(x + r * dim);
        const float4* gr = reinterpret_cast<const float4*>(gy + r * dim);
        float4 v[NV], g[NV], ad[WG ? 1 : NV];
        float s = 0.f;
#pragma unroll
        for (int j = 0; j < NV; ++j) {
            int c = lane + 32 * j;
            v[j] = (c < dim4) ? ld_stream(xr + c) : make_float4(0.f, 0.f, 0.f, 0.f);
            g[j] = (c < dim4) ? ld_stream(gr + c) : make_float4(0.f, 0.f, 0.f, 0.f);
            // dx-only form: the waiting gradient is requested with the other two streams, not after the reductions
            if (!WG) ad[j] = (addend && c < dim4) ? ld_stream(reinterpret_cast<const float4*>(addend + r * dim) + c) : make_float4(0.f, 0.f, 0.f, 0.f);
            s += v[j].x + v[j].y + v[j].z + v[j].w;
        }
        const float mean = warp_sum(s) / float(dim);
        float q = 0.f;
#pragma unroll
        for (int j = 0; j < NV; ++j) {
            int c = lane + 32 * j;
            if (c < dim4) {
                float a = v[j].x - mean, b = v[j].y - mean, cc = v[j].z - mean, d = v[j].w - mean;
                q += a * a + b * b + cc * cc + d * d;
            }
        }
        const float rstd = 1.0f / sqrtf(warp_sum(q) / float(dim) + eps);
        float s1 = 0.f, s2 = 0.f;
#pragma unroll
        for (int j = 0; j < NV; ++j) {   // v becomes x_hat, g stays gy
            v[j].x = (v[j].x - mean) * rstd; v[j].y = (v[j].y - mean) * rstd;
            v[j].z = (v[j].z - mean) * rstd; v[j].w = (v[j].w - mean) * rstd;
            int c = lane + 32 * j;
            if (c < dim4) {
                float a = g[j].x * gam[j].x, b = g[j].y * gam[j].y, cc = g[j].z * gam[j].z, d = g[j].w * gam[j].w;
                s1 += a + b + cc + d;
                s2 += a * v[j].x + b * v[j].y + cc * v[j].z + d * v[j].w;
                if (WG) {
                    dg[j].x += g[j].x * v[j].x; dg[j].y += g[j].y * v[j].y; dg[j].z += g[j].z * v[j].z; dg[j].w += g[j].w * v[j].w;
                    db[j].x += g[j].x; db[j].y += g[j].y; db[j].z += g[j].z; db[j].w += g[j].w;
                }
            }
        }
        s1 = warp_sum(s1);
        s2 = warp_sum(s2);
        const float k = rstd / float(dim), D = float(dim);
        float4* or_ = reinterpret_cast<float4*>(gx + r * dim);
#pragma unroll
        for (int j = 0; j < NV; ++j) {
            int c = lane + 32 * j;
            if (c < dim4) {
                float4 o;
                o.x = k * (D * (g[j].x * gam[j].x) - s1 - v[j].x * s2);
                o.y = k * (D * (g[j].y * gam[j].y) - s1 - v[j].y * s2);
                o.z = k * (D * (g[j].z * gam[j].z) - s1 - v[j].z * s2);
                o.w = k * (D * (g[j].w * gam[j].w) - s1 - v[j].w * s2);
                if (!WG) {       // explicit adds (no contraction into an FMA with the product above): bit-identical to a separate add kernel
                    o.x = __fadd_rn(o.x, ad[j].x); o.y = __fadd_rn(o.y, ad[j].y); o.z = __fadd_rn(o.z, ad[j].z); o.w = __fadd_rn(o.w, ad[j].w);
                }
                else if (addend) {
                    const float4 a = ld_stream(reinterpret_cast<const float4*>(addend + r * dim) + c);
                    o.x += a.x; o.y += a.y; o.z += a.z; o.w += a.w;
                }
                st_stream(or_ + c, o);
                if (gx16) st_bf16x4(gx16 + r * ld16 + c * 4, o);
            }
        }
    }
    if (!WG) return;   // frozen gamma/beta: only dx was wanted
    // CTA reduction of the per-warp column partials, then one row of `partial` per CTA.
    float4* sg = reinterpret_cast<float4*>(smem) + size_t(warp) * 2 * dim4;
#pragma unroll
    for (int j = 0; j < NV; ++j) {
        int c = lane + 32 * j;
        if (c < dim4) { sg[c] = dg[j]; sg[dim4 + c] = db[j]; }
    }
    __syncthreads();
    float* out = partial + size_t(blockIdx.x) * 2 * dim;
    for (int i = threadIdx.x; i < 2 * dim; i += blockDim.x) {
        float s = 0.f;
#pragma unroll
        for (int w = 0; w < kNormWarps; ++w) s += smem[size_t(w) * 2 * dim + i];
        out[i] = s;
    }
}

__global__ void layernorm_bwd_generic(const float* __restrict__ x, const float* __restrict__ gamma,
                                      const float* __restrict__ gy, float* __restrict__ gx, int64_t rows, int dim, float eps) {
    pdl_sync();
    const int lane = threadIdx.x & 31;
    const int64_t warps = (int64_t(gridDim.x) * blockDim.x) >> 5;
    for (int64_t r = (blockIdx.x * int64_t(blockDim.x) + threadIdx.x) >> 5; r < rows; r += warps) {
        const float* xr = x + r * dim;
        const float* gr = gy + r * dim;
        float s = 0.f;
        for (int i = lane; i < dim; i += 32) s += xr[i];
        const float mean = warp_sum(s) / float(dim);
        float q = 0.f;
        for (int i = lane; i < dim; i += 32) { float d = xr[i] - mean; q += d * d; }
        const float rstd = 1.0f / sqrtf(warp_sum(q) / float(dim) + eps);
        float s1 = 0.f, s2 = 0.f;
        for (int i = lane; i < dim; i += 32) {
            float gh = gr[i] * gamma[i], xh = (xr[i] - mean) * rstd;
            s1 += gh; s2 += gh * xh;
        }
        s1 = warp_sum(s1); s2 = warp_sum(s2);
        for (int i = lane; i < dim; i += 32) {
            float gh = gr[i] * gamma[i], xh = (xr[i] - mean) * rstd;
            gx[r * dim + i] = (rstd / float(dim)) * (float(dim) * gh - s1 - xh * s2);
        }
    }
}
// generic weight-grad partials: CTA b handles rows [b*chunk, (b+1)*chunk), thread per column.
__global__ void layernorm_wgrad_generic(const float* __restrict__ x, const float* __restrict__ gy,
                                        float* __restrict__ partial, int64_t rows, int dim, float eps, int64_t chunk) {
    pdl_sync();
    extern __shared__ float stats[];  // [chunk][2] mean, rstd
    const int64_t r0 = int64_t(blockIdx.x) * chunk;
    int64_t r1 = r0 + chunk; if (r1 > rows) r1 = rows;
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5, nw = blockDim.x >> 5;
    for (int64_t r = r0 + warp; r < r1; r += nw) {
        const float* xr = x + r * dim;
        float s = 0.f;
        for (int i = lane; i < dim; i += 32) s += xr[i];
        float mean = warp_sum(s) / float(dim);
        float q = 0.f;
        for (int i = lane; i < dim; i += 32) { float d = xr[i] - mean; q += d * d; }
        float rstd = 1.0f / sqrtf(warp_sum(q) / float(dim) + eps);
        if (lane == 0) { stats[(r - r0) * 2] = mean; stats[(r - r0) * 2 + 1] = rstd; }
    }
    __syncthreads();
    for (int c = threadIdx.x; c < dim; c += blockDim.x) {
        float a = 0.f, b = 0.f;
        for (int64_t r = r0; r < r1; ++r) {
            float g = gy[r * dim + c];
            a += g * (x[r * dim + c] - stats[(r - r0) * 2]) * stats[(r - r0) * 2 + 1];
            b += g;
        }
        partial[size_t(blockIdx.x) * 2 * dim + c] = a;
        partial[size_t(blockIdx.x) * 2 * dim + dim + c] = b;
    }
}

// Second stage shared by LayerNorm (two outputs) and RMSNorm (one): 32 columns x 8 row-groups per CTA, each thread sums
// every 8th partial row, then the 8 groups are added in a fixed order (deterministic).
__global__ void __launch_bounds__(256)
norm_wgrad_final(const float* __restrict__ partial, float* __restrict__ out0, float* __restrict__ out1,
                 int parts, int dim, int per_part) {
    pdl_sync();
    __shared__ float ra[8][33], rb[8][33];
    const int tx = threadIdx.x & 31, ty = threadIdx.x >> 5;
    const int c = blockIdx.x * 32 + tx;
    float a = 0.f, b = 0.f;
    if (c < dim)
        for (int p = ty; p < parts; p += 8) {
            a += partial[size_t(p) * per_part * dim + c];
            if (out1) b += partial[size_t(p) * per_part * dim + dim + c];
        }
    ra[ty][tx] = a; rb[ty][tx] = b;
    __syncthreads();
    if (ty == 0 && c < dim) {
        float sa = 0.f, sb = 0.f;
#pragma unroll
        for (int j = 0; j < 8; ++j) { sa += ra[j][tx]; sb += rb[j][tx]; }
        out0[c] = sa;
        if (out1) out1[c] = sb;
    }
}

}  // namespace mft

namespace mft {

// ---------------------------------------------------------------- RMSNorm (1+w), double statistics
// One warp per row, generic over dim (row re-read hits L1); Gemma dims are 640 (hidden) and 256 (q/k norm).
__global__ void rmsnorm_fwd_kernel(const float* __restrict__ x, const float* __restrict__ w, float* __restrict__ y,
                                   int64_t rows, int dim, float eps, int plus_one,
                                   __nv_bfloat16* __restrict__ y16, int64_t ld16) {
    pdl_sync();   // y16 != NULL: bf16 copy of y
    const int lane = threadIdx.x & 31;
    const int64_t warps = (int64_t(gridDim.x) * blockDim.x) >> 5;
    for (int64_t r = (blockIdx.x * int64_t(blockDim.x) + threadIdx.x) >> 5; r < rows; r += warps) {
        const float* xr = x + r * dim;
        double sq = 0.0;
        for (int i = lane; i < dim; i += 32) { double v = xr[i]; sq += v * v; }
        sq = warp_sum(sq);
        const float rms = float(sqrt(sq / double(dim) + double(eps)));
        for (int i = lane; i < dim; i += 32) {
            float sc = plus_one ? 1.0f + w[i] : w[i];
            const float o = (xr[i] / rms) * sc;
            y[r * dim + i] = o;
            if (y16) y16[r * ld16 + i] = __float2bfloat16_rn(o);
        }
    }
}

// dx per row (double math as the reference) + per-CTA partial dw -> partial[cta][dim].
__global__ void __launch_bounds__(kNormWarps * 32)
rmsnorm_bwd_kernel(const float* __restrict__ x, const float* __restrict__ w, const float* __restrict__ gy,
                   float* __restrict__ gx, float* __restrict__ partial, int64_t rows, int dim, float eps, int plus_one,
                   __nv_bfloat16* __restrict__ gx16, int64_t ld16) {
    pdl_sync();          // gx16 != NULL: bf16 copy of gx
    extern __shared__ float smem[];  // [kNormWarps][dim]
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    float* my = smem + size_t(warp) * dim;
    for (int i = lane; i < dim; i += 32) my[i] = 0.f;
    const int64_t warps = int64_t(gridDim.x) * kNormWarps;
    for (int64_t r = int64_t(blockIdx.x) * kNormWarps + warp; r < rows; r += warps) {
        const float* xr = x + r * dim;
        const float* gr = gy + r * dim;
        double sq = 0.0, dot = 0.0;
        for (int i = lane; i < dim; i += 32) {
            double v = xr[i];
            double we = plus_one ? 1.0 + double(w[i]) : double(w[i]);
            sq += v * v;
            dot += double(gr[i]) * we * v;
        }
        sq = warp_sum(sq);
        dot = warp_sum(dot);
        const double inv = 1.0 / sqrt(sq / double(dim) + double(eps));
        const double coeff = inv * inv * inv / double(dim);
        for (int i = lane; i < dim; i += 32) {
            double we = plus_one ? 1.0 + double(w[i]) : double(w[i]);
            double g = gr[i], v = xr[i];
            const float o = float(g * we * inv - v * coeff * dot);
            gx[r * dim + i] = o;
            if (gx16) gx16[r * ld16 + i] = __float2bfloat16_rn(o);
            if (partial) my[i] += float(g * v * inv);
        }
    }
    if (!partial) return;          // frozen weight: only dx was wanted (uniform over the grid)
    __syncthreads();
    for (int i = threadIdx.x; i < dim; i += blockDim.x) {
        float s = 0.f;
#pragma unroll
        for (int q = 0; q < kNormWarps; ++q) s += smem[size_t(q) * dim + i];
        partial[size_t(blockIdx.x) * dim + i] = s;
    }
}

// ---------------------------------------------------------------- softmax over the last dim
__global__ void softmax_fwd_kernel(const float* __restrict__ x, float* __restrict__ y, int64_t rows, int dim) {
    pdl_sync();
    const int lane = threadIdx.x & 31;
    const int64_t warps = (int64_t(gridDim.x) * blockDim.x) >> 5;
    for (int64_t r = (blockIdx.x * int64_t(blockDim.x) + threadIdx.x) >> 5; r < rows; r += warps) {
        const float* xr = x + r * dim;
        float m = -INFINITY;
        for (int i = lane; i < dim; i += 32) m = fmaxf(m, xr[i]);
        m = warp_max(m);
        float s = 0.f;
        for (int i = lane; i < dim; i += 32) s += expf(xr[i] - m);
        s = warp_sum(s);
        for (int i = lane; i < dim; i += 32) y[r * dim + i] = expf(xr[i] - m) / s;
    }
}
__global__ void softmax_bwd_kernel(const float* __restrict__ y, const float* __restrict__ gy, float* __restrict__ gx,
                                   int64_t rows, int dim) {
    pdl_sync();
    const int lane = threadIdx.x & 31;
    const int64_t warps = (int64_t(gridDim.x) * blockDim.x) >> 5;
    for (int64_t r = (blockIdx.x * int64_t(blockDim.x) + threadIdx.x) >> 5; r < rows; r += warps) {
        const float* yr = y + r * dim;
        const float* gr = gy + r * dim;
        double dot = 0.0;
        for (int i = lane; i < dim; i += 32) dot += double(gr[i]) * double(yr[i]);
        const float d = float(warp_sum(dot));
        for (int i = lane; i < dim; i += 32) gx[r * dim + i] = yr[i] * (gr[i] - d);
    }
}

static int norm_bwd_ctas(int64_t rows) {
    int64_t c = ceil_div(rows, kNormWarps);
    if (c > kNormMaxCtas) c = kNormMaxCtas;
    if (c < 1) c = 1;
    return int(c);
}

template <int NV>
static int launch_ln_fwd(const float* x, const float* g, const float* b, float* y, int64_t rows, int dim, float eps, cudaStream_t st,
                         const float* add = nullptr, float* sum_out = nullptr, void* y16 = nullptr, int64_t ld16 = 0) {
    // four rows (warps) per CTA: 4096 rows spread as 6-7 CTAs per SM instead of 3-4 eight-row CTAs (balance 99 % instead of 86 %)
    constexpr int kFwdWarps = 4;
    int blocks = int(ceil_div(rows, kFwdWarps));
    if (blocks > kNumSMs * 16) blocks = kNumSMs * 16;
    pdl_launch(layernorm_fwd_kernel<NV>, blocks, kFwdWarps * 32, 0, st)(x, g, b, y, rows, dim, eps, add, sum_out, static_cast<__nv_bfloat16*>(y16), ld16);
    return 0;
}
template <int NV>
static int launch_ln_bwd(const float* x, const float* g, const float* gy, float* gx, float* partial, int64_t rows,
                         int dim, float eps, int ctas, cudaStream_t st, const float* addend = nullptr, void* gx16v = nullptr, int64_t ld16 = 0) {
    __nv_bfloat16* gx16 = static_cast<__nv_bfloat16*>(gx16v);
    if (!partial) {   // dx only: four rows per CTA (see launch_ln_fwd)
        constexpr int kDxWarps = 4;
        int blocks = int(ceil_div(rows, kDxWarps));
        if (blocks > kNumSMs * 16) blocks = kNumSMs * 16;
        pdl_launch(layernorm_bwd_kernel<NV, false>, blocks, kDxWarps * 32, 0, st)(x, g, gy, gx, nullptr, rows, dim, eps, addend, gx16, ld16);
        return 0;
    }
    size_t smem = size_t(kNormWarps) * 2 * dim * sizeof(float);
    static bool attr_set[kMaxDevices] = {};
    if (first_use_on_device(attr_set))
        cudaFuncSetAttribute(layernorm_bwd_kernel<NV, true>, cudaFuncAttributeMaxDynamicSharedMemorySize, 200 * 1024);
    pdl_launch(layernorm_bwd_kernel<NV, true>, ctas, kNormWarps * 32, smem, st)(x, g, gy, gx, partial, rows, dim, eps, addend, gx16, ld16);
    return 0;
}

}  // namespace mft

using namespace mft;

// add != NULL: y = LayerNorm(x + add) and sum_out = x + add, one pass (the residual add in front of ln_2 / the next block's ln_1)
static int ln_fwd_impl(const char* who, const float* x, const float* add, float* sum_out, const float* gamma, const float* beta, float* y,
                       int64_t rows, int64_t dim, float eps, mft_stream_t stream, void* y16 = nullptr, int64_t ld16 = 0) {
    if (rows == 0 || dim == 0) return 0;
    if (!x || !gamma || !beta || (!y && !y16) || (add && !sum_out)) return set_error(1, "%s: null pointer", who);
    cudaStream_t st = as_stream(stream);
    const bool vec = (dim % 4 == 0) && dim <= 2048 && aligned16(x) && (!y || aligned16(y)) && aligned16(gamma) && aligned16(beta) &&
                     (!add || (aligned16(add) && aligned16(sum_out))) &&
                     (!y16 || ((reinterpret_cast<uintptr_t>(y16) & 7u) == 0 && (ld16 & 3) == 0));
    if (vec) {
        int nv = int(ceil_div(dim, 128));
        if (nv <= 1) launch_ln_fwd<1>(x, gamma, beta, y, rows, int(dim), eps, st, add, sum_out, y16, ld16);
        else if (nv <= 2) launch_ln_fwd<2>(x, gamma, beta, y, rows, int(dim), eps, st, add, sum_out, y16, ld16);
        else if (nv <= 4) launch_ln_fwd<4>(x, gamma, beta, y, rows, int(dim), eps, st, add, sum_out, y16, ld16);
        else if (nv <= 6) launch_ln_fwd<6>(x, gamma, beta, y, rows, int(dim), eps, st, add, sum_out, y16, ld16);
        else if (nv <= 8) launch_ln_fwd<8>(x, gamma, beta, y, rows, int(dim), eps, st, add, sum_out, y16, ld16);
        else launch_ln_fwd<16>(x, gamma, beta, y, rows, int(dim), eps, st, add, sum_out, y16, ld16);
    } else {
        if (!y) return set_error(1, "%s: a bf16-only result needs dim %% 4 == 0, dim <= 2048 and 16-byte aligned tensors", who);
        const float* src = x;
        if (add) {   // unaligned / wide rows: separate add, then the generic kernel on the sum
            int rc = mft_binary(0, x, add, sum_out, 1, 1, rows * dim, 0, 1, stream);
            if (rc) return rc;
            src = sum_out;
        }
        pdl_launch(layernorm_fwd_generic, stream_grid(rows * 32, 256), 256, 0, st)(src, gamma, beta, y, rows, int(dim), eps);
        if (y16) { int rc = mft_cast_bf16(y, dim, y16, ld16, rows, dim, stream); if (rc) return rc; }
    }
    MFT_CHECK_LAUNCH(who);
    return 0;
}

extern "C" int mft_layernorm_fwd(const float* x, const float* gamma, const float* beta, float* y, int64_t rows,
                                 int64_t dim, float eps, mft_stream_t stream) {
    return ln_fwd_impl("mft_layernorm_fwd", x, nullptr, nullptr, gamma, beta, y, rows, dim, eps, stream);
}

extern "C" int mft_add_layernorm_fwd(const float* x, const float* add, float* sum_out, const float* gamma, const float* beta, float* y,
                                     int64_t rows, int64_t dim, float eps, mft_stream_t stream) {
    if (!add || !sum_out) return set_error(1, "mft_add_layernorm_fwd: null pointer");
    return ln_fwd_impl("mft_add_layernorm_fwd", x, add, sum_out, gamma, beta, y, rows, dim, eps, stream);
}

extern "C" size_t mft_norm_bwd_workspace_bytes(int64_t rows, int64_t dim) {
    // room for either path: vectorised (ctas x 2 x dim) or generic (chunks of 64 rows x 2 x dim)
    int64_t parts = norm_bwd_ctas(rows);
    int64_t chunks = ceil_div(rows, 64);
    if (chunks > parts) parts = chunks;
    return size_t(parts) * 2 * size_t(dim) * sizeof(float);
}

// addend != NULL: gx = dX + addend — the gradient that already reached x through the residual branch is added while dX is
// still in registers instead of by a separate accumulation pass.
static int ln_bwd_impl(const float* x, const float* gamma, const float* gy, const float* addend, float* gx, float* dgamma,
                       float* dbeta, int64_t rows, int64_t dim, float eps, void* workspace, size_t workspace_bytes, mft_stream_t stream,
                       void* gx16 = nullptr, int64_t ld16 = 0) {
    if (dim == 0) return 0;
    if (!x || !gamma || !gy || !gx) return set_error(1, "mft_layernorm_bwd: null pointer");
    if ((dgamma == nullptr) != (dbeta == nullptr)) return set_error(1, "mft_layernorm_bwd: dgamma and dbeta must both be given or both be NULL");
    const bool want_wgrad = dgamma != nullptr;
    if (rows == 0) { if (want_wgrad) { mft_fill(dgamma, dim, 0.f, stream); return mft_fill(dbeta, dim, 0.f, stream); } return 0; }
    if (want_wgrad && (!workspace || workspace_bytes < mft_norm_bwd_workspace_bytes(rows, dim)))
        return set_error(1, "mft_layernorm_bwd: workspace too small (%zu < %zu)", workspace_bytes,
                         mft_norm_bwd_workspace_bytes(rows, dim));
    cudaStream_t st = as_stream(stream);
    float* partial = want_wgrad ? static_cast<float*>(workspace) : nullptr;
    const bool vec = (dim % 4 == 0) && dim <= 2048 && aligned16(x) && aligned16(gy) && aligned16(gx) && aligned16(gamma) &&
                     (!addend || aligned16(addend)) && (!gx16 || ((reinterpret_cast<uintptr_t>(gx16) & 7u) == 0 && (ld16 & 3) == 0));
    int parts;
    if (vec) {
        parts = norm_bwd_ctas(rows);
        int nv = int(ceil_div(dim, 128));
        if (nv <= 1) launch_ln_bwd<1>(x, gamma, gy, gx, partial, rows, int(dim), eps, parts, st, addend, gx16, ld16);
        else if (nv <= 2) launch_ln_bwd<2>(x, gamma, gy, gx, partial, rows, int(dim), eps, parts, st, addend, gx16, ld16);
        else if (nv <= 4) launch_ln_bwd<4>(x, gamma, gy, gx, partial, rows, int(dim), eps, parts, st, addend, gx16, ld16);
        else if (nv <= 6) launch_ln_bwd<6>(x, gamma, gy, gx, partial, rows, int(dim), eps, parts, st, addend, gx16, ld16);
        else if (nv <= 8) launch_ln_bwd<8>(x, gamma, gy, gx, partial, rows, int(dim), eps, parts, st, addend, gx16, ld16);
        else launch_ln_bwd<16>(x, gamma, gy, gx, partial, rows, int(dim), eps, parts, st, addend, gx16, ld16);
    } else {
        pdl_launch(layernorm_bwd_generic, stream_grid(rows * 32, 256), 256, 0, st)(x, gamma, gy, gx, rows, int(dim), eps);
        if (addend) { int rc = mft_accumulate(gx, addend, rows * dim, stream); if (rc) return rc; }
        if (gx16) { int rc = mft_cast_bf16(gx, dim, gx16, ld16, rows, dim, stream); if (rc) return rc; }
        parts = int(ceil_div(rows, 64));
        if (want_wgrad) pdl_launch(layernorm_wgrad_generic, parts, 256, 64 * 2 * sizeof(float), st)(x, gy, partial, rows, int(dim), eps, 64);
    }
    if (want_wgrad) {
        pdl_launch(norm_wgrad_final, unsigned(ceil_div(dim, 32)), 256, 0, st)(partial, dgamma, dbeta, parts, int(dim), 2);
        add_launches(1);
    }
    MFT_CHECK_LAUNCH("mft_layernorm_bwd");
    return 0;
}

extern "C" int mft_layernorm_bwd(const float* x, const float* gamma, const float* gy, float* gx, float* dgamma,
                                 float* dbeta, int64_t rows, int64_t dim, float eps, void* workspace,
                                 size_t workspace_bytes, mft_stream_t stream) {
    return ln_bwd_impl(x, gamma, gy, nullptr, gx, dgamma, dbeta, rows, dim, eps, workspace, workspace_bytes, stream);
}

// Supersets used by the bf16 GEMM mode: optional residual operands as above plus an optional bf16 copy of the result
// (y_bf16 / gx_bf16, row pitch ld_bf16 elements, multiple of 4, 8-byte aligned) for the GEMM that consumes it next.
extern "C" int mft_layernorm_fwd_ex(const float* x, const float* add, float* sum_out, const float* gamma, const float* beta, float* y,
                                    void* y_bf16, int64_t ld_bf16, int64_t rows, int64_t dim, float eps, mft_stream_t stream) {
    return ln_fwd_impl("mft_layernorm_fwd_ex", x, add, sum_out, gamma, beta, y, rows, dim, eps, stream, y_bf16, ld_bf16);
}
extern "C" int mft_layernorm_bwd_ex(const float* x, const float* gamma, const float* gy, const float* addend, float* gx, void* gx_bf16,
                                    int64_t ld_bf16, float* dgamma, float* dbeta, int64_t rows, int64_t dim, float eps, void* workspace,
                                    size_t workspace_bytes, mft_stream_t stream) {
    return ln_bwd_impl(x, gamma, gy, addend, gx, dgamma, dbeta, rows, dim, eps, workspace, workspace_bytes, stream, gx_bf16, ld_bf16);
}

extern "C" int mft_layernorm_bwd_add(const float* x, const float* gamma, const float* gy, const float* addend, float* gx, float* dgamma,
                                     float* dbeta, int64_t rows, int64_t dim, float eps, void* workspace,
                                     size_t workspace_bytes, mft_stream_t stream) {
    if (!addend) return set_error(1, "mft_layernorm_bwd_add: null addend");
    return ln_bwd_impl(x, gamma, gy, addend, gx, dgamma, dbeta, rows, dim, eps, workspace, workspace_bytes, stream);
}

extern "C" int mft_rmsnorm_fwd(const float* x, const float* w, float* y, int64_t rows, int64_t dim, float eps,
                               int plus_one, mft_stream_t stream) {
    if (rows == 0 || dim == 0) return 0;
    if (!x || !w || !y) return set_error(1, "mft_rmsnorm_fwd: null pointer");
    pdl_launch(rmsnorm_fwd_kernel, stream_grid(rows * 32, 256), 256, 0, as_stream(stream))(x, w, y, rows, int(dim), eps, plus_one, nullptr, 0);
    MFT_CHECK_LAUNCH("mft_rmsnorm_fwd");
    return 0;
}

// same, plus a bf16 copy of y (row pitch ld_bf16 >= dim) for the bf16 GEMMs that consume the normalised activations
extern "C" int mft_rmsnorm_fwd_ex(const float* x, const float* w, float* y, void* y_bf16, int64_t ld_bf16, int64_t rows, int64_t dim,
                                  float eps, int plus_one, mft_stream_t stream) {
    if (rows == 0 || dim == 0) return 0;
    if (!x || !w || !y) return set_error(1, "mft_rmsnorm_fwd_ex: null pointer");
    if (y_bf16 && ld_bf16 < dim) return set_error(1, "mft_rmsnorm_fwd_ex: bf16 pitch smaller than dim");
    pdl_launch(rmsnorm_fwd_kernel, stream_grid(rows * 32, 256), 256, 0, as_stream(stream))(x, w, y, rows, int(dim), eps, plus_one,
                                                                                   static_cast<__nv_bfloat16*>(y_bf16), ld_bf16);
    MFT_CHECK_LAUNCH("mft_rmsnorm_fwd_ex");
    return 0;
}

static int rmsnorm_bwd_impl(const float* x, const float* w, const float* gy, float* gx, float* dw, int64_t rows, int64_t dim, float eps,
                            int plus_one, void* workspace, size_t workspace_bytes, mft_stream_t stream, void* gx16, int64_t ld16);

extern "C" int mft_rmsnorm_bwd(const float* x, const float* w, const float* gy, float* gx, float* dw, int64_t rows,
                               int64_t dim, float eps, int plus_one, void* workspace, size_t workspace_bytes,
                               mft_stream_t stream) {
    return rmsnorm_bwd_impl(x, w, gy, gx, dw, rows, dim, eps, plus_one, workspace, workspace_bytes, stream, nullptr, 0);
}
// same, plus a bf16 copy of gx (row pitch ld_bf16 elements) for the dX products that consume the gradient
extern "C" int mft_rmsnorm_bwd_ex(const float* x, const float* w, const float* gy, float* gx, void* gx_bf16, int64_t ld_bf16, float* dw,
                                  int64_t rows, int64_t dim, float eps, int plus_one, void* workspace, size_t workspace_bytes,
                                  mft_stream_t stream) {
    if (gx_bf16 && ld_bf16 < dim) return set_error(1, "mft_rmsnorm_bwd_ex: bf16 pitch < dim");
    return rmsnorm_bwd_impl(x, w, gy, gx, dw, rows, dim, eps, plus_one, workspace, workspace_bytes, stream, gx_bf16, ld_bf16);
}

static int rmsnorm_bwd_impl(const float* x, const float* w, const float* gy, float* gx, float* dw, int64_t rows, int64_t dim, float eps,
                            int plus_one, void* workspace, size_t workspace_bytes, mft_stream_t stream, void* gx16, int64_t ld16) {
    if (dim == 0) return 0;
    if (!x || !w || !gy || !gx) return set_error(1, "mft_rmsnorm_bwd: null pointer");
    if (rows == 0) return dw ? mft_fill(dw, dim, 0.f, stream) : 0;
    if (dw && (!workspace || workspace_bytes < mft_norm_bwd_workspace_bytes(rows, dim)))
        return set_error(1, "mft_rmsnorm_bwd: workspace too small");
    if (size_t(kNormWarps) * dim * sizeof(float) > 200 * 1024) return set_error(1, "mft_rmsnorm_bwd: dim %lld too large", (long long)dim);
    cudaStream_t st = as_stream(stream);
    static bool attr_set[kMaxDevices] = {};
    if (first_use_on_device(attr_set))
        cudaFuncSetAttribute(rmsnorm_bwd_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, 200 * 1024);
    int parts = dw ? norm_bwd_ctas(rows) : int(std::min<int64_t>(ceil_div(rows, kNormWarps), int64_t(kNumSMs) * 8));
    float* partial = dw ? static_cast<float*>(workspace) : nullptr;     // dw == NULL (frozen weight): dx only, no second stage
    pdl_launch(rmsnorm_bwd_kernel, parts, kNormWarps * 32, size_t(kNormWarps) * dim * sizeof(float), st)(
        x, w, gy, gx, partial, rows, int(dim), eps, plus_one, static_cast<__nv_bfloat16*>(gx16), ld16);
    if (dw) {
        pdl_launch(norm_wgrad_final, unsigned(ceil_div(dim, 32)), 256, 0, st)(partial, dw, nullptr, parts, int(dim), 1);
        add_launches(1);
    }
    MFT_CHECK_LAUNCH("mft_rmsnorm_bwd");
    return 0;
}

extern "C" int mft_softmax_fwd(const float* x, float* y, int64_t rows, int64_t dim, mft_stream_t stream) {
    if (rows == 0 || dim == 0) return 0;
    if (!x || !y) return set_error(1, "mft_softmax_fwd: null pointer");
    pdl_launch(softmax_fwd_kernel, stream_grid(rows * 32, 256), 256, 0, as_stream(stream))(x, y, rows, int(dim));
    MFT_CHECK_LAUNCH("mft_softmax_fwd");
    return 0;
}

extern "C" int mft_softmax_bwd(const float* y, const float* gy, float* gx, int64_t rows, int64_t dim, mft_stream_t stream) {
    if (rows == 0 || dim == 0) return 0;
    if (!y || !gy || !gx) return set_error(1, "mft_softmax_bwd: null pointer");
    pdl_launch(softmax_bwd_kernel, stream_grid(rows * 32, 256), 256, 0, as_stream(stream))(y, gy, gx, rows, int(dim));
    MFT_CHECK_LAUNCH("mft_softmax_bwd");
    return 0;
}

// ---------------------------------------------------------------- LayerNorm backward without caller scratch
// Some users of this library cannot hand us workspace: a cgo/JNI/ctypes front end that only forwards the tensors
// its host application gave it owns no device allocator, and a steady-state training loop may forbid cudaMalloc
// because it synchronises the whole device. For them the column-reduction partials live in this fixed array, which
// every GPU that loads the library keeps resident until the process exits.
#define MFT_LN_BWD_STATIC_FLOATS (size_t(mft::kNormMaxCtas) * 2 * 2048)      /* 1,212,416 floats = 4,849,664 bytes */
static __device__ __align__(256) float g_ln_bwd_static_scratch[MFT_LN_BWD_STATIC_FLOATS];

namespace {
float* g_ln_bwd_static_ptr[kMaxDevices] = {nullptr};   // device address of the array, resolved once per device
}  // namespace

extern "C" int mft_layernorm_bwd_static(const float* x, const float* gamma, const float* gy, float* gx, float* dgamma,
                                        float* dbeta, int64_t rows, int64_t dim, float eps, mft_stream_t stream) {
    const size_t need = mft_norm_bwd_workspace_bytes(rows, dim);
    const size_t have = MFT_LN_BWD_STATIC_FLOATS * sizeof(float);
    if (need > have)
        return set_error(1, "mft_layernorm_bwd_static: rows=%lld dim=%lld needs %zu bytes of scratch, the built-in array holds %zu; "
                            "use mft_layernorm_bwd with caller-provided workspace", (long long)rows, (long long)dim, need, have);
    int dev = 0;
    MFT_CUDA_OK(cudaGetDevice(&dev), "mft_layernorm_bwd_static");
    if (dev < 0 || dev >= kMaxDevices) return set_error(1, "mft_layernorm_bwd_static: device index %d out of range", dev);
    if (!g_ln_bwd_static_ptr[dev]) {
        // In host code the array's name is only a handle; the address kernels can dereference comes from the runtime,
        // separately on each device (done here on first use with that device current).
        float* p = nullptr;
        MFT_CUDA_OK(cudaGetSymbolAddress(reinterpret_cast<void**>(&p), g_ln_bwd_static_scratch), "mft_layernorm_bwd_static(cudaGetSymbolAddress)");
        if (!p) return set_error(2, "mft_layernorm_bwd_static: cudaGetSymbolAddress returned a null address");
        g_ln_bwd_static_ptr[dev] = p;
    }
    // The partial sums are fully rewritten by the first stage on every call, so stale contents from earlier calls are
    // harmless and nothing has to be cleared.
    return mft_layernorm_bwd(x, gamma, gy, gx, dgamma, dbeta, rows, dim, eps, g_ln_bwd_static_ptr[dev], have, stream);
}
