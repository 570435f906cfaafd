// Shifted language-model cross-entropy over a 50k..262k vocabulary (core/lm_loss.cpp:19-210).
// HBM-bound: forward reads each valid logits row once from HBM (the second pass hits L2), backward
// reads it once more and writes dlogits once. Rows whose label is ignored (and every last position)
// are never read. Per-row losses are reduced in a fixed order (deterministic).
#include "common.cuh"

namespace mft {

__device__ __forceinline__ float exp2f_approx(float x) {          // one MUFU.EX2
    float y;
    asm("ex2.approx.ftz.f32 %0, %1;" : "=f"(y) : "f"(x));
    return y;
}


constexpr int kCeThreads = 512;

__device__ __forceinline__ int ce_label(const int32_t* labels, int64_t row, int S, int V, int ignore_index) {
    // row = b*S + s is scored against labels[b, s+1]; returns -1 when the row is skipped.
    int s = int(row % S);
    if (s + 1 >= S) return -1;
    int y = labels[row + 1];
    if (y == ignore_index || y < 0 || y >= V) return -1;
    return y;
}

__global__ void __launch_bounds__(kCeThreads)
lm_ce_fwd_kernel(const float* __restrict__ logits, int64_t ld, const int32_t* __restrict__ labels, int64_t rows, int S,
                 int V, int ignore_index, float* __restrict__ row_lse, double* __restrict__ row_nll) {
    pdl_sync();
    __shared__ float sred[32];
    __shared__ double dred[32];
    const bool vec = ((ld & 3) == 0) && ((reinterpret_cast<uintptr_t>(logits) & 15u) == 0);
    for (int64_t row = blockIdx.x; row < rows; row += gridDim.x) {
        const int y = ce_label(labels, row, S, V, ignore_index);
        if (y < 0) {
            if (threadIdx.x == 0) { row_nll[row] = 0.0; row_lse[row] = 0.f; }
            continue;
        }
        const float* x = logits + row * ld;
        float m = -INFINITY;
        if (vec) {
            const float4* x4 = reinterpret_cast<const float4*>(x);
            const int v4 = V >> 2;
            for (int i = threadIdx.x; i < v4; i += kCeThreads) {
                float4 a = x4[i];
                m = fmaxf(m, fmaxf(fmaxf(a.x, a.y), fmaxf(a.z, a.w)));
            }
            for (int i = (v4 << 2) + threadIdx.x; i < V; i += kCeThreads) m = fmaxf(m, x[i]);
        } else {
            for (int i = threadIdx.x; i < V; i += kCeThreads) m = fmaxf(m, x[i]);
        }
        m = block_max(m, sred);
        float s = 0.f;
        if (vec) {
            const float4* x4 = reinterpret_cast<const float4*>(x);
            const int v4 = V >> 2;
            for (int i = threadIdx.x; i < v4; i += kCeThreads) {
                float4 a = x4[i];
                s += expf(a.x - m) + expf(a.y - m) + expf(a.z - m) + expf(a.w - m);
            }
            for (int i = (v4 << 2) + threadIdx.x; i < V; i += kCeThreads) s += expf(x[i] - m);
        } else {
            for (int i = threadIdx.x; i < V; i += kCeThreads) s += expf(x[i] - m);
        }
        const double denom = block_sum<double>(double(s), dred);
        if (threadIdx.x == 0) {
            double lse = double(m) + log(denom);
            row_lse[row] = float(lse);
            row_nll[row] = lse - double(x[y]);
        }
    }
}

// Single-CTA fixed-order reduction of the per-row losses and of the valid-row count.
__global__ void __launch_bounds__(1024)
lm_ce_reduce_kernel(const double* __restrict__ row_nll, const int32_t* __restrict__ labels, int64_t rows, int S, int V,
                    int ignore_index, int reduction, float* __restrict__ loss_out, int32_t* __restrict__ valid_out) {
    pdl_sync();
    __shared__ double dred[32];
    __shared__ double cred[32];
    double s = 0.0, c = 0.0;
    for (int64_t r = threadIdx.x; r < rows; r += blockDim.x) {
        if (ce_label(labels, r, S, V, ignore_index) >= 0) { s += row_nll[r]; c += 1.0; }
    }
    s = block_sum<double>(s, dred);
    c = block_sum<double>(c, cred);
    if (threadIdx.x == 0) {
        int64_t cnt = int64_t(c + 0.5);
        float out;
        if (reduction == 1) out = float(s);
        else out = cnt > 0 ? float(s / double(cnt)) : 0.0f;
        loss_out[0] = out;
        valid_out[0] = int32_t(cnt);
    }
}

// softmax(x) * scale for four logits; FAST (the gradient leaves as bf16 only): one ex2.approx per element
template <bool FAST>
__device__ __forceinline__ float4 ce_grad4(float4 a, float lse, float lse_log2, float scale) {
    float4 r;
    if (FAST) {
        const float c = 1.4426950408889634f;
        r.x = exp2f_approx(fmaf(a.x, c, -lse_log2)) * scale; r.y = exp2f_approx(fmaf(a.y, c, -lse_log2)) * scale;
        r.z = exp2f_approx(fmaf(a.z, c, -lse_log2)) * scale; r.w = exp2f_approx(fmaf(a.w, c, -lse_log2)) * scale;
    } else {
        r.x = expf(a.x - lse) * scale; r.y = expf(a.y - lse) * scale;
        r.z = expf(a.z - lse) * scale; r.w = expf(a.w - lse) * scale;
    }
    return r;
}
__device__ __forceinline__ void ce_fix_label(float4& r, int base, int y, float scale) {
    if (y >= base && y < base + 4) {
        if (y == base) r.x -= scale; else if (y == base + 1) r.y -= scale;
        else if (y == base + 2) r.z -= scale; else r.w -= scale;
    }
}

template <bool FAST>
__global__ void __launch_bounds__(kCeThreads)
lm_ce_bwd_kernel(const float* __restrict__ logits, int64_t ld, const int32_t* __restrict__ labels,
                 const float* __restrict__ row_lse, const int32_t* __restrict__ valid, const float* __restrict__ gout,
                 int64_t rows, int S, int V, int ignore_index, int reduction, float* __restrict__ dlogits, int64_t ldg,
                 __nv_bfloat16* __restrict__ d16, int64_t ld16) {
    pdl_sync();   // d16 != NULL: bf16 copy of dlogits (vectorised path only)
    // dlogits == NULL: the bf16 copy is the only result (launcher guarantees the vectorised layout then)
    const bool vec = ((ld & 3) == 0) && (!dlogits || (ldg & 3) == 0) &&
                     (((reinterpret_cast<uintptr_t>(logits) | reinterpret_cast<uintptr_t>(dlogits)) & 15u) == 0);
    const int cnt = valid[0];
    const float g = gout ? gout[0] : 1.0f;
    const float scale = (reduction == 1) ? g : (cnt > 0 ? g / float(cnt) : 0.0f);
    for (int64_t row = blockIdx.x; row < rows; row += gridDim.x) {
        const int y = ce_label(labels, row, S, V, ignore_index);
        float* d = dlogits ? dlogits + row * ldg : nullptr;
        __nv_bfloat16* h = d16 ? d16 + row * ld16 : nullptr;
        if (y < 0) {   // skipped rows get an all-zero gradient (core/lm_loss.cpp:45-52)
            if (vec) {
                float4* d4 = reinterpret_cast<float4*>(d);
                const int v4 = V >> 2;
                const float4 z = make_float4(0.f, 0.f, 0.f, 0.f);
                for (int i = threadIdx.x; i < v4; i += kCeThreads) { if (d) st_stream(d4 + i, z); if (h) st_bf16x4(h + i * 4, z); }
                for (int i = (v4 << 2) + threadIdx.x; i < V; i += kCeThreads) { if (d) d[i] = 0.f; if (h) h[i] = __float2bfloat16_rn(0.f); }
            } else {
                for (int i = threadIdx.x; i < V; i += kCeThreads) d[i] = 0.f;
            }
            continue;
        }
        const float* x = logits + row * ld;
        const float lse = row_lse[row], lse_log2 = lse * 1.4426950408889634f;
        if (vec) {
            const float4* x4 = reinterpret_cast<const float4*>(x);
            float4* d4 = reinterpret_cast<float4*>(d);
            const int v4 = V >> 2;
            int i = threadIdx.x;
            for (; i + 3 * kCeThreads < v4; i += 4 * kCeThreads) {      // four independent 16-byte loads in flight per thread
                float4 a[4];
#pragma unroll
                for (int u = 0; u < 4; ++u) a[u] = ld_stream(x4 + i + u * kCeThreads);
#pragma unroll
                for (int u = 0; u < 4; ++u) {
                    const int iu = i + u * kCeThreads;
                    float4 r = ce_grad4<FAST>(a[u], lse, lse_log2, scale);
                    ce_fix_label(r, iu << 2, y, scale);
                    if (d) st_stream(d4 + iu, r);
                    if (h) st_bf16x4(h + iu * 4, r);
                }
            }
            for (; i < v4; i += kCeThreads) {
                float4 r = ce_grad4<FAST>(ld_stream(x4 + i), lse, lse_log2, scale);
                ce_fix_label(r, i << 2, y, scale);
                if (d) st_stream(d4 + i, r);
                if (h) st_bf16x4(h + i * 4, r);
            }
            for (int j = (v4 << 2) + threadIdx.x; j < V; j += kCeThreads) {
                float r = expf(x[j] - lse) * scale;
                if (j == y) r -= scale;
                if (d) d[j] = r;
                if (h) h[j] = __float2bfloat16_rn(r);
            }
        } else {
            for (int i = threadIdx.x; i < V; i += kCeThreads) {
                float r = expf(x[i] - lse) * scale;
                if (i == y) r -= scale;
                d[i] = r;
            }
        }
    }
}

}  // namespace mft

using namespace mft;

extern "C" size_t mft_lm_ce_workspace_bytes(int64_t rows) { return size_t(rows) * sizeof(double); }

extern "C" int mft_lm_ce_fwd(const float* logits, int64_t ld, const int32_t* labels, int32_t B, int32_t S, int32_t V,
                             int32_t ignore_index, int32_t reduction, float* loss_out, int32_t* valid_out,
                             float* row_lse, void* workspace, size_t workspace_bytes, mft_stream_t stream) {
    if (!logits || !labels || !loss_out || !valid_out || !row_lse) return set_error(1, "mft_lm_ce_fwd: null pointer");
    if (ld < V) return set_error(1, "mft_lm_ce_fwd: ld (%lld) < V (%d)", (long long)ld, V);
    if (reduction != 0 && reduction != 1) return set_error(1, "mft_lm_ce_fwd: reduction must be 0 (mean) or 1 (sum)");
    const int64_t rows = int64_t(B) * S;
    if (!workspace || workspace_bytes < mft_lm_ce_workspace_bytes(rows))
        return set_error(1, "mft_lm_ce_fwd: workspace too small");
    cudaStream_t st = as_stream(stream);
    double* row_nll = static_cast<double*>(workspace);
    if (rows > 0 && V > 0) {
        int grid = int(rows < int64_t(kNumSMs) * 4 ? rows : int64_t(kNumSMs) * 4);
        pdl_launch(lm_ce_fwd_kernel, grid, kCeThreads, 0, st)(logits, ld, labels, rows, S, V, ignore_index, row_lse, row_nll);
    }
    pdl_launch(lm_ce_reduce_kernel, 1, 1024, 0, st)(row_nll, labels, rows, S, V, ignore_index, reduction, loss_out, valid_out);
    add_launches(1);
    MFT_CHECK_LAUNCH("mft_lm_ce_fwd");
    return 0;
}

static int ce_bwd_impl(const float* logits, int64_t ld, const int32_t* labels, const float* row_lse,
                       const int32_t* valid, const float* gout, int32_t B, int32_t S, int32_t V,
                       int32_t ignore_index, int32_t reduction, float* dlogits, int64_t ldg, void* d16, int64_t ld16, mft_stream_t stream) {
    if (!logits || !labels || !row_lse || !valid || (!dlogits && !d16)) return set_error(1, "mft_lm_ce_bwd: null pointer");
    if (ld < V || (dlogits && ldg < V) || (d16 && ld16 < V)) return set_error(1, "mft_lm_ce_bwd: leading dimension < V");
    const int64_t rows = int64_t(B) * S;
    if (rows == 0 || V == 0) return 0;
    // the bf16 copy is written by the vectorised path; other layouts get a separate cast
    const bool vec = ((ld & 3) == 0) && (!dlogits || (ldg & 3) == 0) && aligned16(logits) && (!dlogits || aligned16(dlogits));
    const bool fused16 = d16 && vec && (ld16 & 3) == 0 && (reinterpret_cast<uintptr_t>(d16) & 7u) == 0;
    if (!dlogits && !fused16) return set_error(1, "mft_lm_ce_bwd_ex: a bf16-only gradient needs 16-byte aligned logits with a pitch %% 4 == 0 and an 8-byte aligned bf16 buffer with a pitch %% 4 == 0");
    int grid = int(rows < int64_t(kNumSMs) * 4 ? rows : int64_t(kNumSMs) * 4);
    __nv_bfloat16* h16 = fused16 ? static_cast<__nv_bfloat16*>(d16) : nullptr;
    if (!dlogits) pdl_launch(lm_ce_bwd_kernel<true>, grid, kCeThreads, 0, as_stream(stream))(logits, ld, labels, row_lse, valid, gout, rows, S, V, ignore_index, reduction, dlogits, ldg, h16, ld16);
    else pdl_launch(lm_ce_bwd_kernel<false>, grid, kCeThreads, 0, as_stream(stream))(logits, ld, labels, row_lse, valid, gout, rows, S, V, ignore_index, reduction, dlogits, ldg, h16, ld16);
    MFT_CHECK_LAUNCH("mft_lm_ce_bwd");
    if (d16 && !fused16) return mft_cast_bf16(dlogits, ldg, d16, ld16, rows, V, stream);
    return 0;
}

extern "C" int mft_lm_ce_bwd(const float* logits, int64_t ld, const int32_t* labels, const float* row_lse,
                             const int32_t* valid, const float* gout, int32_t B, int32_t S, int32_t V,
                             int32_t ignore_index, int32_t reduction, float* dlogits, int64_t ldg, mft_stream_t stream) {
    return ce_bwd_impl(logits, ld, labels, row_lse, valid, gout, B, S, V, ignore_index, reduction, dlogits, ldg, nullptr, 0, stream);
}

// same, plus a bf16 copy of dlogits (row pitch ld_bf16) for the bf16 lm_head products that consume it
extern "C" int mft_lm_ce_bwd_ex(const float* logits, int64_t ld, const int32_t* labels, const float* row_lse,
                                const int32_t* valid, const float* gout, int32_t B, int32_t S, int32_t V,
                                int32_t ignore_index, int32_t reduction, float* dlogits, int64_t ldg, void* dlogits_bf16, int64_t ld_bf16,
                                mft_stream_t stream) {
    return ce_bwd_impl(logits, ld, labels, row_lse, valid, gout, B, S, V, ignore_index, reduction, dlogits, ldg, dlogits_bf16, ld_bf16, stream);
}
