// Per-head RMSNorm (1 + w) followed by interleaved-pair RoPE, fused, for the query and key projections of one layer in ONE
// launch; and its backward (inverse rotation, then RMSNorm backward). Replaces, per layer and per tensor, the reference's
// permute -> rms_norm (core/ops.cpp:1489-1537) -> apply_rope (core/ops.cpp:2151-2225) nodes of graph/gemma_model.cpp:383-420
// and RMSNormBackward / ApplyRoPEBackward (core/backward_functions.cpp:483-523, 718-763).
// Rows are heads: x viewed as [tokens * heads, Hd] in the projection's own [B,S,H,Hd] layout (no permute); the position of
// row r is (r / heads) % S. Arithmetic as the reference: sum of squares and the backward dot in double, rms in float,
// freq = 1 / powf(theta, 2d/Hd), angle = pos * freq, cosf / sinf. HBM-bound: 8 B per element forward, 12 B backward.
#include "common.cuh"

namespace mft {

struct NormRopeSide {
    const float* x; const float* w; const float* gy; float* y; int64_t rows; int heads;
};

template <int MAXP>   // pairs per lane: Hd <= 64 * MAXP
__global__ void __launch_bounds__(256)
qk_norm_rope_fwd_kernel(NormRopeSide a, NormRopeSide b, int S, int Hd, float eps, float theta, int do_norm) {
    pdl_sync();
    const int lane = threadIdx.x & 31;
    const int half = Hd >> 1;
    const int64_t warps = (int64_t(gridDim.x) * blockDim.x) >> 5, total = a.rows + b.rows;
    for (int64_t r = (blockIdx.x * int64_t(blockDim.x) + threadIdx.x) >> 5; r < total; r += warps) {
        const NormRopeSide& sd = r < a.rows ? a : b;
        const int64_t row = r < a.rows ? r : r - a.rows;
        const float2* xr = reinterpret_cast<const float2*>(sd.x + row * Hd);
        float2 v[MAXP];
        double sq = 0.0;
#pragma unroll
        for (int i = 0; i < MAXP; ++i) {
            const int d = lane + 32 * i;
            v[i] = d < half ? xr[d] : make_float2(0.f, 0.f);
            sq += double(v[i].x) * double(v[i].x) + double(v[i].y) * double(v[i].y);
        }
        sq = warp_sum(sq);
        const float rms = float(sqrt(sq / double(Hd) + double(eps)));
        const float pos = float((row / sd.heads) % S);
        float2* yr = reinterpret_cast<float2*>(sd.y + row * Hd);
#pragma unroll
        for (int i = 0; i < MAXP; ++i) {
            const int d = lane + 32 * i;
            if (d >= half) continue;
            float n0 = v[i].x, n1 = v[i].y;
            if (do_norm) {
                const float2 w2 = reinterpret_cast<const float2*>(sd.w)[d];
                n0 = (n0 / rms) * (1.0f + w2.x);
                n1 = (n1 / rms) * (1.0f + w2.y);
            }
            const float freq = 1.0f / powf(theta, 2.0f * float(d) / float(Hd));
            const float angle = pos * freq;
            const float c = cosf(angle), s = sinf(angle);
            yr[d] = make_float2(n0 * c - n1 * s, n0 * s + n1 * c);
        }
    }
}

// gx = RMSNormBackward(x, w, RoPE^-1(gy)); the weight is frozen on this path (its gradient is not produced)
template <int MAXP>
__global__ void __launch_bounds__(256)
qk_norm_rope_bwd_kernel(NormRopeSide a, int S, int Hd, float eps, float theta, int do_norm) {
    pdl_sync();
    const int lane = threadIdx.x & 31;
    const int half = Hd >> 1;
    const int64_t warps = (int64_t(gridDim.x) * blockDim.x) >> 5;
    for (int64_t row = (blockIdx.x * int64_t(blockDim.x) + threadIdx.x) >> 5; row < a.rows; row += warps) {
        const float2* xr = reinterpret_cast<const float2*>(a.x + row * Hd);
        const float2* gr = reinterpret_cast<const float2*>(a.gy + row * Hd);
        const float pos = float((row / a.heads) % S);
        float2 xv[MAXP], dn[MAXP], we[MAXP];
        double sq = 0.0, dot = 0.0;
#pragma unroll
        for (int i = 0; i < MAXP; ++i) {
            const int d = lane + 32 * i;
            xv[i] = dn[i] = make_float2(0.f, 0.f);
            we[i] = make_float2(1.f, 1.f);
            if (d < half) {
                xv[i] = xr[d];
                const float2 gv = gr[d];
                const float freq = 1.0f / powf(theta, 2.0f * float(d) / float(Hd));
                const float angle = pos * freq;
                const float c = cosf(angle), s = -sinf(angle);           // rotation by -angle
                dn[i] = make_float2(gv.x * c - gv.y * s, gv.x * s + gv.y * c);
                if (do_norm) { const float2 w2 = reinterpret_cast<const float2*>(a.w)[d]; we[i] = make_float2(1.0f + w2.x, 1.0f + w2.y); }
            }
            sq += double(xv[i].x) * double(xv[i].x) + double(xv[i].y) * double(xv[i].y);
            dot += double(dn[i].x) * double(we[i].x) * double(xv[i].x) + double(dn[i].y) * double(we[i].y) * double(xv[i].y);
        }
        float2* gx = reinterpret_cast<float2*>(a.y + row * Hd);
        if (!do_norm) {
#pragma unroll
            for (int i = 0; i < MAXP; ++i) { const int d = lane + 32 * i; if (d < half) gx[d] = dn[i]; }
            continue;
        }
        sq = warp_sum(sq);
        dot = warp_sum(dot);
        const double inv = 1.0 / sqrt(sq / double(Hd) + double(eps));
        const double coeff = inv * inv * inv / double(Hd);
#pragma unroll
        for (int i = 0; i < MAXP; ++i) {
            const int d = lane + 32 * i;
            if (d >= half) continue;
            gx[d] = make_float2(float(double(dn[i].x) * double(we[i].x) * inv - double(xv[i].x) * coeff * dot),
                                float(double(dn[i].y) * double(we[i].y) * inv - double(xv[i].y) * coeff * dot));
        }
    }
}

}  // namespace mft

using namespace mft;

extern "C" int mft_qk_norm_rope_fwd(const float* q, const float* wq, float* yq, int64_t q_tokens, int32_t q_heads,
                                    const float* k, const float* wk, float* yk, int64_t k_tokens, int32_t k_heads,
                                    int32_t S, int32_t Hd, float eps, float theta, int do_norm, mft_stream_t stream) {
    const int64_t rq = q ? q_tokens * q_heads : 0, rk = k ? k_tokens * k_heads : 0;
    if (rq + rk == 0) return 0;
    if ((q && !yq) || (k && !yk) || (do_norm && ((q && !wq) || (k && !wk)))) return set_error(1, "mft_qk_norm_rope_fwd: null pointer");
    if (Hd < 2 || (Hd & 1) || Hd > 512) return set_error(1, "mft_qk_norm_rope_fwd: head_dim must be even and <= 512");
    if (S < 1) return set_error(1, "mft_qk_norm_rope_fwd: S must be positive");
    NormRopeSide a{q, wq, nullptr, yq, rq, q_heads > 0 ? q_heads : 1}, b{k, wk, nullptr, yk, rk, k_heads > 0 ? k_heads : 1};
    const int grid = stream_grid((rq + rk) * 32, 256);
    if (Hd <= 128) pdl_launch(qk_norm_rope_fwd_kernel<2>, grid, 256, 0, as_stream(stream))(a, b, S, Hd, eps, theta, do_norm);
    else pdl_launch(qk_norm_rope_fwd_kernel<8>, grid, 256, 0, as_stream(stream))(a, b, S, Hd, eps, theta, do_norm);
    MFT_CHECK_LAUNCH("mft_qk_norm_rope_fwd");
    return 0;
}

extern "C" int mft_qk_norm_rope_bwd(const float* x, const float* w, const float* gy, float* gx, int64_t tokens, int32_t heads,
                                    int32_t S, int32_t Hd, float eps, float theta, int do_norm, mft_stream_t stream) {
    const int64_t rows = tokens * heads;
    if (rows == 0) return 0;
    if (!x || !gy || !gx || (do_norm && !w)) return set_error(1, "mft_qk_norm_rope_bwd: null pointer");
    if (Hd < 2 || (Hd & 1) || Hd > 512) return set_error(1, "mft_qk_norm_rope_bwd: head_dim must be even and <= 512");
    if (S < 1 || heads < 1) return set_error(1, "mft_qk_norm_rope_bwd: S and heads must be positive");
    NormRopeSide a{x, w, gy, gx, rows, heads};
    const int grid = stream_grid(rows * 32, 256);
    if (Hd <= 128) pdl_launch(qk_norm_rope_bwd_kernel<2>, grid, 256, 0, as_stream(stream))(a, S, Hd, eps, theta, do_norm);
    else pdl_launch(qk_norm_rope_bwd_kernel<8>, grid, 256, 0, as_stream(stream))(a, S, Hd, eps, theta, do_norm);
    MFT_CHECK_LAUNCH("mft_qk_norm_rope_bwd");
    return 0;
}
