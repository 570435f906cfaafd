// Causal self-attention forward (GPT-2 semantics: core/memory_efficient_attention.cpp:40-185).
//
// Flash-style: one CTA = 64 query rows of one (batch, head); 4 warps x 16 rows; K/V streamed in
// 64-key tiles through shared memory with cp.async double buffering; scores and P*V on the tensor
// pipe (mma.sync m16n8k8 TF32) with the 3xTF32 split (hi/lo) so results stay at fp32 accuracy
// (the reference is fp32; attention is <5% of the step's FLOPs so the 3x cost is immaterial).
// No S x S matrix is ever written to HBM. q/k/v are read with arbitrary (batch, seq, head) strides,
// so the fused [B,S,3C] qkv projection output is consumed in place and ctx is written as [B,S,C].
#include <cstdlib>
#include "common.cuh"

namespace mft {

constexpr int kAttBQ = 64;   // query rows per CTA
constexpr int kAttBK = 64;   // keys per tile
constexpr int kAttPad = 4;   // smem row padding (floats): conflict-free B-fragment reads

__device__ __forceinline__ void cp_async16(void* smem, const void* gmem, bool pred) {
    uint32_t s = static_cast<uint32_t>(__cvta_generic_to_shared(smem));
    int bytes = pred ? 16 : 0;   // src-size 0 => zero fill
    asm volatile("cp.async.cg.shared.global [%0], [%1], 16, %2;" :: "r"(s), "l"(gmem), "r"(bytes) : "memory");
}
__device__ __forceinline__ void cp_async_commit() { asm volatile("cp.async.commit_group;" ::: "memory"); }
template <int N>
__device__ __forceinline__ void cp_async_wait() { asm volatile("cp.async.wait_group %0;" :: "n"(N) : "memory"); }

// PRECISE = true : 3xTF32 split (hi/lo) on both products, fp32-accurate (exact mode / parity tests).
// PRECISE = false: plain TF32 operands (same input precision as the tcgen05 GEMMs), 3x fewer MMAs. K and V go to the tensor
//                  pipe as raw fp32 bit patterns (the pipe reads the upper 19 bits: truncation instead of a cvt.rna per
//                  fragment element, which would cost more issue slots than the MMA it feeds); Q and P are rounded.
// MT = 16-row query slabs per warp: every K/V fragment read from shared memory feeds MT MMAs.
template <int HD, bool PRECISE, int MT>
__global__ void __launch_bounds__(128)
causal_attention_fwd_kernel(const float* __restrict__ q, const float* __restrict__ k, const float* __restrict__ v,
                            float* __restrict__ ctx, int H, int S,
                            int64_t qsb, int64_t qss, int64_t qsh, int64_t ksb, int64_t kss, int64_t ksh,
                            int64_t osb, int64_t oss, int64_t osh, float scale, int window,
                            __nv_bfloat16* __restrict__ ctx16) {
    pdl_sync();   // ctx16 != NULL: bf16 copy of ctx, same element strides
    constexpr int LD = HD + kAttPad;
    constexpr int KS = HD / 8;    // k-slabs over the head dim (QK^T)
    constexpr int NT = HD / 8;    // n-tiles over the head dim (PV)
    constexpr int BQ = 64 * MT;   // query rows per CTA (4 warps x MT slabs x 16 rows)
    constexpr float kLog2e = 1.4426950408889634f;
    extern __shared__ __align__(16) float smem[];
    float* sK = smem;                          // [2][kAttBK][LD]
    float* sV = smem + 2 * kAttBK * LD;        // [2][kAttBK][LD]

    const int qt = gridDim.x - 1 - blockIdx.x; // heavy (late) query tiles first
    const int bh = blockIdx.y;
    const int b = bh / H, h = bh % H;
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const int g = lane >> 2, t = lane & 3;
    const int q0 = qt * BQ;
    const int wrow = q0 + warp * 16 * MT;      // first query row of this warp
    const float* qb = q + b * qsb + h * qsh;
    const float* kb = k + b * ksb + h * ksh;
    const float* vb = v + b * ksb + h * ksh;
    // fast mode: scores are kept in the log2 domain (scale * log2(e) folded into Q) so the softmax uses ex2 directly
    const float qscale = PRECISE ? scale : scale * kLog2e;

    // Q fragments (A operand, rows g / g+8 of each 16-row slab), pre-scaled, split hi/lo in precise mode.
    uint32_t qhi[MT][KS][4], qlo[PRECISE ? MT : 1][PRECISE ? KS : 1][4];
#pragma unroll
    for (int m = 0; m < MT; ++m) {
        const int r0 = wrow + m * 16 + g, r1 = r0 + 8;
#pragma unroll
        for (int s = 0; s < KS; ++s) {
            float a0 = (r0 < S) ? qb[int64_t(r0) * qss + s * 8 + t] * qscale : 0.f;
            float a1 = (r1 < S) ? qb[int64_t(r1) * qss + s * 8 + t] * qscale : 0.f;
            float a2 = (r0 < S) ? qb[int64_t(r0) * qss + s * 8 + t + 4] * qscale : 0.f;
            float a3 = (r1 < S) ? qb[int64_t(r1) * qss + s * 8 + t + 4] * qscale : 0.f;
            if (PRECISE) {
                split_tf32(a0, qhi[m][s][0], qlo[m][s][0]); split_tf32(a1, qhi[m][s][1], qlo[m][s][1]);
                split_tf32(a2, qhi[m][s][2], qlo[m][s][2]); split_tf32(a3, qhi[m][s][3], qlo[m][s][3]);
            } else {
                qhi[m][s][0] = f2tf32(a0); qhi[m][s][1] = f2tf32(a1); qhi[m][s][2] = f2tf32(a2); qhi[m][s][3] = f2tf32(a3);
            }
        }
    }

    float o[MT][NT][4];
    float mrow[MT][2], lrow[MT][2];
#pragma unroll
    for (int m = 0; m < MT; ++m) {
#pragma unroll
        for (int n = 0; n < NT; ++n) { o[m][n][0] = o[m][n][1] = o[m][n][2] = o[m][n][3] = 0.f; }
        mrow[m][0] = mrow[m][1] = -INFINITY;
        lrow[m][0] = lrow[m][1] = 0.f;
    }

    // key-tile range: causal upper bound (last key of the CTA's last row), optional sliding-window lower bound
    const int last_row = min(q0 + BQ, S) - 1;
    const int kt_hi = last_row / kAttBK;               // last tile index (inclusive)
    int kt_lo = 0;
    if (window > 0) { int lo = q0 - (window - 1); if (lo > 0) kt_lo = lo / kAttBK; }

    auto load_tile = [&](int kt, int buf) {
        const int key0 = kt * kAttBK;
        float* dk = sK + buf * kAttBK * LD;
        float* dv = sV + buf * kAttBK * LD;
        constexpr int CH = HD / 4;                    // 16-byte chunks per row
        for (int c = threadIdx.x; c < kAttBK * CH; c += 128) {
            int r = c / CH, cc = c % CH;
            bool ok = (key0 + r) < S;
            int64_t off = int64_t(ok ? key0 + r : 0) * kss + cc * 4;
            cp_async16(dk + r * LD + cc * 4, kb + off, ok);
            cp_async16(dv + r * LD + cc * 4, vb + off, ok);
        }
        cp_async_commit();
    };

    load_tile(kt_lo, 0);
    for (int kt = kt_lo; kt <= kt_hi; ++kt) {
        const int buf = (kt - kt_lo) & 1;
        if (kt + 1 <= kt_hi) { load_tile(kt + 1, buf ^ 1); cp_async_wait<1>(); } else { cp_async_wait<0>(); }
        __syncthreads();
        const float* tk = sK + buf * kAttBK * LD;
        const float* tv = sV + buf * kAttBK * LD;
        const int key0 = kt * kAttBK;
        // a warp whose rows all precede this key tile has nothing to add (only possible when MT > 1)
        const bool active = key0 <= wrow + 16 * MT - 1;
        if (active) {
            // ---- S = (Q*scale) K^T  : (16*MT) x 64 per warp
            float sc[MT][8][4];
#pragma unroll
            for (int n = 0; n < 8; ++n) {
#pragma unroll
                for (int m = 0; m < MT; ++m) sc[m][n][0] = sc[m][n][1] = sc[m][n][2] = sc[m][n][3] = 0.f;
#pragma unroll
                for (int s = 0; s < KS; ++s) {
                    const float k0 = tk[(n * 8 + g) * LD + s * 8 + t];
                    const float k1 = tk[(n * 8 + g) * LD + s * 8 + t + 4];
                    if (PRECISE) {
                        uint32_t bh0, bl0, bh1, bl1;
                        split_tf32(k0, bh0, bl0); split_tf32(k1, bh1, bl1);
#pragma unroll
                        for (int m = 0; m < MT; ++m) {
                            mma_tf32(sc[m][n], qlo[m][s], bh0, bh1);
                            mma_tf32(sc[m][n], qhi[m][s], bl0, bl1);
                            mma_tf32(sc[m][n], qhi[m][s], bh0, bh1);
                        }
                    } else {
#pragma unroll
                        for (int m = 0; m < MT; ++m) mma_tf32(sc[m][n], qhi[m][s], __float_as_uint(k0), __float_as_uint(k1));
                    }
                }
            }
            // ---- masks (causal / window / sequence tail) + online softmax
            uint32_t phi[MT][8][4];
#pragma unroll
            for (int m = 0; m < MT; ++m) {
                const int r0 = wrow + m * 16 + g, r1 = r0 + 8;
                const bool need_mask = (key0 + kAttBK - 1 > wrow + m * 16) || (window > 0) || (key0 + kAttBK > S);
                float mx0 = -INFINITY, mx1 = -INFINITY;
#pragma unroll
                for (int n = 0; n < 8; ++n) {
                    if (need_mask) {
                        int c0 = key0 + n * 8 + 2 * t, c1 = c0 + 1;
                        bool w00 = c0 <= r0 && c0 < S && (window <= 0 || r0 - c0 < window);
                        bool w01 = c1 <= r0 && c1 < S && (window <= 0 || r0 - c1 < window);
                        bool w10 = c0 <= r1 && c0 < S && (window <= 0 || r1 - c0 < window);
                        bool w11 = c1 <= r1 && c1 < S && (window <= 0 || r1 - c1 < window);
                        if (!w00) sc[m][n][0] = -INFINITY;
                        if (!w01) sc[m][n][1] = -INFINITY;
                        if (!w10) sc[m][n][2] = -INFINITY;
                        if (!w11) sc[m][n][3] = -INFINITY;
                    }
                    mx0 = fmaxf(mx0, fmaxf(sc[m][n][0], sc[m][n][1]));
                    mx1 = fmaxf(mx1, fmaxf(sc[m][n][2], sc[m][n][3]));
                }
                mx0 = fmaxf(mx0, __shfl_xor_sync(0xffffffffu, mx0, 1)); mx0 = fmaxf(mx0, __shfl_xor_sync(0xffffffffu, mx0, 2));
                mx1 = fmaxf(mx1, __shfl_xor_sync(0xffffffffu, mx1, 1)); mx1 = fmaxf(mx1, __shfl_xor_sync(0xffffffffu, mx1, 2));
                const float nm0 = fmaxf(mrow[m][0], mx0), nm1 = fmaxf(mrow[m][1], mx1);
                const float ref0 = (nm0 == -INFINITY) ? 0.f : nm0, ref1 = (nm1 == -INFINITY) ? 0.f : nm1;
                // exp(-inf) = 0 on the first tile
                const float cor0 = PRECISE ? expf(mrow[m][0] - ref0) : exp2f(mrow[m][0] - ref0);
                const float cor1 = PRECISE ? expf(mrow[m][1] - ref1) : exp2f(mrow[m][1] - ref1);
                mrow[m][0] = nm0; mrow[m][1] = nm1;
                float rs0 = 0.f, rs1 = 0.f;
#pragma unroll
                for (int n = 0; n < 8; ++n) {
                    if (PRECISE) {
                        sc[m][n][0] = expf(sc[m][n][0] - ref0); sc[m][n][1] = expf(sc[m][n][1] - ref0);
                        sc[m][n][2] = expf(sc[m][n][2] - ref1); sc[m][n][3] = expf(sc[m][n][3] - ref1);
                    } else {
                        sc[m][n][0] = exp2f(sc[m][n][0] - ref0); sc[m][n][1] = exp2f(sc[m][n][1] - ref0);
                        sc[m][n][2] = exp2f(sc[m][n][2] - ref1); sc[m][n][3] = exp2f(sc[m][n][3] - ref1);
                    }
                    rs0 += sc[m][n][0] + sc[m][n][1];
                    rs1 += sc[m][n][2] + sc[m][n][3];
                }
                lrow[m][0] = lrow[m][0] * cor0 + rs0;   // per-thread partial row sums; quad-reduced at the end
                lrow[m][1] = lrow[m][1] * cor1 + rs1;
#pragma unroll
                for (int n = 0; n < NT; ++n) { o[m][n][0] *= cor0; o[m][n][1] *= cor0; o[m][n][2] *= cor1; o[m][n][3] *= cor1; }
                if (!PRECISE) {
                    // The C-fragment of S holds columns (2t, 2t+1) where an A-fragment wants (t, t+4): relabel the contraction
                    // index (key 2t <-> slot t, key 2t+1 <-> slot t+4) and read V rows to match.
#pragma unroll
                    for (int n = 0; n < 8; ++n) {
                        phi[m][n][0] = f2tf32(sc[m][n][0]); phi[m][n][1] = f2tf32(sc[m][n][2]);
                        phi[m][n][2] = f2tf32(sc[m][n][1]); phi[m][n][3] = f2tf32(sc[m][n][3]);
                    }
                }
            }

            // ---- O += P V
#pragma unroll
            for (int s = 0; s < 8; ++s) {           // 8 key-slabs of 8 keys
                uint32_t ph[MT][4], pl[PRECISE ? MT : 1][4];
#pragma unroll
                for (int m = 0; m < MT; ++m) {
                    if (PRECISE) {
                        split_tf32(sc[m][s][0], ph[m][0], pl[m][0]);   // (row g,   key 2t)   -> a0
                        split_tf32(sc[m][s][2], ph[m][1], pl[m][1]);   // (row g+8, key 2t)   -> a1
                        split_tf32(sc[m][s][1], ph[m][2], pl[m][2]);   // (row g,   key 2t+1) -> a2
                        split_tf32(sc[m][s][3], ph[m][3], pl[m][3]);   // (row g+8, key 2t+1) -> a3
                    } else {
                        ph[m][0] = phi[m][s][0]; ph[m][1] = phi[m][s][1]; ph[m][2] = phi[m][s][2]; ph[m][3] = phi[m][s][3];
                    }
                }
#pragma unroll
                for (int n = 0; n < NT; ++n) {
                    const float v0 = tv[(s * 8 + 2 * t) * LD + n * 8 + g];
                    const float v1 = tv[(s * 8 + 2 * t + 1) * LD + n * 8 + g];
                    if (PRECISE) {
                        uint32_t bh0, bl0, bh1, bl1;
                        split_tf32(v0, bh0, bl0); split_tf32(v1, bh1, bl1);
#pragma unroll
                        for (int m = 0; m < MT; ++m) {
                            mma_tf32(o[m][n], pl[m], bh0, bh1);
                            mma_tf32(o[m][n], ph[m], bl0, bl1);
                            mma_tf32(o[m][n], ph[m], bh0, bh1);
                        }
                    } else {
#pragma unroll
                        for (int m = 0; m < MT; ++m) mma_tf32(o[m][n], ph[m], __float_as_uint(v0), __float_as_uint(v1));
                    }
                }
            }
        }
        __syncthreads();   // everyone done with this buffer before it is refilled
    }

    float* ob = ctx + b * osb + h * osh;
    __nv_bfloat16* ob16 = ctx16 ? ctx16 + b * osb + h * osh : nullptr;
#pragma unroll
    for (int m = 0; m < MT; ++m) {
        float l0 = lrow[m][0], l1 = lrow[m][1];
        l0 += __shfl_xor_sync(0xffffffffu, l0, 1); l0 += __shfl_xor_sync(0xffffffffu, l0, 2);
        l1 += __shfl_xor_sync(0xffffffffu, l1, 1); l1 += __shfl_xor_sync(0xffffffffu, l1, 2);
        const float inv0 = l0 > 0.f ? 1.0f / l0 : 0.f, inv1 = l1 > 0.f ? 1.0f / l1 : 0.f;
        const int r0 = wrow + m * 16 + g, r1 = r0 + 8;
#pragma unroll
        for (int n = 0; n < NT; ++n) {
            if (r0 < S) *reinterpret_cast<float2*>(ob + int64_t(r0) * oss + n * 8 + 2 * t) = make_float2(o[m][n][0] * inv0, o[m][n][1] * inv0);
            if (r1 < S) *reinterpret_cast<float2*>(ob + int64_t(r1) * oss + n * 8 + 2 * t) = make_float2(o[m][n][2] * inv1, o[m][n][3] * inv1);
            if (ob16) {
                if (r0 < S) *reinterpret_cast<__nv_bfloat162*>(ob16 + int64_t(r0) * oss + n * 8 + 2 * t) = __floats2bfloat162_rn(o[m][n][0] * inv0, o[m][n][1] * inv0);
                if (r1 < S) *reinterpret_cast<__nv_bfloat162*>(ob16 + int64_t(r1) * oss + n * 8 + 2 * t) = __floats2bfloat162_rn(o[m][n][2] * inv1, o[m][n][3] * inv1);
            }
        }
    }
}

template <int HD, bool PRECISE, int MT>
static int launch_attention_mt(const float* q, const float* k, const float* v, float* ctx, int B, int H, int S,
                               int64_t qsb, int64_t qss, int64_t qsh, int64_t ksb, int64_t kss, int64_t ksh,
                               int64_t osb, int64_t oss, int64_t osh, float scale, int window, cudaStream_t st, void* ctx16) {
    constexpr size_t smem = size_t(4) * kAttBK * (HD + kAttPad) * sizeof(float);
    static bool attr_set[kMaxDevices] = {};
    if (first_use_on_device(attr_set))
        cudaFuncSetAttribute(causal_attention_fwd_kernel<HD, PRECISE, MT>, cudaFuncAttributeMaxDynamicSharedMemorySize, int(smem));
    dim3 grid(unsigned(ceil_div(S, 64 * MT)), unsigned(B * H));
    pdl_launch(causal_attention_fwd_kernel<HD, PRECISE, MT>, grid, 128, smem, st)(q, k, v, ctx, H, S, qsb, qss, qsh, ksb, kss, ksh, osb, oss,
                                                                 osh, scale, window, static_cast<__nv_bfloat16*>(ctx16));
    return 0;
}

template <int HD, bool PRECISE>
static int launch_attention(const float* q, const float* k, const float* v, float* ctx, int B, int H, int S,
                            int64_t qsb, int64_t qss, int64_t qsh, int64_t ksb, int64_t kss, int64_t ksh,
                            int64_t osb, int64_t oss, int64_t osh, float scale, int window, cudaStream_t st, void* ctx16) {
    // two 16-row slabs per warp halve the shared-memory reads per MMA; worth it once there are enough 128-row CTAs to
    // fill the chip, and only where the fragments still fit in registers (fast mode, head_dim <= 64)
    if (!PRECISE && HD <= 64 && int64_t(ceil_div(S, 128)) * B * H >= 2 * kNumSMs)
        return launch_attention_mt<HD, PRECISE, (!PRECISE && HD <= 64) ? 2 : 1>(q, k, v, ctx, B, H, S, qsb, qss, qsh, ksb, kss, ksh, osb, oss, osh,
                                                                               scale, window, st, ctx16);
    return launch_attention_mt<HD, PRECISE, 1>(q, k, v, ctx, B, H, S, qsb, qss, qsh, ksb, kss, ksh, osb, oss, osh, scale, window, st, ctx16);
}

// ---------------------------------------------------------------------------------------------------------------------
// bf16-operand variant (bf16 GEMM mode): the same flash-style loop with mma.sync m16n8k16 — q, k, v stay fp32 in HBM and in
// the shared-memory tiles, fragments are packed to bf16x2 as they are built (one cvt per pair), accumulation, scores and
// the softmax stay fp32. Half the MMAs of the TF32 kernel for the same FLOPs. For this shape the score tile's C-fragment
// (row g: keys 2t, 2t+1) is already the A-fragment layout of the P*V product, so no relabelling is needed.
// K tile pitch HD+8 floats (conflict-free 64-bit B-fragment reads), V tile pitch HD+4 (conflict-free 32-bit reads).
__device__ __forceinline__ uint32_t pack_bf16(float lo, float hi) {
    uint32_t r;
    asm("cvt.rn.bf16x2.f32 %0, %1, %2;" : "=r"(r) : "f"(hi), "f"(lo));
    return r;
}
__device__ __forceinline__ void mma_bf16(float (&c)[4], const uint32_t (&a)[4], uint32_t b0, uint32_t b1) {
    asm volatile(
        "mma.sync.aligned.m16n8k16.row.col.f32.bf16.bf16.f32 {%0,%1,%2,%3}, {%4,%5,%6,%7}, {%8,%9}, {%0,%1,%2,%3};"
        : "+f"(c[0]), "+f"(c[1]), "+f"(c[2]), "+f"(c[3])
        : "r"(a[0]), "r"(a[1]), "r"(a[2]), "r"(a[3]), "r"(b0), "r"(b1));
}

template <int HD, int MT>
__global__ void __launch_bounds__(128)
causal_attention_bf16_kernel(const float* __restrict__ q, const float* __restrict__ k, const float* __restrict__ v,
                             float* __restrict__ ctx, int H, int S,
                             int64_t qsb, int64_t qss, int64_t qsh, int64_t ksb, int64_t kss, int64_t ksh,
                             int64_t osb, int64_t oss, int64_t osh, float scale, int window,
                             __nv_bfloat16* __restrict__ ctx16) {
    pdl_sync();
    constexpr int LDK = HD + 8, LDV = HD + 4;
    constexpr int KS = HD / 16;   // k16 slabs over the head dim (QK^T)
    constexpr int NT = HD / 8;    // n-tiles over the head dim (PV)
    constexpr int BQ = 64 * MT;
    constexpr float kLog2e = 1.4426950408889634f;
    extern __shared__ __align__(16) float smem[];
    float* sK = smem;                           // [2][kAttBK][LDK]
    float* sV = smem + 2 * kAttBK * LDK;        // [2][kAttBK][LDV]

    const int qt = gridDim.x - 1 - blockIdx.x;
    const int bh = blockIdx.y;
    const int b = bh / H, h = bh % H;
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const int g = lane >> 2, t = lane & 3;
    const int q0 = qt * BQ;
    const int wrow = q0 + warp * 16 * MT;
    const float* qb = q + b * qsb + h * qsh;
    const float* kb = k + b * ksb + h * ksh;
    const float* vb = v + b * ksb + h * ksh;
    const float qscale = scale * kLog2e;        // scores in the log2 domain

    uint32_t qa[MT][KS][4];
#pragma unroll
    for (int m = 0; m < MT; ++m) {
        const int r0 = wrow + m * 16 + g, r1 = r0 + 8;
#pragma unroll
        for (int s = 0; s < KS; ++s) {
            const int c = s * 16 + 2 * t;
            const float2 z = make_float2(0.f, 0.f);
            const float2 x00 = r0 < S ? *reinterpret_cast<const float2*>(qb + int64_t(r0) * qss + c) : z;
            const float2 x10 = r1 < S ? *reinterpret_cast<const float2*>(qb + int64_t(r1) * qss + c) : z;
            const float2 x01 = r0 < S ? *reinterpret_cast<const float2*>(qb + int64_t(r0) * qss + c + 8) : z;
            const float2 x11 = r1 < S ? *reinterpret_cast<const float2*>(qb + int64_t(r1) * qss + c + 8) : z;
            qa[m][s][0] = pack_bf16(x00.x * qscale, x00.y * qscale);
            qa[m][s][1] = pack_bf16(x10.x * qscale, x10.y * qscale);
            qa[m][s][2] = pack_bf16(x01.x * qscale, x01.y * qscale);
            qa[m][s][3] = pack_bf16(x11.x * qscale, x11.y * qscale);
        }
    }
    float o[MT][NT][4];
    float mrow[MT][2], lrow[MT][2];
#pragma unroll
    for (int m = 0; m < MT; ++m) {
#pragma unroll
        for (int n = 0; n < NT; ++n) { o[m][n][0] = o[m][n][1] = o[m][n][2] = o[m][n][3] = 0.f; }
        mrow[m][0] = mrow[m][1] = -INFINITY;
        lrow[m][0] = lrow[m][1] = 0.f;
    }
    const int last_row = min(q0 + BQ, S) - 1;
    const int kt_hi = last_row / kAttBK;
    int kt_lo = 0;
    if (window > 0) { int lo = q0 - (window - 1); if (lo > 0) kt_lo = lo / kAttBK; }

    auto load_tile = [&](int kt, int buf) {
        const int key0 = kt * kAttBK;
        float* dk = sK + buf * kAttBK * LDK;
        float* dv = sV + buf * kAttBK * LDV;
        constexpr int CH = HD / 4;
        for (int c = threadIdx.x; c < kAttBK * CH; c += 128) {
            int r = c / CH, cc = c % CH;
            bool ok = (key0 + r) < S;
            int64_t off = int64_t(ok ? key0 + r : 0) * kss + cc * 4;
            cp_async16(dk + r * LDK + cc * 4, kb + off, ok);
            cp_async16(dv + r * LDV + cc * 4, vb + off, ok);
        }
        cp_async_commit();
    };
    load_tile(kt_lo, 0);
    for (int kt = kt_lo; kt <= kt_hi; ++kt) {
        const int buf = (kt - kt_lo) & 1;
        if (kt + 1 <= kt_hi) { load_tile(kt + 1, buf ^ 1); cp_async_wait<1>(); } else { cp_async_wait<0>(); }
        __syncthreads();
        const float* tk = sK + buf * kAttBK * LDK;
        const float* tv = sV + buf * kAttBK * LDV;
        const int key0 = kt * kAttBK;
        const bool active = key0 <= wrow + 16 * MT - 1;
        if (active) {
            // ---- S = (Q*scale) K^T : (16*MT) x 64 per warp
            float sc[MT][8][4];
#pragma unroll
            for (int n = 0; n < 8; ++n) {
#pragma unroll
                for (int m = 0; m < MT; ++m) sc[m][n][0] = sc[m][n][1] = sc[m][n][2] = sc[m][n][3] = 0.f;
#pragma unroll
                for (int s = 0; s < KS; ++s) {
                    const float2 k0 = *reinterpret_cast<const float2*>(tk + (n * 8 + g) * LDK + s * 16 + 2 * t);
                    const float2 k1 = *reinterpret_cast<const float2*>(tk + (n * 8 + g) * LDK + s * 16 + 2 * t + 8);
                    const uint32_t b0 = pack_bf16(k0.x, k0.y), b1 = pack_bf16(k1.x, k1.y);
#pragma unroll
                    for (int m = 0; m < MT; ++m) mma_bf16(sc[m][n], qa[m][s], b0, b1);
                }
            }
            // ---- masks + online softmax (fp32, log2 domain); P packed to bf16 A-fragments
            uint32_t pa[MT][4][4];
#pragma unroll
            for (int m = 0; m < MT; ++m) {
                const int r0 = wrow + m * 16 + g, r1 = r0 + 8;
                const bool need_mask = (key0 + kAttBK - 1 > wrow + m * 16) || (window > 0) || (key0 + kAttBK > S);
                float mx0 = -INFINITY, mx1 = -INFINITY;
#pragma unroll
                for (int n = 0; n < 8; ++n) {
                    if (need_mask) {
                        int c0 = key0 + n * 8 + 2 * t, c1 = c0 + 1;
                        bool w00 = c0 <= r0 && c0 < S && (window <= 0 || r0 - c0 < window);
                        bool w01 = c1 <= r0 && c1 < S && (window <= 0 || r0 - c1 < window);
                        bool w10 = c0 <= r1 && c0 < S && (window <= 0 || r1 - c0 < window);
                        bool w11 = c1 <= r1 && c1 < S && (window <= 0 || r1 - c1 < window);
                        if (!w00) sc[m][n][0] = -INFINITY;
                        if (!w01) sc[m][n][1] = -INFINITY;
                        if (!w10) sc[m][n][2] = -INFINITY;
                        if (!w11) sc[m][n][3] = -INFINITY;
                    }
                    mx0 = fmaxf(mx0, fmaxf(sc[m][n][0], sc[m][n][1]));
                    mx1 = fmaxf(mx1, fmaxf(sc[m][n][2], sc[m][n][3]));
                }
                mx0 = fmaxf(mx0, __shfl_xor_sync(0xffffffffu, mx0, 1)); mx0 = fmaxf(mx0, __shfl_xor_sync(0xffffffffu, mx0, 2));
                mx1 = fmaxf(mx1, __shfl_xor_sync(0xffffffffu, mx1, 1)); mx1 = fmaxf(mx1, __shfl_xor_sync(0xffffffffu, mx1, 2));
                const float nm0 = fmaxf(mrow[m][0], mx0), nm1 = fmaxf(mrow[m][1], mx1);
                const float ref0 = (nm0 == -INFINITY) ? 0.f : nm0, ref1 = (nm1 == -INFINITY) ? 0.f : nm1;
                const float cor0 = exp2f(mrow[m][0] - ref0), cor1 = exp2f(mrow[m][1] - ref1);
                mrow[m][0] = nm0; mrow[m][1] = nm1;
                float rs0 = 0.f, rs1 = 0.f;
#pragma unroll
                for (int n = 0; n < 8; ++n) {
                    sc[m][n][0] = exp2f(sc[m][n][0] - ref0); sc[m][n][1] = exp2f(sc[m][n][1] - ref0);
                    sc[m][n][2] = exp2f(sc[m][n][2] - ref1); sc[m][n][3] = exp2f(sc[m][n][3] - ref1);
                    rs0 += sc[m][n][0] + sc[m][n][1];
                    rs1 += sc[m][n][2] + sc[m][n][3];
                }
                lrow[m][0] = lrow[m][0] * cor0 + rs0;
                lrow[m][1] = lrow[m][1] * cor1 + rs1;
#pragma unroll
                for (int n = 0; n < NT; ++n) { o[m][n][0] *= cor0; o[m][n][1] *= cor0; o[m][n][2] *= cor1; o[m][n][3] *= cor1; }
#pragma unroll
                for (int s = 0; s < 4; ++s) {       // 16-key slab s = score n-tiles 2s, 2s+1
                    pa[m][s][0] = pack_bf16(sc[m][2 * s][0], sc[m][2 * s][1]);          // row g,   keys 2t, 2t+1
                    pa[m][s][1] = pack_bf16(sc[m][2 * s][2], sc[m][2 * s][3]);          // row g+8
                    pa[m][s][2] = pack_bf16(sc[m][2 * s + 1][0], sc[m][2 * s + 1][1]);  // row g,   keys 8+2t, 9+2t
                    pa[m][s][3] = pack_bf16(sc[m][2 * s + 1][2], sc[m][2 * s + 1][3]);  // row g+8
                }
            }
            // ---- O += P V
#pragma unroll
            for (int s = 0; s < 4; ++s) {
#pragma unroll
                for (int n = 0; n < NT; ++n) {
                    const float* vr = tv + (s * 16 + 2 * t) * LDV + n * 8 + g;
                    const uint32_t b0 = pack_bf16(vr[0], vr[LDV]), b1 = pack_bf16(vr[8 * LDV], vr[9 * LDV]);
#pragma unroll
                    for (int m = 0; m < MT; ++m) mma_bf16(o[m][n], pa[m][s], b0, b1);
                }
            }
        }
        __syncthreads();
    }

    float* ob = ctx + b * osb + h * osh;
    __nv_bfloat16* ob16 = ctx16 ? ctx16 + b * osb + h * osh : nullptr;
#pragma unroll
    for (int m = 0; m < MT; ++m) {
        float l0 = lrow[m][0], l1 = lrow[m][1];
        l0 += __shfl_xor_sync(0xffffffffu, l0, 1); l0 += __shfl_xor_sync(0xffffffffu, l0, 2);
        l1 += __shfl_xor_sync(0xffffffffu, l1, 1); l1 += __shfl_xor_sync(0xffffffffu, l1, 2);
        const float inv0 = l0 > 0.f ? 1.0f / l0 : 0.f, inv1 = l1 > 0.f ? 1.0f / l1 : 0.f;
        const int r0 = wrow + m * 16 + g, r1 = r0 + 8;
#pragma unroll
        for (int n = 0; n < NT; ++n) {
            if (r0 < S) *reinterpret_cast<float2*>(ob + int64_t(r0) * oss + n * 8 + 2 * t) = make_float2(o[m][n][0] * inv0, o[m][n][1] * inv0);
            if (r1 < S) *reinterpret_cast<float2*>(ob + int64_t(r1) * oss + n * 8 + 2 * t) = make_float2(o[m][n][2] * inv1, o[m][n][3] * inv1);
            if (ob16) {
                if (r0 < S) *reinterpret_cast<__nv_bfloat162*>(ob16 + int64_t(r0) * oss + n * 8 + 2 * t) = __floats2bfloat162_rn(o[m][n][0] * inv0, o[m][n][1] * inv0);
                if (r1 < S) *reinterpret_cast<__nv_bfloat162*>(ob16 + int64_t(r1) * oss + n * 8 + 2 * t) = __floats2bfloat162_rn(o[m][n][2] * inv1, o[m][n][3] * inv1);
            }
        }
    }
}

template <int HD, int MT>
static int launch_attention_bf16(const float* q, const float* k, const float* v, float* ctx, int B, int H, int S,
                                 int64_t qsb, int64_t qss, int64_t qsh, int64_t ksb, int64_t kss, int64_t ksh,
                                 int64_t osb, int64_t oss, int64_t osh, float scale, int window, cudaStream_t st, void* ctx16) {
    constexpr size_t smem = size_t(2) * kAttBK * ((HD + 8) + (HD + 4)) * sizeof(float);
    static bool attr_set[kMaxDevices] = {};
    if (first_use_on_device(attr_set))
        cudaFuncSetAttribute(causal_attention_bf16_kernel<HD, MT>, cudaFuncAttributeMaxDynamicSharedMemorySize, int(smem));
    dim3 grid(unsigned(ceil_div(S, 64 * MT)), unsigned(B * H));
    pdl_launch(causal_attention_bf16_kernel<HD, MT>, grid, 128, smem, st)(q, k, v, ctx, H, S, qsb, qss, qsh, ksb, kss, ksh, osb, oss, osh, scale,
                                                            window, static_cast<__nv_bfloat16*>(ctx16));
    return 0;
}

static bool env_attention_bf16() {
    const char* e = getenv("MFT_GEMM_PRECISION");
    return e && (e[0] == 'b' || e[0] == 'B');
}
static bool g_attention_bf16 = env_attention_bf16();     // mft_attention_bf16(): follows the bf16 GEMM operand mode

static bool env_attention_precise() {
    const char* e = getenv("MFT_USE_TENSOR_CORES");
    return e && e[0] == '0';        // exact mode: everything at fp32 accuracy
}
static bool g_attention_precise = env_attention_precise();
bool attention_precise_enabled() { return g_attention_precise; }     // shared with gqa_attention.cu / attention_tc.cu

}  // namespace mft

using namespace mft;

extern "C" int mft_attention_bf16(int enable) {
    int prev = g_attention_bf16 ? 1 : 0;
    g_attention_bf16 = enable != 0;
    return prev;
}

extern "C" int mft_attention_mode(void) { return (g_attention_precise ? 1 : 0) | (g_attention_bf16 ? 2 : 0); }

extern "C" int mft_attention_precise(int enable) {
    int prev = g_attention_precise ? 1 : 0;
    g_attention_precise = enable != 0;
    return prev;
}

static int attention_impl(const float* q, const float* k, const float* v, float* ctx, void* ctx16, int32_t B, int32_t H,
                                        int32_t S, int32_t Hd, int64_t q_stride_b, int64_t q_stride_s, int64_t q_stride_h,
                                        int64_t kv_stride_b, int64_t kv_stride_s, int64_t kv_stride_h,
                                        int64_t o_stride_b, int64_t o_stride_s, int64_t o_stride_h,
                                        float scale, int32_t window, mft_stream_t stream) {
    if (B == 0 || H == 0 || S == 0) return 0;
    if (!q || !k || !v || !ctx) return set_error(1, "mft_causal_attention_fwd: null pointer");
    if (int64_t(B) * H > 65535) return set_error(1, "mft_causal_attention_fwd: B*H too large");
    if (!(aligned16(k) && aligned16(v)) || (kv_stride_b % 4) || (kv_stride_s % 4) || (kv_stride_h % 4))
        return set_error(1, "mft_causal_attention_fwd: k/v must be 16-byte aligned with strides %% 4 == 0");
    if ((reinterpret_cast<uintptr_t>(ctx) & 7u) || (o_stride_b % 2) || (o_stride_s % 2) || (o_stride_h % 2))
        return set_error(1, "mft_causal_attention_fwd: ctx must be 8-byte aligned with even strides");
    if (ctx16 && ((reinterpret_cast<uintptr_t>(ctx16) & 3u) || (Hd & 1)))
        return set_error(1, "mft_causal_attention_fwd_ex: the bf16 output must be 4-byte aligned");
    cudaStream_t st = as_stream(stream);
    if (g_attention_bf16 && !g_attention_precise && Hd == 64 && (q_stride_s % 2) == 0 && (q_stride_b % 2) == 0 && (q_stride_h % 2) == 0 &&
        (reinterpret_cast<uintptr_t>(q) & 7u) == 0) {
        if (int64_t(ceil_div(S, 128)) * B * H >= 2 * kNumSMs)
            launch_attention_bf16<64, 2>(q, k, v, ctx, B, H, S, q_stride_b, q_stride_s, q_stride_h, kv_stride_b, kv_stride_s, kv_stride_h, o_stride_b,
                                         o_stride_s, o_stride_h, scale, window, st, ctx16);
        else
            launch_attention_bf16<64, 1>(q, k, v, ctx, B, H, S, q_stride_b, q_stride_s, q_stride_h, kv_stride_b, kv_stride_s, kv_stride_h, o_stride_b,
                                         o_stride_s, o_stride_h, scale, window, st, ctx16);
        MFT_CHECK_LAUNCH("mft_causal_attention_fwd");
        return 0;
    }
#define MFT_ATT(HD_) (g_attention_precise ? launch_attention<HD_, true>(q, k, v, ctx, B, H, S, q_stride_b, q_stride_s, q_stride_h, kv_stride_b, \
                                            kv_stride_s, kv_stride_h, o_stride_b, o_stride_s, o_stride_h, scale, window, st, ctx16) \
                                          : launch_attention<HD_, false>(q, k, v, ctx, B, H, S, q_stride_b, q_stride_s, q_stride_h, kv_stride_b, \
                                            kv_stride_s, kv_stride_h, o_stride_b, o_stride_s, o_stride_h, scale, window, st, ctx16))
    switch (Hd) {
        case 16: MFT_ATT(16); break;
        case 32: MFT_ATT(32); break;
        case 64: MFT_ATT(64); break;
        case 128: MFT_ATT(128); break;
        default: return set_error(1, "mft_causal_attention_fwd: head_dim %d not supported (16/32/64/128)", Hd);
    }
#undef MFT_ATT
    MFT_CHECK_LAUNCH("mft_causal_attention_fwd");
    return 0;
}

extern "C" int mft_causal_attention_fwd(const float* q, const float* k, const float* v, float* ctx, int32_t B, int32_t H,
                                        int32_t S, int32_t Hd, int64_t q_stride_b, int64_t q_stride_s, int64_t q_stride_h,
                                        int64_t kv_stride_b, int64_t kv_stride_s, int64_t kv_stride_h,
                                        int64_t o_stride_b, int64_t o_stride_s, int64_t o_stride_h,
                                        float scale, int32_t window, mft_stream_t stream) {
    return attention_impl(q, k, v, ctx, nullptr, B, H, S, Hd, q_stride_b, q_stride_s, q_stride_h, kv_stride_b, kv_stride_s, kv_stride_h,
                          o_stride_b, o_stride_s, o_stride_h, scale, window, stream);
}

// same, plus a bf16 copy of ctx (same element strides as ctx) for the bf16 GEMM of the output projection
extern "C" int mft_causal_attention_fwd_ex(const float* q, const float* k, const float* v, float* ctx, void* ctx_bf16, int32_t B, int32_t H,
                                           int32_t S, int32_t Hd, int64_t q_stride_b, int64_t q_stride_s, int64_t q_stride_h,
                                           int64_t kv_stride_b, int64_t kv_stride_s, int64_t kv_stride_h,
                                           int64_t o_stride_b, int64_t o_stride_s, int64_t o_stride_h,
                                           float scale, int32_t window, mft_stream_t stream) {
    return attention_impl(q, k, v, ctx, ctx_bf16, B, H, S, Hd, q_stride_b, q_stride_s, q_stride_h, kv_stride_b, kv_stride_s, kv_stride_h,
                          o_stride_b, o_stride_s, o_stride_h, scale, window, stream);
}
