// Grouped-query causal / sliding-window attention, forward AND backward, without the [B,H,S,S] score tensor in HBM.
//
// Replaces the reference's Gemma attention chain (graph/gemma_model.cpp:383-514):
//   repeat_kv_heads (core/ops.cpp:2072-2149) -> Q K^T -> * query_pre_attn_scalar^-1/2 -> + mask [S,S] (+ pad [B,1,1,S])
//   -> softmax (core/ops.cpp:1119-1175) -> P V, and in backward MatmulBackward, SoftmaxBackward
//   (core/backward_functions.cpp:297-340), ScaleBackward, RepeatKVHeadsBackward (:683-716: sum over the group).
// The same kernels give GPT-2 a real attention backward (KV == H, head_dim 64) for the true-gradient mode.
//
// Layout: q [B,S,H,Hd], k/v [B,S,KV,Hd] addressed by (batch, seq, head) strides, so projection outputs are consumed in
// place and ctx is written as [B,S,H*Hd]; query head h reads kv head h / (H/KV) — nothing is repeated in memory.
// Forward: one CTA = BQ query rows of one (b, h); keys stream through shared memory in tiles of 32; scores, softmax
// statistics and P live in shared memory / registers only; per-row log-sum-exp is saved for the backward.
// Backward (deterministic, no atomics): kernel 1 (CTA per query tile) recomputes P from the saved statistics and
// produces dQ and D = rowsum(dO * O); kernel 2 (CTA per key tile and query head) produces that head's dK / dV; with
// grouped heads a third kernel sums the group's partial dK / dV in a fixed order.
// Products run on the tensor pipe as mma.sync m16n8k8 TF32; the exact mode (mft_attention_precise) uses the 3xTF32
// split on every product so results stay at fp32 accuracy (parity tests: 1e-5 against the fp64 oracle).
#include <cstdlib>
#include "common.cuh"

namespace mft {

constexpr int kGBK = 32;            // keys per tile (and query rows per tile in the backward kernels)
constexpr float kGMask = -1e10f;    // the reference's finite mask value (graph/gemma_model.cpp)

struct GqaArgs {
    const float *q, *k, *v, *key_bias;     // key_bias: optional additive [B,S] (0 / -1e10 padding mask)
    float* o; float* lse; __nv_bfloat16* o16;
    const float *dout, *out;               // backward: dO and O (O strides)
    float *dq, *dk, *dv, *dsum;            // dk/dv: final buffers (kv strides) or per-query-head partials [B,H,S,Hd]
    int B, H, KV, S;
    int64_t qsb, qss, qsh, ksb, kss, ksh, osb, oss, osh;
    int64_t dqsb, dqss, dqsh, dksb, dkss, dksh;
    float scale; int window;
    int partial;                           // dkv kernel: 1 = write per-head partials (dense [B,H,S,Hd])
};

// acc[NT][4] += A[16 x K] * B^T with A rows and B rows both K-contiguous in shared memory (B row n = output column n)
template <int NT, bool PRECISE>
__device__ __forceinline__ void warp_mma_nt(float (&acc)[NT][4], const float* __restrict__ A, int lda, const float* __restrict__ Bm,
                                            int ldb, int K, int lane) {
    const int g = lane >> 2, t = lane & 3;
    for (int k0 = 0; k0 < K; k0 += 8) {
        const float a[4] = {A[g * lda + k0 + t], A[(g + 8) * lda + k0 + t], A[g * lda + k0 + t + 4], A[(g + 8) * lda + k0 + t + 4]};
        uint32_t ah[4], al[4];
#pragma unroll
        for (int i = 0; i < 4; ++i) { if (PRECISE) split_tf32(a[i], ah[i], al[i]); else { ah[i] = f2tf32(a[i]); al[i] = 0; } }
#pragma unroll
        for (int n = 0; n < NT; ++n) {
            const float b0 = Bm[(n * 8 + g) * ldb + k0 + t], b1 = Bm[(n * 8 + g) * ldb + k0 + t + 4];
            if (PRECISE) {
                uint32_t bh0, bl0, bh1, bl1;
                split_tf32(b0, bh0, bl0); split_tf32(b1, bh1, bl1);
                mma_tf32(acc[n], al, bh0, bh1); mma_tf32(acc[n], ah, bl0, bl1); mma_tf32(acc[n], ah, bh0, bh1);
            } else mma_tf32(acc[n], ah, f2tf32(b0), f2tf32(b1));
        }
    }
}
// acc[NT][4] += A[16 x K] * B[K x 8*NT] with B stored row-major [k][n]
template <int NT, bool PRECISE>
__device__ __forceinline__ void warp_mma_nn(float (&acc)[NT][4], const float* __restrict__ A, int lda, const float* __restrict__ Bm,
                                            int ldb, int K, int lane) {
    const int g = lane >> 2, t = lane & 3;
    for (int k0 = 0; k0 < K; k0 += 8) {
        const float a[4] = {A[g * lda + k0 + t], A[(g + 8) * lda + k0 + t], A[g * lda + k0 + t + 4], A[(g + 8) * lda + k0 + t + 4]};
        uint32_t ah[4], al[4];
#pragma unroll
        for (int i = 0; i < 4; ++i) { if (PRECISE) split_tf32(a[i], ah[i], al[i]); else { ah[i] = f2tf32(a[i]); al[i] = 0; } }
#pragma unroll
        for (int n = 0; n < NT; ++n) {
            const float b0 = Bm[(k0 + t) * ldb + n * 8 + g], b1 = Bm[(k0 + t + 4) * ldb + n * 8 + g];
            if (PRECISE) {
                uint32_t bh0, bl0, bh1, bl1;
                split_tf32(b0, bh0, bl0); split_tf32(b1, bh1, bl1);
                mma_tf32(acc[n], al, bh0, bh1); mma_tf32(acc[n], ah, bl0, bl1); mma_tf32(acc[n], ah, bh0, bh1);
            } else mma_tf32(acc[n], ah, f2tf32(b0), f2tf32(b1));
        }
    }
}

// rows [row0, row0 + nrows) of a [S, HD] strided matrix -> shared memory (pitch LD floats); rows >= S are zero-filled
template <int HD, int THREADS>
__device__ __forceinline__ void load_rows(float* __restrict__ dst, int ld, const float* __restrict__ src, int64_t row_stride, int row0,
                                          int nrows, int S) {
    constexpr int CH = HD / 4;
    for (int c = threadIdx.x; c < nrows * CH; c += THREADS) {
        const int r = c / CH, cc = c % CH;
        float4 v = make_float4(0.f, 0.f, 0.f, 0.f);
        if (row0 + r < S) v = *reinterpret_cast<const float4*>(src + int64_t(row0 + r) * row_stride + cc * 4);
        *reinterpret_cast<float4*>(dst + r * ld + cc * 4) = v;
    }
}

// score of (query i, key j) as the reference builds it: dot * scale, + mask value where the causal / window mask forbids
// the pair, + the padding bias of key j
__device__ __forceinline__ float masked_score(float dot, float scale, int i, int j, int window, const float* __restrict__ kbias) {
    float s = dot * scale;
    if (j > i || (window > 0 && i - j >= window)) s += kGMask;
    if (kbias) s += kbias[j];
    return s;
}

template <int HD, int BQ, int NW, bool PRECISE>
__global__ void __launch_bounds__(NW * 32)
gqa_fwd_kernel(GqaArgs p) {
    pdl_sync();
    constexpr int LD = HD + 4, LDP = kGBK + 4;
    constexpr int RG = BQ / 16, CG = NW / RG, NT_S = (kGBK / 8) / CG, NT_O = HD / (8 * CG);
    constexpr int THREADS = NW * 32, TPR = THREADS / BQ, EPT = kGBK / TPR;
    static_assert(NT_S >= 1 && NT_O >= 1 && EPT >= 1, "tile split");
    extern __shared__ __align__(16) float sm[];
    float* sQ = sm;
    float* sK = sQ + BQ * LD;
    float* sV = sK + kGBK * LD;
    float* sP = sV + kGBK * LD;
    float* sM = sP + BQ * LDP;
    float* sL = sM + BQ;
    float* sA = sL + BQ;

    const int qt = gridDim.x - 1 - blockIdx.x;       // heavy (late) query tiles first
    const int b = blockIdx.y / p.H, h = blockIdx.y % p.H, kvh = h / (p.H / p.KV);
    const int S = p.S, q0 = qt * BQ;
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31, g = lane >> 2, t = lane & 3;
    const int rg = warp % RG, cg = warp / RG;
    const float* qb = p.q + b * p.qsb + h * p.qsh;
    const float* kb = p.k + b * p.ksb + kvh * p.ksh;
    const float* vb = p.v + b * p.ksb + kvh * p.ksh;
    const float* kbias = p.key_bias ? p.key_bias + int64_t(b) * S : nullptr;

    load_rows<HD, THREADS>(sQ, LD, qb, p.qss, q0, BQ, S);
    if (threadIdx.x < BQ) { sM[threadIdx.x] = -INFINITY; sL[threadIdx.x] = 0.f; }
    float o[NT_O][4];
#pragma unroll
    for (int n = 0; n < NT_O; ++n) o[n][0] = o[n][1] = o[n][2] = o[n][3] = 0.f;

    const int last_row = min(q0 + BQ, S) - 1;
    const int kt_hi = last_row / kGBK;
    int kt_lo = 0;
    if (p.window > 0) { const int lo = q0 - (p.window - 1); if (lo > 0) kt_lo = lo / kGBK; }

    for (int kt = kt_lo; kt <= kt_hi; ++kt) {
        const int key0 = kt * kGBK;
        __syncthreads();                              // the previous tile's P*V has read sP / sV
        load_rows<HD, THREADS>(sK, LD, kb, p.kss, key0, kGBK, S);
        load_rows<HD, THREADS>(sV, LD, vb, p.kss, key0, kGBK, S);
        __syncthreads();
        {   // ---- scores of this warp's 16 x (8*NT_S) patch
            float sc[NT_S][4];
#pragma unroll
            for (int n = 0; n < NT_S; ++n) sc[n][0] = sc[n][1] = sc[n][2] = sc[n][3] = 0.f;
            warp_mma_nt<NT_S, PRECISE>(sc, sQ + rg * 16 * LD, LD, sK + cg * NT_S * 8 * LD, LD, HD, lane);
#pragma unroll
            for (int n = 0; n < NT_S; ++n)
#pragma unroll
                for (int e = 0; e < 4; ++e) {
                    const int row = rg * 16 + g + (e >= 2 ? 8 : 0), col = cg * NT_S * 8 + n * 8 + 2 * t + (e & 1);
                    const int i = q0 + row, j = key0 + col;
                    sP[row * LDP + col] = j < S ? masked_score(sc[n][e], p.scale, i, j, p.window, kbias) : -INFINITY;
                }
        }
        __syncthreads();
        {   // ---- online softmax, TPR threads per row
            const int row = threadIdx.x / TPR, sub = threadIdx.x % TPR;
            float vals[EPT], mx = -INFINITY;
            const float m_old = sM[row], l_old = sL[row];
#pragma unroll
            for (int e = 0; e < EPT; ++e) { vals[e] = sP[row * LDP + sub * EPT + e]; mx = fmaxf(mx, vals[e]); }
#pragma unroll
            for (int off = TPR / 2; off > 0; off >>= 1) mx = fmaxf(mx, __shfl_xor_sync(0xffffffffu, mx, off));
            const float m_new = fmaxf(m_old, mx);
            const float alpha = expf(m_old - m_new);  // exp(-inf) = 0 on the first tile
            float sum = 0.f;
#pragma unroll
            for (int e = 0; e < EPT; ++e) { const float pv = expf(vals[e] - m_new); sP[row * LDP + sub * EPT + e] = pv; sum += pv; }
#pragma unroll
            for (int off = TPR / 2; off > 0; off >>= 1) sum += __shfl_xor_sync(0xffffffffu, sum, off);
            if (sub == 0) { sM[row] = m_new; sL[row] = l_old * alpha + sum; sA[row] = alpha; }
        }
        __syncthreads();
        {   // ---- O = O * alpha + P V
            const float a0 = sA[rg * 16 + g], a1 = sA[rg * 16 + g + 8];
#pragma unroll
            for (int n = 0; n < NT_O; ++n) { o[n][0] *= a0; o[n][1] *= a0; o[n][2] *= a1; o[n][3] *= a1; }
            warp_mma_nn<NT_O, PRECISE>(o, sP + rg * 16 * LDP, LDP, sV + cg * NT_O * 8, LD, kGBK, lane);
        }
    }
    __syncthreads();
    const int lr0 = rg * 16 + g, lr1 = lr0 + 8, r0 = q0 + lr0, r1 = q0 + lr1;
    const float l0 = sL[lr0], l1 = sL[lr1];
    const float inv0 = l0 > 0.f ? 1.0f / l0 : 0.f, inv1 = l1 > 0.f ? 1.0f / l1 : 0.f;
    float* ob = p.o + b * p.osb + h * p.osh;
    __nv_bfloat16* ob16 = p.o16 ? p.o16 + b * p.osb + h * p.osh : nullptr;
#pragma unroll
    for (int n = 0; n < NT_O; ++n) {
        const int col = cg * NT_O * 8 + n * 8 + 2 * t;
        if (r0 < S) {
            *reinterpret_cast<float2*>(ob + int64_t(r0) * p.oss + col) = make_float2(o[n][0] * inv0, o[n][1] * inv0);
            if (ob16) *reinterpret_cast<__nv_bfloat162*>(ob16 + int64_t(r0) * p.oss + col) = __floats2bfloat162_rn(o[n][0] * inv0, o[n][1] * inv0);
        }
        if (r1 < S) {
            *reinterpret_cast<float2*>(ob + int64_t(r1) * p.oss + col) = make_float2(o[n][2] * inv1, o[n][3] * inv1);
            if (ob16) *reinterpret_cast<__nv_bfloat162*>(ob16 + int64_t(r1) * p.oss + col) = __floats2bfloat162_rn(o[n][2] * inv1, o[n][3] * inv1);
        }
    }
    if (p.lse && cg == 0 && t == 0) {
        float* lp = p.lse + (int64_t(b) * p.H + h) * S;
        if (r0 < S) lp[r0] = sM[lr0] + logf(l0);
        if (r1 < S) lp[r1] = sM[lr1] + logf(l1);
    }
}

// ---- backward, kernel 1: dQ and D = rowsum(dO * O) for one tile of 32 queries
template <int HD, int NW, bool PRECISE>
__global__ void __launch_bounds__(NW * 32)
gqa_bwd_dq_kernel(GqaArgs p) {
    pdl_sync();
    constexpr int BQ = kGBK, LD = HD + 4, LDP = kGBK + 4;
    constexpr int RG = BQ / 16, CG = NW / RG, NT_S = (kGBK / 8) / CG, NT_O = HD / (8 * CG);
    constexpr int THREADS = NW * 32, TPR = THREADS / BQ;
    static_assert(NT_S >= 1 && NT_O >= 1, "tile split");
    extern __shared__ __align__(16) float sm[];
    float* sQ = sm;
    float* sdO = sQ + BQ * LD;
    float* sK = sdO + BQ * LD;
    float* sV = sK + kGBK * LD;
    float* sP = sV + kGBK * LD;
    float* sLse = sP + BQ * LDP;
    float* sD = sLse + BQ;

    const int qt = gridDim.x - 1 - blockIdx.x;
    const int b = blockIdx.y / p.H, h = blockIdx.y % p.H, kvh = h / (p.H / p.KV);
    const int S = p.S, q0 = qt * BQ;
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31, g = lane >> 2, t = lane & 3;
    const int rg = warp % RG, cg = warp / RG;
    const float* qb = p.q + b * p.qsb + h * p.qsh;
    const float* kb = p.k + b * p.ksb + kvh * p.ksh;
    const float* vb = p.v + b * p.ksb + kvh * p.ksh;
    const float* dob = p.dout + b * p.osb + h * p.osh;
    const float* ob = p.out + b * p.osb + h * p.osh;
    const float* kbias = p.key_bias ? p.key_bias + int64_t(b) * S : nullptr;
    const int64_t stat = (int64_t(b) * p.H + h) * S;

    load_rows<HD, THREADS>(sQ, LD, qb, p.qss, q0, BQ, S);
    load_rows<HD, THREADS>(sdO, LD, dob, p.oss, q0, BQ, S);
    load_rows<HD, THREADS>(sK, LD, ob, p.oss, q0, BQ, S);        // O, only for D
    __syncthreads();
    {
        const int row = threadIdx.x / TPR, sub = threadIdx.x % TPR;
        float acc = 0.f;
        for (int c = sub; c < HD; c += TPR) acc += sdO[row * LD + c] * sK[row * LD + c];
#pragma unroll
        for (int off = TPR / 2; off > 0; off >>= 1) acc += __shfl_xor_sync(0xffffffffu, acc, off);
        if (sub == 0) {
            sD[row] = acc;
            const bool in = q0 + row < S;
            sLse[row] = in ? p.lse[stat + q0 + row] : 0.f;
            if (in) p.dsum[stat + q0 + row] = acc;
        }
    }
    float dq[NT_O][4];
#pragma unroll
    for (int n = 0; n < NT_O; ++n) dq[n][0] = dq[n][1] = dq[n][2] = dq[n][3] = 0.f;
    const int last_row = min(q0 + BQ, S) - 1;
    const int kt_hi = last_row / kGBK;
    int kt_lo = 0;
    if (p.window > 0) { const int lo = q0 - (p.window - 1); if (lo > 0) kt_lo = lo / kGBK; }
    for (int kt = kt_lo; kt <= kt_hi; ++kt) {
        const int key0 = kt * kGBK;
        __syncthreads();
        load_rows<HD, THREADS>(sK, LD, kb, p.kss, key0, kGBK, S);
        load_rows<HD, THREADS>(sV, LD, vb, p.kss, key0, kGBK, S);
        __syncthreads();
        {
            float sc[NT_S][4], dp[NT_S][4];
#pragma unroll
            for (int n = 0; n < NT_S; ++n) { sc[n][0] = sc[n][1] = sc[n][2] = sc[n][3] = 0.f; dp[n][0] = dp[n][1] = dp[n][2] = dp[n][3] = 0.f; }
            warp_mma_nt<NT_S, PRECISE>(sc, sQ + rg * 16 * LD, LD, sK + cg * NT_S * 8 * LD, LD, HD, lane);
            warp_mma_nt<NT_S, PRECISE>(dp, sdO + rg * 16 * LD, LD, sV + cg * NT_S * 8 * LD, LD, HD, lane);
#pragma unroll
            for (int n = 0; n < NT_S; ++n)
#pragma unroll
                for (int e = 0; e < 4; ++e) {
                    const int row = rg * 16 + g + (e >= 2 ? 8 : 0), col = cg * NT_S * 8 + n * 8 + 2 * t + (e & 1);
                    const int i = q0 + row, j = key0 + col;
                    float ds = 0.f;
                    if (i < S && j < S) {
                        const float pr = expf(masked_score(sc[n][e], p.scale, i, j, p.window, kbias) - sLse[row]);
                        ds = pr * (dp[n][e] - sD[row]);
                    }
                    sP[row * LDP + col] = ds;
                }
        }
        __syncthreads();
        warp_mma_nn<NT_O, PRECISE>(dq, sP + rg * 16 * LDP, LDP, sK + cg * NT_O * 8, LD, kGBK, lane);
    }
    float* dqb = p.dq + b * p.dqsb + h * p.dqsh;
    const int r0 = q0 + rg * 16 + g, r1 = r0 + 8;
#pragma unroll
    for (int n = 0; n < NT_O; ++n) {
        const int col = cg * NT_O * 8 + n * 8 + 2 * t;
        if (r0 < S) *reinterpret_cast<float2*>(dqb + int64_t(r0) * p.dqss + col) = make_float2(dq[n][0] * p.scale, dq[n][1] * p.scale);
        if (r1 < S) *reinterpret_cast<float2*>(dqb + int64_t(r1) * p.dqss + col) = make_float2(dq[n][2] * p.scale, dq[n][3] * p.scale);
    }
}

// ---- backward, kernel 2: dK and dV of one tile of 32 keys, contributions of ONE query head
template <int HD, int NW, bool PRECISE>
__global__ void __launch_bounds__(NW * 32)
gqa_bwd_dkv_kernel(GqaArgs p) {
    pdl_sync();
    constexpr int LD = HD + 4, LDP = kGBK + 4;
    constexpr int RG = kGBK / 16, CG = NW / RG, NT_S = (kGBK / 8) / CG, NT_O = HD / (8 * CG);
    constexpr int THREADS = NW * 32;
    extern __shared__ __align__(16) float sm[];
    float* sK = sm;
    float* sV = sK + kGBK * LD;
    float* sQ = sV + kGBK * LD;
    float* sdO = sQ + kGBK * LD;
    float* sPt = sdO + kGBK * LD;
    float* sdSt = sPt + kGBK * LDP;
    float* sLse = sdSt + kGBK * LDP;
    float* sD = sLse + kGBK;

    const int kt = blockIdx.x;
    const int b = blockIdx.y / p.H, h = blockIdx.y % p.H, kvh = h / (p.H / p.KV);
    const int S = p.S, key0 = kt * kGBK;
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31, g = lane >> 2, t = lane & 3;
    const int rg = warp % RG, cg = warp / RG;
    const float* qb = p.q + b * p.qsb + h * p.qsh;
    const float* kb = p.k + b * p.ksb + kvh * p.ksh;
    const float* vb = p.v + b * p.ksb + kvh * p.ksh;
    const float* dob = p.dout + b * p.osb + h * p.osh;
    const float* kbias = p.key_bias ? p.key_bias + int64_t(b) * S : nullptr;
    const int64_t stat = (int64_t(b) * p.H + h) * S;

    load_rows<HD, THREADS>(sK, LD, kb, p.kss, key0, kGBK, S);
    load_rows<HD, THREADS>(sV, LD, vb, p.kss, key0, kGBK, S);
    float dk[NT_O][4], dv[NT_O][4];
#pragma unroll
    for (int n = 0; n < NT_O; ++n) { dk[n][0] = dk[n][1] = dk[n][2] = dk[n][3] = 0.f; dv[n][0] = dv[n][1] = dv[n][2] = dv[n][3] = 0.f; }
    const int qt_lo = kt;                                  // first query tile that can see these keys (causal)
    int qt_hi = (S - 1) / kGBK;
    if (p.window > 0) qt_hi = min(qt_hi, (key0 + kGBK - 1 + p.window - 1) / kGBK);
    for (int qt = qt_lo; qt <= qt_hi; ++qt) {
        const int q0 = qt * kGBK;
        __syncthreads();
        load_rows<HD, THREADS>(sQ, LD, qb, p.qss, q0, kGBK, S);
        load_rows<HD, THREADS>(sdO, LD, dob, p.oss, q0, kGBK, S);
        if (threadIdx.x < kGBK) {
            const bool in = q0 + int(threadIdx.x) < S;
            sLse[threadIdx.x] = in ? p.lse[stat + q0 + threadIdx.x] : 0.f;
            sD[threadIdx.x] = in ? p.dsum[stat + q0 + threadIdx.x] : 0.f;
        }
        __syncthreads();
        {   // transposed patches: rows = keys, columns = queries
            float st[NT_S][4], dpt[NT_S][4];
#pragma unroll
            for (int n = 0; n < NT_S; ++n) { st[n][0] = st[n][1] = st[n][2] = st[n][3] = 0.f; dpt[n][0] = dpt[n][1] = dpt[n][2] = dpt[n][3] = 0.f; }
            warp_mma_nt<NT_S, PRECISE>(st, sK + rg * 16 * LD, LD, sQ + cg * NT_S * 8 * LD, LD, HD, lane);
            warp_mma_nt<NT_S, PRECISE>(dpt, sV + rg * 16 * LD, LD, sdO + cg * NT_S * 8 * LD, LD, HD, lane);
#pragma unroll
            for (int n = 0; n < NT_S; ++n)
#pragma unroll
                for (int e = 0; e < 4; ++e) {
                    const int krow = rg * 16 + g + (e >= 2 ? 8 : 0), qcol = cg * NT_S * 8 + n * 8 + 2 * t + (e & 1);
                    const int j = key0 + krow, i = q0 + qcol;
                    float pr = 0.f, ds = 0.f;
                    if (i < S && j < S) {
                        pr = expf(masked_score(st[n][e], p.scale, i, j, p.window, kbias) - sLse[qcol]);
                        ds = pr * (dpt[n][e] - sD[qcol]);
                    }
                    sPt[krow * LDP + qcol] = pr;
                    sdSt[krow * LDP + qcol] = ds;
                }
        }
        __syncthreads();
        warp_mma_nn<NT_O, PRECISE>(dv, sPt + rg * 16 * LDP, LDP, sdO + cg * NT_O * 8, LD, kGBK, lane);
        warp_mma_nn<NT_O, PRECISE>(dk, sdSt + rg * 16 * LDP, LDP, sQ + cg * NT_O * 8, LD, kGBK, lane);
    }
    float *dkb, *dvb;
    int64_t rs;
    if (p.partial) {
        dkb = p.dk + (int64_t(b) * p.H + h) * S * HD; dvb = p.dv + (int64_t(b) * p.H + h) * S * HD; rs = HD;
    } else {
        dkb = p.dk + b * p.dksb + kvh * p.dksh; dvb = p.dv + b * p.dksb + kvh * p.dksh; rs = p.dkss;
    }
    const int r0 = key0 + rg * 16 + g, r1 = r0 + 8;
#pragma unroll
    for (int n = 0; n < NT_O; ++n) {
        const int col = cg * NT_O * 8 + n * 8 + 2 * t;
        if (r0 < S) {
            *reinterpret_cast<float2*>(dkb + int64_t(r0) * rs + col) = make_float2(dk[n][0] * p.scale, dk[n][1] * p.scale);
            *reinterpret_cast<float2*>(dvb + int64_t(r0) * rs + col) = make_float2(dv[n][0], dv[n][1]);
        }
        if (r1 < S) {
            *reinterpret_cast<float2*>(dkb + int64_t(r1) * rs + col) = make_float2(dk[n][2] * p.scale, dk[n][3] * p.scale);
            *reinterpret_cast<float2*>(dvb + int64_t(r1) * rs + col) = make_float2(dv[n][2], dv[n][3]);
        }
    }
}

// ---- backward, kernel 3 (grouped heads only): dK / dV of a kv head = sum of its query heads' partials, in head order
// (RepeatKVHeadsBackward, core/backward_functions.cpp:683-716)
__global__ void gqa_group_sum_kernel(const float* __restrict__ pk, const float* __restrict__ pv, float* __restrict__ dk, float* __restrict__ dv,
                                     int B, int H, int KV, int S, int HD, int64_t dksb, int64_t dkss, int64_t dksh) {
    pdl_sync();
    const int rep = H / KV, c4 = HD / 4;
    const int64_t n = int64_t(B) * KV * S * c4;
    for (int64_t idx = blockIdx.x * int64_t(blockDim.x) + threadIdx.x; idx < n; idx += int64_t(gridDim.x) * blockDim.x) {
        const int c = int(idx % c4);
        const int s = int((idx / c4) % S);
        const int kvh = int((idx / (int64_t(c4) * S)) % KV);
        const int b = int(idx / (int64_t(c4) * S * KV));
        float4 ak = make_float4(0.f, 0.f, 0.f, 0.f), av = ak;
        for (int r = 0; r < rep; ++r) {
            const int64_t off = ((int64_t(b) * H + kvh * rep + r) * S + s) * HD + c * 4;
            const float4 a = *reinterpret_cast<const float4*>(pk + off), bb = *reinterpret_cast<const float4*>(pv + off);
            ak.x += a.x; ak.y += a.y; ak.z += a.z; ak.w += a.w;
            av.x += bb.x; av.y += bb.y; av.z += bb.z; av.w += bb.w;
        }
        const int64_t o = b * dksb + int64_t(s) * dkss + kvh * dksh + c * 4;
        *reinterpret_cast<float4*>(dk + o) = ak;
        *reinterpret_cast<float4*>(dv + o) = av;
    }
}

bool attention_precise_enabled();      // attention.cu (mft_attention_precise state)

template <int HD, bool PRECISE>
static int launch_gqa_fwd(const GqaArgs& a, cudaStream_t st) {
    constexpr int BQ = 64, NW = 8;
    constexpr size_t smem = (size_t(BQ) * (HD + 4) + 2 * size_t(kGBK) * (HD + 4) + size_t(BQ) * (kGBK + 4) + 3 * BQ) * sizeof(float);
    static bool attr_set[kMaxDevices] = {};
    if (first_use_on_device(attr_set))
        cudaFuncSetAttribute(gqa_fwd_kernel<HD, BQ, NW, PRECISE>, cudaFuncAttributeMaxDynamicSharedMemorySize, int(smem));
    dim3 grid(unsigned(ceil_div(a.S, BQ)), unsigned(a.B * a.H));
    pdl_launch(gqa_fwd_kernel<HD, BQ, NW, PRECISE>, grid, NW * 32, smem, st)(a);
    return 0;
}
template <int HD, bool PRECISE>
static int launch_gqa_bwd(GqaArgs a, float* part_k, float* part_v, cudaStream_t st) {
    constexpr int NW = HD >= 32 ? 8 : 4;
    constexpr size_t smem_dq = (4 * size_t(kGBK) * (HD + 4) + size_t(kGBK) * (kGBK + 4) + 2 * kGBK) * sizeof(float);
    constexpr size_t smem_kv = (4 * size_t(kGBK) * (HD + 4) + 2 * size_t(kGBK) * (kGBK + 4) + 2 * kGBK) * sizeof(float);
    static bool attr_set[kMaxDevices] = {};
    if (first_use_on_device(attr_set)) {
        cudaFuncSetAttribute(gqa_bwd_dq_kernel<HD, NW, PRECISE>, cudaFuncAttributeMaxDynamicSharedMemorySize, int(smem_dq));
        cudaFuncSetAttribute(gqa_bwd_dkv_kernel<HD, NW, PRECISE>, cudaFuncAttributeMaxDynamicSharedMemorySize, int(smem_kv));
    }
    dim3 grid(unsigned(ceil_div(a.S, kGBK)), unsigned(a.B * a.H));
    pdl_launch(gqa_bwd_dq_kernel<HD, NW, PRECISE>, grid, NW * 32, smem_dq, st)(a);
    float* dk_final = a.dk;
    float* dv_final = a.dv;
    if (a.H != a.KV) { a.partial = 1; a.dk = part_k; a.dv = part_v; }
    pdl_launch(gqa_bwd_dkv_kernel<HD, NW, PRECISE>, grid, NW * 32, smem_kv, st)(a);
    add_launches(1);
    if (a.H != a.KV) {
        const int64_t n = int64_t(a.B) * a.KV * a.S * (HD / 4);
        pdl_launch(gqa_group_sum_kernel, stream_grid(n, 256), 256, 0, st)(part_k, part_v, dk_final, dv_final, a.B, a.H, a.KV, a.S, HD, a.dksb, a.dkss, a.dksh);
        add_launches(1);
    }
    return 0;
}

}  // namespace mft

using namespace mft;

static int gqa_check_common(const char* who, const void* q, const void* k, const void* v, int32_t B, int32_t H, int32_t KV, int32_t S, int32_t Hd,
                            const int64_t* strides, int nstrides) {
    if (!q || !k || !v) return set_error(1, "%s: null pointer", who);
    if (KV < 1 || H % KV) return set_error(1, "%s: H (%d) must be a multiple of KV (%d)", who, H, KV);
    if (int64_t(B) * H > 65535) return set_error(1, "%s: B*H too large", who);
    if (Hd != 16 && Hd != 32 && Hd != 64 && Hd != 128 && Hd != 256) return set_error(1, "%s: head_dim %d not supported (16/32/64/128/256)", who, Hd);
    if (!aligned16(q) || !aligned16(k) || !aligned16(v)) return set_error(1, "%s: q/k/v must be 16-byte aligned", who);
    for (int i = 0; i < nstrides; ++i) if (strides[i] % 4) return set_error(1, "%s: strides must be multiples of 4 elements", who);
    (void)S;
    return 0;
}

#define MFT_GQA_DISPATCH(CALL)                                              \
    switch (Hd) {                                                           \
        case 16: if (precise) { CALL(16, true); } else { CALL(16, false); } break;    \
        case 32: if (precise) { CALL(32, true); } else { CALL(32, false); } break;    \
        case 64: if (precise) { CALL(64, true); } else { CALL(64, false); } break;    \
        case 128: if (precise) { CALL(128, true); } else { CALL(128, false); } break; \
        default: if (precise) { CALL(256, true); } else { CALL(256, false); } break;  \
    }

extern "C" int mft_gqa_attention_fwd(const float* q, const float* k, const float* v, const float* key_bias, float* ctx, float* lse, void* ctx_bf16,
                                     int32_t B, int32_t H, int32_t KV, int32_t S, int32_t Hd,
                                     int64_t q_stride_b, int64_t q_stride_s, int64_t q_stride_h,
                                     int64_t kv_stride_b, int64_t kv_stride_s, int64_t kv_stride_h,
                                     int64_t o_stride_b, int64_t o_stride_s, int64_t o_stride_h,
                                     float scale, int32_t window, mft_stream_t stream) {
    if (B == 0 || H == 0 || S == 0) return 0;
    const int64_t strides[9] = {q_stride_b, q_stride_s, q_stride_h, kv_stride_b, kv_stride_s, kv_stride_h, o_stride_b, o_stride_s, o_stride_h};
    if (int rc = gqa_check_common("mft_gqa_attention_fwd", q, k, v, B, H, KV, S, Hd, strides, 9)) return rc;
    if (!ctx || !aligned16(ctx)) return set_error(1, "mft_gqa_attention_fwd: ctx must be a 16-byte aligned pointer");
    GqaArgs a{};
    a.q = q; a.k = k; a.v = v; a.key_bias = key_bias; a.o = ctx; a.lse = lse; a.o16 = static_cast<__nv_bfloat16*>(ctx_bf16);
    a.B = B; a.H = H; a.KV = KV; a.S = S;
    a.qsb = q_stride_b; a.qss = q_stride_s; a.qsh = q_stride_h; a.ksb = kv_stride_b; a.kss = kv_stride_s; a.ksh = kv_stride_h;
    a.osb = o_stride_b; a.oss = o_stride_s; a.osh = o_stride_h; a.scale = scale; a.window = window;
    const bool precise = attention_precise_enabled();
    cudaStream_t st = as_stream(stream);
#define CALL(HD_, P_) launch_gqa_fwd<HD_, P_>(a, st)
    MFT_GQA_DISPATCH(CALL)
#undef CALL
    MFT_CHECK_LAUNCH("mft_gqa_attention_fwd");
    return 0;
}

extern "C" size_t mft_gqa_attention_bwd_workspace_bytes(int32_t B, int32_t H, int32_t KV, int32_t S, int32_t Hd) {
    size_t n = size_t(B) * H * S * sizeof(float);                         // D = rowsum(dO * O)
    n = (n + 255) & ~size_t(255);
    if (H != KV) n += 2 * size_t(B) * H * S * Hd * sizeof(float);         // per-query-head dK / dV partials
    return n;
}

extern "C" int mft_gqa_attention_bwd(const float* q, const float* k, const float* v, const float* key_bias, const float* ctx, const float* lse,
                                     const float* dctx, float* dq, float* dk, float* dv,
                                     int32_t B, int32_t H, int32_t KV, int32_t S, int32_t Hd,
                                     int64_t q_stride_b, int64_t q_stride_s, int64_t q_stride_h,
                                     int64_t kv_stride_b, int64_t kv_stride_s, int64_t kv_stride_h,
                                     int64_t o_stride_b, int64_t o_stride_s, int64_t o_stride_h,
                                     int64_t dq_stride_b, int64_t dq_stride_s, int64_t dq_stride_h,
                                     int64_t dkv_stride_b, int64_t dkv_stride_s, int64_t dkv_stride_h,
                                     float scale, int32_t window, void* workspace, size_t workspace_bytes, mft_stream_t stream) {
    if (B == 0 || H == 0 || S == 0) return 0;
    const int64_t strides[15] = {q_stride_b, q_stride_s, q_stride_h, kv_stride_b, kv_stride_s, kv_stride_h, o_stride_b, o_stride_s, o_stride_h,
                                 dq_stride_b, dq_stride_s, dq_stride_h, dkv_stride_b, dkv_stride_s, dkv_stride_h};
    if (int rc = gqa_check_common("mft_gqa_attention_bwd", q, k, v, B, H, KV, S, Hd, strides, 15)) return rc;
    if (!ctx || !lse || !dctx || !dq || !dk || !dv) return set_error(1, "mft_gqa_attention_bwd: null pointer");
    if (!aligned16(ctx) || !aligned16(dctx) || !aligned16(dq) || !aligned16(dk) || !aligned16(dv))
        return set_error(1, "mft_gqa_attention_bwd: tensors must be 16-byte aligned");
    const size_t need = mft_gqa_attention_bwd_workspace_bytes(B, H, KV, S, Hd);
    if (!workspace || workspace_bytes < need || !aligned16(workspace)) return set_error(1, "mft_gqa_attention_bwd: workspace too small (%zu < %zu)", workspace_bytes, need);
    GqaArgs a{};
    a.q = q; a.k = k; a.v = v; a.key_bias = key_bias; a.out = ctx; a.lse = const_cast<float*>(lse); a.dout = dctx;
    a.dq = dq; a.dk = dk; a.dv = dv; a.dsum = static_cast<float*>(workspace);
    a.B = B; a.H = H; a.KV = KV; a.S = S;
    a.qsb = q_stride_b; a.qss = q_stride_s; a.qsh = q_stride_h; a.ksb = kv_stride_b; a.kss = kv_stride_s; a.ksh = kv_stride_h;
    a.osb = o_stride_b; a.oss = o_stride_s; a.osh = o_stride_h;
    a.dqsb = dq_stride_b; a.dqss = dq_stride_s; a.dqsh = dq_stride_h; a.dksb = dkv_stride_b; a.dkss = dkv_stride_s; a.dksh = dkv_stride_h;
    a.scale = scale; a.window = window; a.partial = 0;
    size_t off = (size_t(B) * H * S * sizeof(float) + 255) & ~size_t(255);
    float* pk = reinterpret_cast<float*>(static_cast<char*>(workspace) + off);
    float* pv = pk + size_t(B) * H * S * Hd;
    const bool precise = attention_precise_enabled();
    cudaStream_t st = as_stream(stream);
#define CALL(HD_, P_) launch_gqa_bwd<HD_, P_>(a, pk, pv, st)
    MFT_GQA_DISPATCH(CALL)
#undef CALL
    MFT_CHECK_LAUNCH("mft_gqa_attention_bwd");
    return 0;
}
