// Rank-r LoRA factor products. These are bandwidth-bound (one streaming read of a [tokens, width] activation against a
// matrix with r <= 64 columns) and far too narrow for a 128-row tensor-core tile or a 64x64 SIMT tile, so they get two
// dedicated streaming kernels with enough independent 128-bit loads in flight per SM to approach HBM bandwidth:
//   lora_down : y[T,r]  = alpha * x[T,K] * A            (x*A in the forward pass, dY*B^T in the backward pass)
//   lora_tn   : out     = alpha * S[T,r]^T * W[T,q]      (dB = (xA)^T dY and dA = x^T d(xA): contraction over all tokens)
// Reference: the matmul/MatmulBackward calls LoRALinear::forward generates for slice.A / slice.B
// (nn/lora_linear.cpp:66-72, core/backward_functions.cpp:94-138). Algorithmic bytes: 4*T*K (+4*T*r) and 4*T*q (+4*T*r).
#include <cstdlib>
#include "common.cuh"
#include "gemm_internal.cuh"

namespace mft {

// lora_down: one warp owns RPW token rows; lane l takes k = l, l+32, ... (coalesced 128-byte row segments), keeps
// RPW x RMAX dot products in registers and warp-reduces them once per row. The A chunk sits in shared memory as
// [k][RMAX+4]: a lane reads its k-row with 128-bit loads and the +4 pitch keeps the eight lanes of a quarter-warp on
// distinct banks. The grid is sized by the launcher (RPW) so that every SM holds several CTAs: one CTA's staging of A
// overlaps another's streaming.
template <int RPW, int RMAX, int KC>
__global__ void __launch_bounds__(256, (RPW <= 2 && RMAX <= 16) ? 2 : 1)
lora_down_kernel(const float* __restrict__ x, int64_t ldx, const float* __restrict__ a, int64_t lda, int a_is_kr,
                 float* __restrict__ y, int64_t ldy, int64_t T, int K, int r, float alpha) {
    pdl_sync();
    constexpr int PITCH = RMAX + 4;
    extern __shared__ __align__(16) float At[];     // [KC][PITCH]
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const int64_t row0 = (int64_t(blockIdx.x) * 8 + warp) * RPW;
    const float* xr[RPW];
    bool live[RPW];
#pragma unroll
    for (int i = 0; i < RPW; ++i) { live[i] = row0 + i < T; xr[i] = x + (live[i] ? row0 + i : 0) * ldx; }
    float acc[RPW][RMAX];
#pragma unroll
    for (int i = 0; i < RPW; ++i)
#pragma unroll
        for (int c = 0; c < RMAX; ++c) acc[i][c] = 0.f;

    constexpr int NJ = KC / 32;
    constexpr int NS = KC * RMAX / 256;
    // Software pipeline over the K chunks: the global loads of chunk c+1 (this thread's x values and its share of A) are
    // in flight while chunk c is multiplied out of shared memory, so only the first chunk's latency is exposed.
    float xv[NJ][RPW], stage[NS];
    auto load_chunk = [&](int k0, float (&xo)[NJ][RPW], float (&so)[NS]) {
        const int kc = min(KC, K - k0);
#pragma unroll
        for (int j = 0; j < NJ; ++j) {
            const int kk = j * 32 + lane;
#pragma unroll
            for (int i = 0; i < RPW; ++i) xo[j][i] = (live[i] && kk < kc) ? ld_stream(xr[i] + k0 + kk) : 0.f;
        }
#pragma unroll
        for (int it = 0; it < NS; ++it) {
            const int idx = it * 256 + threadIdx.x;
            int kk, c;
            if (a_is_kr) { c = idx % RMAX; kk = idx / RMAX; } else { kk = idx % KC; c = idx / KC; }
            so[it] = (kk < kc && c < r) ? (a_is_kr ? __ldg(a + int64_t(k0 + kk) * lda + c) : __ldg(a + int64_t(c) * lda + k0 + kk)) : 0.f;
        }
    };
    load_chunk(0, xv, stage);
    for (int k0 = 0; k0 < K; k0 += KC) {
        float xc[NJ][RPW];
#pragma unroll
        for (int j = 0; j < NJ; ++j)
#pragma unroll
            for (int i = 0; i < RPW; ++i) xc[j][i] = xv[j][i];
        __syncthreads();     // the previous chunk's readers are done with At
#pragma unroll
        for (int it = 0; it < NS; ++it) {
            const int idx = it * 256 + threadIdx.x;
            int kk, c;
            if (a_is_kr) { c = idx % RMAX; kk = idx / RMAX; } else { kk = idx % KC; c = idx / KC; }
            At[kk * PITCH + c] = stage[it];
        }
        if (k0 + KC < K) load_chunk(k0 + KC, xv, stage);     // next chunk's loads fly during the FMAs below
        __syncthreads();
#pragma unroll
        for (int j = 0; j < NJ; ++j) {
            const float4* arow = reinterpret_cast<const float4*>(At + (j * 32 + lane) * PITCH);
#pragma unroll
            for (int c4 = 0; c4 < RMAX / 4; ++c4) {
                const float4 av = arow[c4];
#pragma unroll
                for (int i = 0; i < RPW; ++i) {
                    acc[i][c4 * 4 + 0] = fmaf(xc[j][i], av.x, acc[i][c4 * 4 + 0]);
                    acc[i][c4 * 4 + 1] = fmaf(xc[j][i], av.y, acc[i][c4 * 4 + 1]);
                    acc[i][c4 * 4 + 2] = fmaf(xc[j][i], av.z, acc[i][c4 * 4 + 2]);
                    acc[i][c4 * 4 + 3] = fmaf(xc[j][i], av.w, acc[i][c4 * 4 + 3]);
                }
            }
        }
    }
#pragma unroll
    for (int i = 0; i < RPW; ++i) {
#pragma unroll
        for (int c = 0; c < RMAX; ++c) {
            const float v = warp_sum(acc[i][c]);
            if (lane == (c & 31) && c < r && live[i]) y[(row0 + i) * ldy + c] = v * alpha;
        }
    }
}

// lora_down on the legacy tensor pipe (mma.sync m16n8k8 TF32), the default for r <= 16 and 16-byte aligned x.
// A CTA owns 16 token rows; its four warps split every 1024-wide K span into quarters. Everything a span needs is
// requested before anything is used: the span's slice of A goes to shared memory with cp.async ([k][18] floats:
// conflict-free B-fragment reads), and each lane issues all 32 of its 128-bit x loads (64 KB per CTA in flight), so one
// memory latency is exposed per span instead of one per tile. A-fragments are built from those registers: lane (g, t)
// holds x[row g / g+8][k0 + 4t .. 4t+3] per 16-wide k group, with the contraction slots relabelled (slot t <-> k0+4t+{0,2},
// slot t+4 <-> k0+4t+{1,3} for the two k8 steps). PRECISE = 3xTF32 split (exact mode), else TF32 operands like the
// large GEMMs. The FMA kernel above stays as the general path (any alignment, r up to 64).
constexpr int kDmSpan = 1024;         // k per span (4 warps x 256)
constexpr int kDmPitch = 18;
constexpr int kDmSmemBytes = kDmSpan * kDmPitch * 4;

__device__ __forceinline__ void cp_async_zfill(void* smem, const void* gmem, int bytes, int src_bytes) {
    const uint32_t s = static_cast<uint32_t>(__cvta_generic_to_shared(smem));
    if (bytes == 8) asm volatile("cp.async.ca.shared.global [%0], [%1], 8, %2;" :: "r"(s), "l"(gmem), "r"(src_bytes) : "memory");
    else asm volatile("cp.async.ca.shared.global [%0], [%1], 4, %2;" :: "r"(s), "l"(gmem), "r"(src_bytes) : "memory");
}

template <bool PRECISE>
__global__ void __launch_bounds__(128)
lora_down_mma_kernel(const float* __restrict__ x, int64_t ldx, const float* __restrict__ a, int64_t lda, int a_is_kr,
                     float* __restrict__ y, int64_t ldy, int64_t T, int K, int r, float alpha,
                     __nv_bfloat16* __restrict__ y16, int64_t ld16) {
    pdl_sync();   // y16 != NULL: bf16 copy of y
    extern __shared__ __align__(16) float As[];                       // [kDmSpan][18]; reused for the cross-warp reduction
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const int g = lane >> 2, t = lane & 3;
    const int64_t row0 = int64_t(blockIdx.x) * 16;
    const bool live0 = row0 + g < T, live1 = row0 + g + 8 < T;
    const float* x0 = x + (live0 ? row0 + g : 0) * ldx + warp * 256 + t * 4;
    const float* x1 = x + (live1 ? row0 + g + 8 : 0) * ldx + warp * 256 + t * 4;
    float acc[2][4];
#pragma unroll
    for (int n = 0; n < 2; ++n) acc[n][0] = acc[n][1] = acc[n][2] = acc[n][3] = 0.f;
    const bool pair_ok = a_is_kr && (lda & 1) == 0 && (reinterpret_cast<uintptr_t>(a) & 7u) == 0;

    for (int k0 = 0; k0 < K; k0 += kDmSpan) {
        if (k0) __syncthreads();                                      // the previous span's readers are done with As
        // ---- A slice -> shared memory (asynchronous; zero-filled outside K x r)
        if (pair_ok) {
            for (int idx = threadIdx.x; idx < kDmSpan * 8; idx += 128) {        // 8-byte pieces: (k, column pair)
                const int kk = idx >> 3, c = (idx & 7) * 2;
                const bool in = k0 + kk < K;
                const int nb = !in ? 0 : (c + 1 < r ? 8 : (c < r ? 4 : 0));
                cp_async_zfill(As + kk * kDmPitch + c, a + int64_t(in ? k0 + kk : 0) * lda + (c < r ? c : 0), 8, nb);
            }
        } else {
            for (int idx = threadIdx.x; idx < kDmSpan * 16; idx += 128) {
                int kk, c;
                if (a_is_kr) { c = idx & 15; kk = idx >> 4; } else { kk = idx % kDmSpan; c = idx / kDmSpan; }
                const bool in = k0 + kk < K && c < r;
                const float* src = a_is_kr ? a + int64_t(in ? k0 + kk : 0) * lda + (in ? c : 0) : a + int64_t(in ? c : 0) * lda + (in ? k0 + kk : 0);
                cp_async_zfill(As + kk * kDmPitch + c, src, 4, in ? 4 : 0);
            }
        }
        asm volatile("cp.async.commit_group;" ::: "memory");
        // ---- all of this lane's x for the span: 16 groups x 2 row halves
        float4 xq[2][16];
#pragma unroll
        for (int grp = 0; grp < 16; ++grp) {
            const bool in = k0 + warp * 256 + grp * 16 + t * 4 < K;    // K % 4 == 0: a float4 is inside or outside as a whole
            xq[0][grp] = (live0 && in) ? ld_stream(reinterpret_cast<const float4*>(x0 + k0 + grp * 16)) : make_float4(0.f, 0.f, 0.f, 0.f);
            xq[1][grp] = (live1 && in) ? ld_stream(reinterpret_cast<const float4*>(x1 + k0 + grp * 16)) : make_float4(0.f, 0.f, 0.f, 0.f);
        }
        // The barrier's predicate depends on every loaded value: that keeps all 32 loads in front of it (ptxas would
        // otherwise sink each read-only load to its first use, one exposed latency per group instead of one per span).
        uint32_t chk = 0xFFFFFFFFu;
#pragma unroll
        for (int grp = 0; grp < 16; ++grp)
#pragma unroll
            for (int h = 0; h < 2; ++h)
                chk &= __float_as_uint(xq[h][grp].x) & __float_as_uint(xq[h][grp].y) & __float_as_uint(xq[h][grp].z) & __float_as_uint(xq[h][grp].w);
        asm volatile("cp.async.wait_group 0;" ::: "memory");
        (void)__syncthreads_or(chk == 0xFFFFFFFFu);
#pragma unroll
        for (int grp = 0; grp < 16; ++grp) {
            const float* ab = As + (warp * 256 + grp * 16 + t * 4) * kDmPitch + g;
#pragma unroll
            for (int step = 0; step < 2; ++step) {                    // k8 step inside the 16-wide group: k = 4t + {0,1} / {2,3}
                float b[2][2];
#pragma unroll
                for (int n = 0; n < 2; ++n) {
                    b[n][0] = ab[(step * 2) * kDmPitch + n * 8];
                    b[n][1] = ab[(step * 2 + 1) * kDmPitch + n * 8];
                }
                const float a0 = step ? xq[0][grp].z : xq[0][grp].x, a2 = step ? xq[0][grp].w : xq[0][grp].y;
                const float a1 = step ? xq[1][grp].z : xq[1][grp].x, a3 = step ? xq[1][grp].w : xq[1][grp].y;
                if (PRECISE) {
                    uint32_t ah[4], al[4];
                    split_tf32(a0, ah[0], al[0]); split_tf32(a1, ah[1], al[1]); split_tf32(a2, ah[2], al[2]); split_tf32(a3, ah[3], al[3]);
#pragma unroll
                    for (int n = 0; n < 2; ++n) {
                        uint32_t bh0, bl0, bh1, bl1;
                        split_tf32(b[n][0], bh0, bl0); split_tf32(b[n][1], bh1, bl1);
                        mma_tf32(acc[n], al, bh0, bh1);
                        mma_tf32(acc[n], ah, bl0, bl1);
                        mma_tf32(acc[n], ah, bh0, bh1);
                    }
                } else {
                    const uint32_t af[4] = {f2tf32(a0), f2tf32(a1), f2tf32(a2), f2tf32(a3)};
#pragma unroll
                    for (int n = 0; n < 2; ++n) mma_tf32(acc[n], af, f2tf32(b[n][0]), f2tf32(b[n][1]));
                }
            }
        }
    }
    // the four warps hold partial sums over their K quarters: add them in warp order through shared memory
    __syncthreads();
    float* red = As;                                                  // [4 warps][16 rows][16 cols]
#pragma unroll
    for (int n = 0; n < 2; ++n) {
        float* base = red + warp * 256 + n * 8 + 2 * t;
        base[g * 16] = acc[n][0]; base[g * 16 + 1] = acc[n][1];
        base[(g + 8) * 16] = acc[n][2]; base[(g + 8) * 16 + 1] = acc[n][3];
    }
    __syncthreads();
    for (int idx = threadIdx.x; idx < 256; idx += 128) {
        const int rr = idx >> 4, c = idx & 15;
        const int64_t row = row0 + rr;
        if (row < T && c < r) {
            const float v = (((red[idx] + red[256 + idx]) + red[512 + idx]) + red[768 + idx]) * alpha;
            y[row * ldy + c] = v;
            if (y16) y16[row * ld16 + c] = __float2bfloat16_rn(v);
        }
    }
}

// lora_tn: CTA = kTnRG row-groups x kTnCT column-threads over kTnRows token rows; a thread owns CPT adjacent columns of
// W and RMAX accumulators per column. Row-groups are summed through shared memory in a fixed order, so one partial
// [r, cols] tile leaves the CTA; the partials over the token splits are summed (fixed order again) by the reduce kernel.
constexpr int kTnRows = 128, kTnRG = 4, kTnCT = 64;
// One arrival counter per column block of lora_tn; zero between launches (the reducing CTA resets its counter). Launches
// are stream-ordered on the library's single stream, so at most one launch uses the table at a time.
constexpr int kTnMaxTickets = 1024;
__device__ unsigned g_tn_tickets[kTnMaxTickets];

template <int CPT, int RMAX>
__device__ __forceinline__ void lora_tn_body(const float* __restrict__ s, int64_t lds, const float* __restrict__ w, int64_t ldw,
               float* __restrict__ partial, int64_t T, int q, int r, float* __restrict__ out, int64_t ldo, int out_is_rq,
               float alpha, int use_tickets, const int bx, const int by, const int nsplits, unsigned* __restrict__ tickets) {
    extern __shared__ __align__(16) float sm[];
    float* Ss = sm;                                  // [kTnRows][RMAX]
    float* red = sm + kTnRows * RMAX;                // [RMAX][kTnCT * CPT]
    const int ct = threadIdx.x % kTnCT, rg = threadIdx.x / kTnCT;
    const int col = (bx * kTnCT + ct) * CPT;
    const int64_t t0 = int64_t(by) * kTnRows;
    constexpr int RPG = kTnRows / kTnRG;             // rows per row-group (32): every W load of the thread is issued up front
    const bool colok = col < q;                      // q % CPT == 0 is guaranteed by the launcher
    float wv[RPG][CPT];
#pragma unroll
    for (int u = 0; u < RPG; ++u) {
        const int64_t row = t0 + rg * RPG + u;
        const bool ok = colok && row < T;
        if (CPT == 2) {
            const float2 v = ok ? ld_stream(reinterpret_cast<const float2*>(w + row * ldw + col)) : make_float2(0.f, 0.f);
            wv[u][0] = v.x; wv[u][CPT - 1] = v.y;
        } else {
            wv[u][0] = ok ? ld_stream(w + row * ldw + col) : 0.f;
        }
    }
    {
        constexpr int NS = kTnRows * RMAX / (kTnRG * kTnCT);
        float stage[NS];
#pragma unroll
        for (int it = 0; it < NS; ++it) {
            const int idx = it * (kTnRG * kTnCT) + threadIdx.x;
            const int tt = idx / RMAX, c = idx % RMAX;
            stage[it] = (t0 + tt < T && c < r) ? __ldg(s + (t0 + tt) * lds + c) : 0.f;
        }
#pragma unroll
        for (int it = 0; it < NS; ++it) Ss[it * (kTnRG * kTnCT) + threadIdx.x] = stage[it];
    }
    __syncthreads();
    float acc[CPT][RMAX];
#pragma unroll
    for (int e = 0; e < CPT; ++e)
#pragma unroll
        for (int c = 0; c < RMAX; ++c) acc[e][c] = 0.f;
#pragma unroll
    for (int u = 0; u < RPG; ++u) {
        const float4* srow = reinterpret_cast<const float4*>(Ss + (rg * RPG + u) * RMAX);
#pragma unroll
        for (int c4 = 0; c4 < RMAX / 4; ++c4) {
            const float4 sv = srow[c4];              // same address for the whole warp: broadcast
#pragma unroll
            for (int e = 0; e < CPT; ++e) {
                acc[e][c4 * 4 + 0] = fmaf(wv[u][e], sv.x, acc[e][c4 * 4 + 0]);
                acc[e][c4 * 4 + 1] = fmaf(wv[u][e], sv.y, acc[e][c4 * 4 + 1]);
                acc[e][c4 * 4 + 2] = fmaf(wv[u][e], sv.z, acc[e][c4 * 4 + 2]);
                acc[e][c4 * 4 + 3] = fmaf(wv[u][e], sv.w, acc[e][c4 * 4 + 3]);
            }
        }
    }
    // fixed-order reduction over the row-groups
    for (int g = 0; g < kTnRG; ++g) {
        if (rg == g) {
#pragma unroll
            for (int c = 0; c < RMAX; ++c)
#pragma unroll
                for (int e = 0; e < CPT; ++e) {
                    float* dst = red + c * (kTnCT * CPT) + ct * CPT + e;
                    *dst = (g == 0) ? acc[e][c] : *dst + acc[e][c];
                }
        }
        __syncthreads();
    }
    constexpr int W = kTnCT * CPT;                   // columns of this CTA
    float* P = partial + size_t(by) * size_t(r) * q;
    for (int idx = threadIdx.x; idx < r * W; idx += kTnRG * kTnCT) {
        const int c = idx / W, cc = idx % W;
        const int gcol = bx * W + cc;
        if (gcol < q) P[size_t(c) * q + gcol] = red[c * W + cc];
    }
    if (!use_tickets) return;                        // separate reduce kernel follows
    // The CTA that finishes last for this column block sums the token splits in index order (so the result does not
    // depend on which CTA that is), applies alpha and writes the output; it leaves the ticket counter at zero again.
    __shared__ bool last;
    __threadfence();
    __syncthreads();
    if (threadIdx.x == 0) last = atomicAdd(&tickets[bx], 1u) == unsigned(nsplits - 1);
    __syncthreads();
    if (!last) return;
    __threadfence();
    const int splits = nsplits;
    const size_t n = size_t(r) * q;
    for (int idx = threadIdx.x; idx < r * W; idx += kTnRG * kTnCT) {
        const int c = idx / W, cc = idx % W;
        const int gcol = bx * W + cc;
        if (gcol >= q) continue;
        const float* src = partial + size_t(c) * q + gcol;
        float v = 0.f;
        int sp = 0;
        for (; sp + 8 <= splits; sp += 8) {
            float p[8];
#pragma unroll
            for (int u = 0; u < 8; ++u) p[u] = __ldcg(src + size_t(sp + u) * n);
#pragma unroll
            for (int u = 0; u < 8; ++u) v += p[u];
        }
        for (; sp < splits; ++sp) v += __ldcg(src + size_t(sp) * n);
        v *= alpha;
        if (out_is_rq) out[int64_t(c) * ldo + gcol] = v; else out[int64_t(gcol) * ldo + c] = v;
    }
    if (threadIdx.x == 0) tickets[bx] = 0;
}

template <int CPT, int RMAX>
__global__ void __launch_bounds__(kTnRG * kTnCT)
lora_tn_kernel(const float* __restrict__ s, int64_t lds, const float* __restrict__ w, int64_t ldw,
               float* __restrict__ partial, int64_t T, int q, int r, float* __restrict__ out, int64_t ldo, int out_is_rq,
               float alpha, int use_tickets) {
    pdl_sync();
    lora_tn_body<CPT, RMAX>(s, lds, w, ldw, partial, T, q, r, out, ldo, out_is_rq, alpha, use_tickets, int(blockIdx.x), int(blockIdx.y), int(gridDim.y),
                            g_tn_tickets);
}

// Grouped form: blockIdx.z selects one of many independent products (every adapter gradient of a backward pass, deferred
// to its end), each with its own shapes; CTAs outside a problem's (column blocks x token splits) grid exit at once.
constexpr int kTnGroupTickets = 32768;
__device__ unsigned g_tn_group_tickets[kTnGroupTickets];
__global__ void __launch_bounds__(kTnRG * kTnCT)
lora_tn_group_kernel(const mft_lora_tn_problem* __restrict__ probs) {
    pdl_sync();
    const mft_lora_tn_problem p = probs[blockIdx.z];
    if (int(blockIdx.x) >= p.colblocks || int(blockIdx.y) >= p.splits) return;
    lora_tn_body<2, 16>(p.s, p.lds, p.w, p.ldw, p.partial, p.T, int(p.q), p.r, p.out, p.ldo, p.out_is_rq, p.alpha, p.use_tickets, int(blockIdx.x),
                        int(blockIdx.y), p.splits, g_tn_group_tickets + p.ticket_base);
}

// out[c][col] (out_is_rq) or out[col][c] = alpha * sum over splits, fixed order
__global__ void lora_tn_reduce_kernel(const float* __restrict__ partial, float* __restrict__ out, int64_t ldo, int out_is_rq,
                                      int splits, int q, int r, float alpha) {
    pdl_sync();
    const int64_t n = int64_t(r) * q;
    for (int64_t idx = blockIdx.x * int64_t(blockDim.x) + threadIdx.x; idx < n; idx += int64_t(gridDim.x) * blockDim.x) {
        const int c = int(idx / q), col = int(idx % q);
        float v = 0.f;
        int sp = 0;
        for (; sp + 8 <= splits; sp += 8) {          // eight loads in flight, summed in index order
            float pv[8];
#pragma unroll
            for (int u = 0; u < 8; ++u) pv[u] = __ldcg(partial + size_t(sp + u) * n + idx);
#pragma unroll
            for (int u = 0; u < 8; ++u) v += pv[u];
        }
        for (; sp < splits; ++sp) v += __ldcg(partial + size_t(sp) * n + idx);
        v *= alpha;
        if (out_is_rq) out[int64_t(c) * ldo + col] = v; else out[int64_t(col) * ldo + c] = v;
    }
}

// grouped reduce for the problems whose split list is too long for the last-CTA sum
__global__ void lora_tn_group_reduce_kernel(const mft_lora_tn_problem* __restrict__ probs) {
    pdl_sync();
    const mft_lora_tn_problem p = probs[blockIdx.y];
    if (p.use_tickets) return;
    const int64_t n = int64_t(p.r) * p.q;
    for (int64_t idx = blockIdx.x * int64_t(blockDim.x) + threadIdx.x; idx < n; idx += int64_t(gridDim.x) * blockDim.x) {
        const int c = int(idx / p.q), col = int(idx % p.q);
        float v = 0.f;
        int sp = 0;
        for (; sp + 8 <= p.splits; sp += 8) {
            float pv[8];
#pragma unroll
            for (int u = 0; u < 8; ++u) pv[u] = __ldcg(p.partial + size_t(sp + u) * n + idx);
#pragma unroll
            for (int u = 0; u < 8; ++u) v += pv[u];
        }
        for (; sp < p.splits; ++sp) v += __ldcg(p.partial + size_t(sp) * n + idx);
        v *= p.alpha;
        if (p.out_is_rq) p.out[int64_t(c) * p.ldo + col] = v; else p.out[int64_t(col) * p.ldo + c] = v;
    }
}

}  // namespace mft

using namespace mft;

template <int RPW, int RMAX, int KC>
static void launch_down(const float* x, int64_t ldx, const float* a, int64_t lda, int a_is_kr, float* y, int64_t ldy, int64_t T, int K,
                        int r, float alpha, cudaStream_t st) {
    constexpr size_t smem = size_t(KC) * (RMAX + 4) * sizeof(float);
    static bool attr[kMaxDevices] = {};
    if (first_use_on_device(attr)) cudaFuncSetAttribute(lora_down_kernel<RPW, RMAX, KC>, cudaFuncAttributeMaxDynamicSharedMemorySize, int(smem));
    pdl_launch(lora_down_kernel<RPW, RMAX, KC>, unsigned(ceil_div(T, int64_t(8 * RPW))), 256, smem, st)(x, ldx, a, lda, a_is_kr, y, ldy, T, K, r, alpha);
}

static int lora_down_impl(const float* x, int64_t ldx, const float* a, int64_t lda, int a_is_kr, float* y, int64_t ldy,
                          int64_t T, int64_t K, int32_t r, float alpha, mft_stream_t stream, void* y16v, int64_t ld16) {
    __nv_bfloat16* y16 = static_cast<__nv_bfloat16*>(y16v);
    if (T == 0 || r == 0) return 0;
    if (!x || !a || !y) return set_error(1, "mft_lora_down: null pointer");
    if (r < 1 || r > 64) return set_error(1, "mft_lora_down: rank %d outside 1..64", r);
    if (K < 1 || K > (int64_t(1) << 30)) return set_error(1, "mft_lora_down: K out of range");
    if (ceil_div(T, int64_t(8)) > 0x7fffffff) return set_error(1, "mft_lora_down: too many rows");
    cudaStream_t st = as_stream(stream);
    const int k = int(K);
    static const bool mma_on = [] { const char* e = getenv("MFT_LORA_DOWN_MMA"); return !(e && e[0] == '0'); }();
    if (mma_on && r <= 16 && (K & 3) == 0 && (ldx & 3) == 0 && aligned16(x)) {
        const bool precise = !tensor_cores_enabled();          // exact mode: 3xTF32 split, fp32-accurate
        static bool attr[kMaxDevices] = {};
        if (first_use_on_device(attr)) {
            cudaFuncSetAttribute(lora_down_mma_kernel<true>, cudaFuncAttributeMaxDynamicSharedMemorySize, kDmSmemBytes);
            cudaFuncSetAttribute(lora_down_mma_kernel<false>, cudaFuncAttributeMaxDynamicSharedMemorySize, kDmSmemBytes);
        }
        const unsigned grid = unsigned(ceil_div(T, int64_t(16)));
        if (precise) pdl_launch(lora_down_mma_kernel<true>, grid, 128, kDmSmemBytes, st)(x, ldx, a, lda, a_is_kr, y, ldy, T, k, r, alpha, y16, ld16);
        else pdl_launch(lora_down_mma_kernel<false>, grid, 128, kDmSmemBytes, st)(x, ldx, a, lda, a_is_kr, y, ldy, T, k, r, alpha, y16, ld16);
        MFT_CHECK_LAUNCH("mft_lora_down");
        return 0;
    }
    if (r <= 16) {
        // rows per warp: as many as still leave >= 2 CTAs per SM
        if (T >= int64_t(32) * 2 * kNumSMs) launch_down<4, 16, 128>(x, ldx, a, lda, a_is_kr, y, ldy, T, k, r, alpha, st);
        else if (T >= int64_t(16) * kNumSMs) launch_down<2, 16, 256>(x, ldx, a, lda, a_is_kr, y, ldy, T, k, r, alpha, st);
        else launch_down<1, 16, 256>(x, ldx, a, lda, a_is_kr, y, ldy, T, k, r, alpha, st);
    } else {
        launch_down<1, 64, 64>(x, ldx, a, lda, a_is_kr, y, ldy, T, k, r, alpha, st);
    }
    MFT_CHECK_LAUNCH("mft_lora_down");
    if (y16) return mft_cast_bf16(y, ldy, y16, ld16, T, r, stream);      // FMA path: separate cast
    return 0;
}

extern "C" int mft_lora_down(const float* x, int64_t ldx, const float* a, int64_t lda, int a_is_kr, float* y, int64_t ldy,
                             int64_t T, int64_t K, int32_t r, float alpha, mft_stream_t stream) {
    return lora_down_impl(x, ldx, a, lda, a_is_kr, y, ldy, T, K, r, alpha, stream, nullptr, 0);
}

// same, plus a bf16 copy of y (row pitch ld_bf16 elements) — the K-major A operand of the rank-r segment of the bf16 GEMM
extern "C" int mft_lora_down_ex(const float* x, int64_t ldx, const float* a, int64_t lda, int a_is_kr, float* y, int64_t ldy, void* y_bf16,
                                int64_t ld_bf16, int64_t T, int64_t K, int32_t r, float alpha, mft_stream_t stream) {
    return lora_down_impl(x, ldx, a, lda, a_is_kr, y, ldy, T, K, r, alpha, stream, y_bf16, ld_bf16);
}

extern "C" size_t mft_lora_tn_workspace_bytes(int64_t T, int64_t q, int32_t r) {
    return size_t(ceil_div(T, kTnRows)) * size_t(r) * size_t(q) * sizeof(float);
}

extern "C" int mft_lora_tn(const float* s, int64_t lds, const float* w, int64_t ldw, float* out, int64_t ldo, int out_is_rq,
                           int64_t T, int64_t q, int32_t r, float alpha, void* workspace, size_t workspace_bytes, mft_stream_t stream) {
    if (q == 0 || r == 0) return 0;
    if (!s || !w || !out) return set_error(1, "mft_lora_tn: null pointer");
    if (r < 1 || r > 64) return set_error(1, "mft_lora_tn: rank %d outside 1..64", r);
    if (T == 0) return set_error(1, "mft_lora_tn: empty contraction");
    if (!workspace || workspace_bytes < mft_lora_tn_workspace_bytes(T, q, r)) return set_error(1, "mft_lora_tn: workspace too small");
    const int64_t splits = ceil_div(T, kTnRows);
    if (splits > 65535) return set_error(1, "mft_lora_tn: too many tokens");
    cudaStream_t st = as_stream(stream);
    float* partial = static_cast<float*>(workspace);
    const bool vec_layout = (q & 1) == 0 && (ldw & 1) == 0 && (reinterpret_cast<uintptr_t>(w) & 7) == 0;
    if (r > 16 && vec_layout) {
        // ranks above 16: slices of 16 columns of S through the 16-wide kernel (W is re-read per slice from L2 / HBM, which is
        // still several times faster than the one-accumulator-per-rank FMA form: 64 accumulators per thread, one column each)
        for (int32_t r0 = 0; r0 < r; r0 += 16) {
            const int32_t rs = r - r0 < 16 ? r - r0 : 16;
            float* o = out_is_rq ? out + int64_t(r0) * ldo : out + r0;
            int rc = mft_lora_tn(s + r0, lds, w, ldw, o, ldo, out_is_rq, T, q, rs, alpha, workspace, workspace_bytes, stream);
            if (rc) return rc;
        }
        return 0;
    }
    const bool vec = r <= 16 && vec_layout;
    const int64_t colblocks = ceil_div(q, vec ? kTnCT * 2 : kTnCT);
    // the last-arriving CTA sums the splits itself when there are few of them; a long split list is summed faster by a
    // separate, fully parallel launch
    const int tk = (colblocks <= kTnMaxTickets && splits <= 64) ? 1 : 0;      // 64 splits x [r x 128] partials = 512 KB from L2 for the last CTA of a column block
    if (vec) {
        const size_t smem = (size_t(kTnRows) * 16 + size_t(16) * kTnCT * 2) * sizeof(float);
        pdl_launch(lora_tn_kernel<2, 16>, dim3(unsigned(colblocks), unsigned(splits)), kTnRG * kTnCT, smem, st)(s, lds, w, ldw, partial, T, int(q), r, out, ldo,
                                                                                                 out_is_rq, alpha, tk);
    } else {
        const size_t smem = (size_t(kTnRows) * 64 + size_t(64) * kTnCT) * sizeof(float);
        static bool attr[kMaxDevices] = {};
        if (first_use_on_device(attr)) cudaFuncSetAttribute(lora_tn_kernel<1, 64>, cudaFuncAttributeMaxDynamicSharedMemorySize, int(smem));
        pdl_launch(lora_tn_kernel<1, 64>, dim3(unsigned(colblocks), unsigned(splits)), kTnRG * kTnCT, smem, st)(s, lds, w, ldw, partial, T, int(q), r, out, ldo,
                                                                                                 out_is_rq, alpha, tk);
    }
    if (!tk) {
        add_launches(1);
        pdl_launch(lora_tn_reduce_kernel, stream_grid(int64_t(r) * q, 256), 256, 0, st)(partial, out, ldo, out_is_rq, int(splits), int(q), r, alpha);
    }
    MFT_CHECK_LAUNCH("mft_lora_tn");
    return 0;
}

// ---------------------------------------------------------------------------------------------------------------------
// lora_down on bf16 operands (bf16 operand mode): y[T,r] = alpha * x[T,K] * A with x given as the bf16 copy its producer
// wrote (half the bytes of the fp32 rows) and A as its K-major bf16 shadow [r][K] (the same array the rank-r segment of the
// big GEMM reads). A CTA owns 16 token rows at a time and loops over row blocks; its four warps split K into quarters.
// The whole shadow goes to shared memory once per CTA as straight 16-byte cp.async copies (no transposition, 32 KB for
// K = 1024) while the x loads of the first block are already in flight; products run as mma.sync m16n8k16 bf16 with fp32
// accumulation. Fragments are built straight from 16-byte loads by relabelling the contraction index inside each 32-wide k
// group (slot 2t,2t+1 <-> k 8t+0,1; slot 2t+8,2t+9 <-> k 8t+2,3 for the first MMA; +4 for the second) — x and A use the same
// relabelling, so no shuffle is needed. Algorithmic bytes: 2*T*K. HBM-bound.
namespace mft {

__device__ __forceinline__ void mma_bf16_16816(float (&c)[4], uint32_t a0, uint32_t a1, uint32_t a2, uint32_t a3, uint32_t b0, uint32_t b1) {
    asm volatile(
        "mma.sync.aligned.m16n8k16.row.col.f32.bf16.bf16.f32 {%0,%1,%2,%3}, {%4,%5,%6,%7}, {%8,%9}, {%0,%1,%2,%3};"
        : "+f"(c[0]), "+f"(c[1]), "+f"(c[2]), "+f"(c[3])
        : "r"(a0), "r"(a1), "r"(a2), "r"(a3), "r"(b0), "r"(b1));
}

constexpr int kLbPad = 32;      // bf16 elements of padding per shadow row in shared memory (64 bytes: conflict-free 16-byte reads)

template <int NT>     // output columns in tiles of 8: rank <= 8 * NT
__global__ void __launch_bounds__(128)
lora_down_bf16_kernel(const __nv_bfloat16* __restrict__ x, int64_t ldx, const __nv_bfloat16* __restrict__ at, int64_t lda,
                      float* __restrict__ y, int64_t ldy, __nv_bfloat16* __restrict__ y16, int64_t ld16, int64_t T, int K, int r, float alpha) {
    pdl_sync();
    extern __shared__ __align__(16) uint8_t lb_smem[];
    const int KP = K + kLbPad;
    constexpr int RC = 8 * NT;                                                       // shadow rows / output columns held
    __nv_bfloat16* sA = reinterpret_cast<__nv_bfloat16*>(lb_smem);                 // [RC][KP]
    float* red = reinterpret_cast<float*>(lb_smem + size_t(RC) * KP * 2);           // [4 warps][16 rows][RC cols]
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31, g = lane >> 2, t = lane & 3;
    const int KQ = K >> 2;                                                           // this warp's K quarter (multiple of 32)
    // shadow -> shared memory (rows >= r are zero)
    {
        const int chunks = K >> 3;                                                   // 16-byte pieces per row
        for (int idx = threadIdx.x; idx < RC * chunks; idx += 128) {
            const int c = idx / chunks, pc = idx % chunks;
            const uint32_t dst = static_cast<uint32_t>(__cvta_generic_to_shared(sA + c * KP + pc * 8));
            const int nb = c < r ? 16 : 0;
            asm volatile("cp.async.cg.shared.global [%0], [%1], 16, %2;" :: "r"(dst), "l"(at + int64_t(c < r ? c : 0) * lda + pc * 8), "r"(nb) : "memory");
        }
        asm volatile("cp.async.commit_group;" ::: "memory");
    }
    const int64_t nblk = (T + 15) >> 4;
    bool first = true;
    for (int64_t blk = blockIdx.x; blk < nblk; blk += gridDim.x) {
        const int64_t row0 = blk << 4;
        const bool live0 = row0 + g < T, live1 = row0 + g + 8 < T;
        const __nv_bfloat16* x0 = x + (live0 ? row0 + g : 0) * ldx + warp * KQ + 8 * t;
        const __nv_bfloat16* x1 = x + (live1 ? row0 + g + 8 : 0) * ldx + warp * KQ + 8 * t;
        float acc[NT][4];
#pragma unroll
        for (int n = 0; n < NT; ++n) acc[n][0] = acc[n][1] = acc[n][2] = acc[n][3] = 0.f;
        for (int kc = 0; kc < KQ; kc += 256) {
            uint4 xa[8], xb[8];
            const int nit = min(8, (KQ - kc) >> 5);
#pragma unroll
            for (int i = 0; i < 8; ++i) {
                xa[i] = (i < nit && live0) ? __ldcs(reinterpret_cast<const uint4*>(x0 + kc + 32 * i)) : make_uint4(0u, 0u, 0u, 0u);
                xb[i] = (i < nit && live1) ? __ldcs(reinterpret_cast<const uint4*>(x1 + kc + 32 * i)) : make_uint4(0u, 0u, 0u, 0u);
            }
            if (first) {                                                             // the shadow must have landed before its first use
                asm volatile("cp.async.wait_group 0;" ::: "memory");
                __syncthreads();
                first = false;
            }
#pragma unroll
            for (int i = 0; i < 8; ++i) {
                if (i >= nit) break;
                const __nv_bfloat16* bp = sA + g * KP + warp * KQ + kc + 32 * i + 8 * t;
#pragma unroll
                for (int n = 0; n < NT; ++n) {
                    const uint4 b0 = *reinterpret_cast<const uint4*>(bp + n * 8 * KP);   // output column 8n + g
                    mma_bf16_16816(acc[n], xa[i].x, xb[i].x, xa[i].y, xb[i].y, b0.x, b0.y);
                    mma_bf16_16816(acc[n], xa[i].z, xb[i].z, xa[i].w, xb[i].w, b0.z, b0.w);
                }
            }
        }
        // the four warps hold partial sums over their K quarters: add them in warp order through shared memory
#pragma unroll
        for (int n = 0; n < NT; ++n) {
            float* base = red + warp * (16 * RC) + n * 8 + 2 * t;
            base[g * RC] = acc[n][0]; base[g * RC + 1] = acc[n][1];
            base[(g + 8) * RC] = acc[n][2]; base[(g + 8) * RC + 1] = acc[n][3];
        }
        __syncthreads();
        for (int idx = threadIdx.x; idx < 16 * RC; idx += 128) {
            const int rr = idx / RC, c = idx % RC;
            const int64_t row = row0 + rr;
            if (row < T && c < r) {
                const float v = (((red[idx] + red[16 * RC + idx]) + red[2 * 16 * RC + idx]) + red[3 * 16 * RC + idx]) * alpha;
                y[row * ldy + c] = v;
                if (y16) y16[row * ld16 + c] = __float2bfloat16_rn(v);
            }
        }
        __syncthreads();                                                             // `red` is rewritten by the next block
    }
    asm volatile("cp.async.wait_group 0;" ::: "memory");                             // a CTA without any block still drains its copies
}

}  // namespace mft

extern "C" int mft_lora_down_bf16(const void* x_bf16, int64_t ldx, const void* a_kmajor_bf16, int64_t lda, float* y, int64_t ldy, void* y_bf16,
                                  int64_t ld_bf16, int64_t T, int64_t K, int32_t r, float alpha, mft_stream_t stream) {
    if (T == 0 || r == 0) return 0;
    if (!x_bf16 || !a_kmajor_bf16 || !y) return set_error(1, "mft_lora_down_bf16: null pointer");
    if (r < 1 || r > 64) return set_error(1, "mft_lora_down_bf16: rank %d outside 1..64", r);
    const int nt = r <= 8 ? 1 : (r <= 16 ? 2 : (r <= 32 ? 4 : 8));
    const int64_t kmax = nt <= 2 ? 6144 : (nt == 4 ? 3072 : 1536);          // the whole shadow must fit in shared memory
    if (K < 128 || (K & 127) || K > kmax) return set_error(1, "mft_lora_down_bf16: K must be a multiple of 128 in 128..%lld for rank %d", (long long)kmax, r);
    if ((ldx & 7) || (lda & 7) || !aligned16(x_bf16) || !aligned16(a_kmajor_bf16)) return set_error(1, "mft_lora_down_bf16: operands must be 16-byte aligned with pitches that are multiples of 8");
    const size_t smem = size_t(8 * nt) * size_t(K + kLbPad) * 2 + size_t(4) * 16 * 8 * nt * sizeof(float);
    static bool attr_set[kMaxDevices] = {};
    if (first_use_on_device(attr_set)) {
        cudaFuncSetAttribute(lora_down_bf16_kernel<1>, cudaFuncAttributeMaxDynamicSharedMemorySize, 220 * 1024);
        cudaFuncSetAttribute(lora_down_bf16_kernel<2>, cudaFuncAttributeMaxDynamicSharedMemorySize, 220 * 1024);
        cudaFuncSetAttribute(lora_down_bf16_kernel<4>, cudaFuncAttributeMaxDynamicSharedMemorySize, 220 * 1024);
        cudaFuncSetAttribute(lora_down_bf16_kernel<8>, cudaFuncAttributeMaxDynamicSharedMemorySize, 220 * 1024);
    }
    const int64_t nblk = ceil_div(T, int64_t(16));
    const int64_t cap = int64_t(kNumSMs) * 4;
    const unsigned grid = unsigned(nblk < cap ? nblk : cap);
    const __nv_bfloat16* xp = static_cast<const __nv_bfloat16*>(x_bf16);
    const __nv_bfloat16* ap = static_cast<const __nv_bfloat16*>(a_kmajor_bf16);
    __nv_bfloat16* y16 = static_cast<__nv_bfloat16*>(y_bf16);
    cudaStream_t st = as_stream(stream);
    if (nt == 1) pdl_launch(lora_down_bf16_kernel<1>, grid, 128, smem, st)(xp, ldx, ap, lda, y, ldy, y16, ld_bf16, T, int(K), r, alpha);
    else if (nt == 2) pdl_launch(lora_down_bf16_kernel<2>, grid, 128, smem, st)(xp, ldx, ap, lda, y, ldy, y16, ld_bf16, T, int(K), r, alpha);
    else if (nt == 4) pdl_launch(lora_down_bf16_kernel<4>, grid, 128, smem, st)(xp, ldx, ap, lda, y, ldy, y16, ld_bf16, T, int(K), r, alpha);
    else pdl_launch(lora_down_bf16_kernel<8>, grid, 128, smem, st)(xp, ldx, ap, lda, y, ldy, y16, ld_bf16, T, int(K), r, alpha);
    MFT_CHECK_LAUNCH("mft_lora_down_bf16");
    return 0;
}

// ---- grouped lora_tn: many independent products in one launch (problem table in device memory)
extern "C" int mft_lora_tn_group_eligible(int64_t T, int64_t q, int32_t r, const float* w, int64_t ldw) {
    return (r >= 1 && r <= 16 && T > 0 && q > 0 && (q & 1) == 0 && (ldw & 1) == 0 && (reinterpret_cast<uintptr_t>(w) & 7) == 0 &&
            ceil_div(T, kTnRows) <= 65535) ? 1 : 0;
}
extern "C" int32_t mft_lora_tn_group_colblocks(int64_t q) { return int32_t(ceil_div(q, int64_t(kTnCT) * 2)); }
extern "C" int32_t mft_lora_tn_group_splits(int64_t T) { return int32_t(ceil_div(T, kTnRows)); }

extern "C" int mft_lora_tn_group(const mft_lora_tn_problem* dev_problems, int32_t num_problems, int32_t max_colblocks, int32_t max_splits,
                                 int32_t any_long_split, int64_t max_rq, mft_stream_t stream) {
    if (num_problems == 0) return 0;
    if (!dev_problems) return set_error(1, "mft_lora_tn_group: null problem table");
    if (num_problems > 65535 || max_splits > 65535 || max_colblocks < 1 || max_splits < 1) return set_error(1, "mft_lora_tn_group: grid out of range");
    cudaStream_t st = as_stream(stream);
    const size_t smem = (size_t(kTnRows) * 16 + size_t(16) * kTnCT * 2) * sizeof(float);
    pdl_launch(lora_tn_group_kernel, dim3(unsigned(max_colblocks), unsigned(max_splits), unsigned(num_problems)), kTnRG * kTnCT, smem, st)(dev_problems);
    if (any_long_split) {
        add_launches(1);
        int64_t bx = ceil_div(max_rq > 0 ? max_rq : 1, int64_t(256));
        if (bx > 64) bx = 64;
        pdl_launch(lora_tn_group_reduce_kernel, dim3(unsigned(bx), unsigned(num_problems)), 256, 0, st)(dev_problems);
    }
    MFT_CHECK_LAUNCH("mft_lora_tn_group");
    return 0;
}
extern "C" int32_t mft_lora_tn_group_ticket_capacity(void) { return kTnGroupTickets; }
