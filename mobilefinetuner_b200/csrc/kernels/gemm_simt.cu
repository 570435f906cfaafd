// SIMT fp32 GEMM: the general-layout path (any transposes, any alignment, tiny or skinny shapes) and the
// bit-faithful fp32 comparison path for the tensor-core kernel. 64x64 CTA tile, BK = 16, 256 threads,
// 4x4 register micro-tile, global loads mapped so that the contiguous dimension is the fast one.
#include "common.cuh"
#include "gemm_internal.cuh"

namespace mft {

constexpr int SBM = 64, SBN = 64, SBK = 16;

// One thread's share (CNT elements) of an EXT x SBK operand tile, fetched into registers so that the next k-block's
// global loads are in flight while the current one is multiplied (software pipelining; latency-bound otherwise).
// `contig_k` = the operand's contiguous dimension is K (A with trans_a=0 / B with trans_b=1).
template <int CNT> struct TileRegs { float v[CNT]; };
template <int EXT>
__device__ __forceinline__ TileRegs<EXT * SBK / 256> fetch_tile(const float* __restrict__ p, int64_t ld, bool contig_k,
                                                                int mn0, int k0, int mn_max, int k_max) {
    constexpr int CNT = EXT * SBK / 256;
    TileRegs<CNT> r;
#pragma unroll
    for (int j = 0; j < CNT; ++j) {
        int idx = threadIdx.x + 256 * j;
        int kk, mm;
        if (contig_k) { kk = idx % SBK; mm = idx / SBK; } else { mm = idx % EXT; kk = idx / EXT; }
        int gm = mn0 + mm, gk = k0 + kk;
        float v = 0.f;
        if (gm < mn_max && gk < k_max) v = contig_k ? __ldg(p + int64_t(gm) * ld + gk) : __ldg(p + int64_t(gk) * ld + gm);
        r.v[j] = v;
    }
    return r;
}
template <int EXT>
__device__ __forceinline__ void stash_tile(float (*s)[EXT + 4], const TileRegs<EXT * SBK / 256>& r, bool contig_k) {
    constexpr int CNT = EXT * SBK / 256;
#pragma unroll
    for (int j = 0; j < CNT; ++j) {
        int idx = threadIdx.x + 256 * j;
        int kk, mm;
        if (contig_k) { kk = idx % SBK; mm = idx / SBK; } else { mm = idx % EXT; kk = idx / EXT; }
        s[kk][mm] = r.v[j];
    }
}

// CTA tile = (16*MI) x (16*NJ) with MI, NJ in {1, 4}: 64x64 for general products, 64x16 / 16x64 for the rank-r LoRA
// factors (narrow output: a 64-wide tile would waste 3/4 of the FMAs and smem traffic on padding).
template <int MI, int NJ>
__global__ void __launch_bounds__(256)
gemm_simt_kernel(GemmParams p) {
    pdl_sync();
    constexpr int BM = 16 * MI, BN = 16 * NJ;
    __shared__ __align__(16) float As[SBK][BM + 4];
    __shared__ __align__(16) float Bs[SBK][BN + 4];
    const int bz = blockIdx.z / p.ksplit, ks = blockIdx.z % p.ksplit;
    const float* A = static_cast<const float*>(p.a) + bz * p.stride_a;
    const float* B = static_cast<const float*>(p.b) + bz * p.stride_b;
    float* C = p.c + bz * p.stride_c;
    float* G = p.gelu_out ? p.gelu_out + bz * p.stride_c : nullptr;
    const int m0 = blockIdx.y * BM, n0 = blockIdx.x * BN;
    const int tx = threadIdx.x % 16, ty = threadIdx.x / 16;
    float acc[MI][NJ];
#pragma unroll
    for (int i = 0; i < MI; ++i)
#pragma unroll
        for (int j = 0; j < NJ; ++j) acc[i][j] = 0.f;

    // k-blocks of segment 1 owned by this CTA (split-K), followed by all of segment 2 (never split)
    int kb_begin = 0, kb_end = (p.k + SBK - 1) / SBK;
    if (p.ksplit > 1) {
        const int per = (kb_end + p.ksplit - 1) / p.ksplit;
        kb_begin = min(kb_end, ks * per);
        kb_end = min(kb_end, (ks + 1) * per);
    }
    const int nb1 = kb_end - kb_begin;
    const int nb2 = (p.k2 > 0 && p.a2 && p.b2) ? (p.k2 + SBK - 1) / SBK : 0;
    const int nb = nb1 + nb2;
    const float* A2 = static_cast<const float*>(p.a2);
    const float* B2 = static_cast<const float*>(p.b2);
    const bool a_ck1 = p.trans_a == 0, b_ck1 = p.trans_b != 0;

    TileRegs<MI> ra;
    TileRegs<NJ> rb;
    auto fetch = [&](int i) {
        if (i < nb1) {
            const int k0 = (kb_begin + i) * SBK;
            ra = fetch_tile<BM>(A, p.lda, a_ck1, m0, k0, p.m, p.k);
            rb = fetch_tile<BN>(B, p.ldb, b_ck1, n0, k0, p.n, p.k);
        } else {
            const int k0 = (i - nb1) * SBK;
            ra = fetch_tile<BM>(A2, p.lda2, true, m0, k0, p.m, p.k2);
            rb = fetch_tile<BN>(B2, p.ldb2, true, n0, k0, p.n, p.k2);
        }
    };

    if (nb > 0) fetch(0);
    for (int i = 0; i < nb; ++i) {
        const bool seg2 = i >= nb1;
        stash_tile<BM>(As, ra, seg2 ? true : a_ck1);
        stash_tile<BN>(Bs, rb, seg2 ? true : b_ck1);
        __syncthreads();
        if (i + 1 < nb) fetch(i + 1);                  // next tile's loads overlap this tile's FMAs
#pragma unroll
        for (int kk = 0; kk < SBK; ++kk) {
            float ar[MI], br[NJ];
            if (MI == 4) { float4 v = *reinterpret_cast<const float4*>(&As[kk][ty * 4]); ar[0] = v.x; ar[MI > 1 ? 1 : 0] = v.y; ar[MI > 2 ? 2 : 0] = v.z; ar[MI > 3 ? 3 : 0] = v.w; }
            else ar[0] = As[kk][ty];
            if (NJ == 4) { float4 v = *reinterpret_cast<const float4*>(&Bs[kk][tx * 4]); br[0] = v.x; br[NJ > 1 ? 1 : 0] = v.y; br[NJ > 2 ? 2 : 0] = v.z; br[NJ > 3 ? 3 : 0] = v.w; }
            else br[0] = Bs[kk][tx];
#pragma unroll
            for (int ii = 0; ii < MI; ++ii)
#pragma unroll
                for (int jj = 0; jj < NJ; ++jj) acc[ii][jj] = fmaf(ar[ii], br[jj], acc[ii][jj]);
        }
        __syncthreads();
    }
    if (p.ksplit > 1) {   // raw partial sums; alpha/bias/beta are applied by the reduction kernel
        float* P = p.partial + (size_t(bz) * p.ksplit + ks) * size_t(p.m) * p.n;
#pragma unroll
        for (int i = 0; i < MI; ++i) {
            int gm = m0 + ty * MI + i;
            if (gm >= p.m) continue;
#pragma unroll
            for (int j = 0; j < NJ; ++j) {
                int gn = n0 + tx * NJ + j;
                if (gn < p.n) P[size_t(gm) * p.n + gn] = acc[i][j];
            }
        }
        return;
    }
#pragma unroll
    for (int i = 0; i < MI; ++i) {
        int gm = m0 + ty * MI + i;
        if (gm >= p.m) continue;
#pragma unroll
        for (int j = 0; j < NJ; ++j) {
            int gn = n0 + tx * NJ + j;
            if (gn >= p.n) continue;
            float v = acc[i][j] * p.alpha;
            if (p.bias) v += p.bias[gn];
            int64_t off = int64_t(gm) * p.ldc + gn;
            if (p.beta != 0.f) v += p.beta * C[off];
            if (p.addend) v += p.addend[bz * p.stride_c + int64_t(gm) * p.ld_addend + gn];
            if (p.gelu_grad_x) v *= gelu_tanh_grad_f(p.gelu_grad_x[bz * p.stride_c + int64_t(gm) * p.ld_gelu_grad_x + gn]);
            C[off] = v;
            if (G) G[off] = gelu_tanh_f(v);
        }
    }
}

// Fixed-order reduction of the split-K partials + epilogue (deterministic, unlike atomics).
__global__ void gemm_splitk_reduce_kernel(GemmParams p) {
    pdl_sync();
    const int64_t total = int64_t(p.batch) * p.m * p.n;
    for (int64_t idx = blockIdx.x * int64_t(blockDim.x) + threadIdx.x; idx < total; idx += int64_t(gridDim.x) * blockDim.x) {
        const int gn = int(idx % p.n);
        const int64_t t = idx / p.n;
        const int gm = int(t % p.m);
        const int bz = int(t / p.m);
        const float* P = p.partial + size_t(bz) * p.ksplit * size_t(p.m) * p.n + size_t(gm) * p.n + gn;
        float s = 0.f;
        for (int k = 0; k < p.ksplit; ++k) s += P[size_t(k) * p.m * p.n];
        float v = s * p.alpha;
        if (p.bias) v += p.bias[gn];
        const int64_t off = bz * p.stride_c + int64_t(gm) * p.ldc + gn;
        if (p.beta != 0.f) v += p.beta * p.c[off];
        if (p.addend) v += p.addend[bz * p.stride_c + int64_t(gm) * p.ld_addend + gn];
        if (p.gelu_grad_x) v *= gelu_tanh_grad_f(p.gelu_grad_x[bz * p.stride_c + int64_t(gm) * p.ld_gelu_grad_x + gn]);
        p.c[off] = v;
        if (p.gelu_out) p.gelu_out[off] = gelu_tanh_f(v);
    }
}

int launch_gemm_simt(const GemmParams& p, cudaStream_t st) {
    const int bm = p.m <= 16 ? 16 : SBM, bn = p.n <= 16 ? 16 : SBN;
    dim3 grid(unsigned(ceil_div(p.n, bn)), unsigned(ceil_div(p.m, bm)), unsigned(p.batch * p.ksplit));
    if (grid.y > 65535 || grid.z > 65535) return set_error(1, "mft_gemm(simt): grid too large");
    if (bm == 16 && bn == 16) pdl_launch(gemm_simt_kernel<1, 1>, grid, 256, 0, st)(p);
    else if (bm == 16) pdl_launch(gemm_simt_kernel<1, 4>, grid, 256, 0, st)(p);
    else if (bn == 16) pdl_launch(gemm_simt_kernel<4, 1>, grid, 256, 0, st)(p);
    else pdl_launch(gemm_simt_kernel<4, 4>, grid, 256, 0, st)(p);
    add_launches(1);
    if (p.ksplit > 1) {
        pdl_launch(gemm_splitk_reduce_kernel, stream_grid(int64_t(p.batch) * p.m * p.n, 256), 256, 0, st)(p);
        add_launches(1);
    }
    cudaError_t e = cudaGetLastError();
    if (e != cudaSuccess) return set_error(2, "mft_gemm(simt): %s", cudaGetErrorString(e));
    return 0;
}

}  // namespace mft
