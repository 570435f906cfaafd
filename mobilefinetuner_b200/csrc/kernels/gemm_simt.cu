// SIMT fp32 GEMM: the general-layout path (any transposes, any alignment, tiny or skinny shapes) and the
// bit-faithful fp32 comparison path for the tensor-core kernel. 64x64 CTA tile, BK = 16, 256 threads,
// 4x4 register micro-tile, global loads mapped so that the contiguous dimension is the fast one.
#include "common.cuh"
#include "gemm_internal.cuh"

namespace mft {

constexpr int SBM = 64, SBN = 64, SBK = 16;

// Loads a (rows = MN extent 64) x (SBK) tile of an operand into smem as s[k][mn].
// `contig_k` = the operand's contiguous dimension is K (A with trans_a=0 / B with trans_b=1).
__device__ __forceinline__ void load_tile(float (*s)[SBM + 4], const float* __restrict__ p, int64_t ld, bool contig_k,
                                          int mn0, int k0, int mn_max, int k_max) {
#pragma unroll
    for (int j = 0; j < 4; ++j) {
        int idx = threadIdx.x + 256 * j;
        int kk, mm;
        if (contig_k) { kk = idx % SBK; mm = idx / SBK; } else { mm = idx % SBM; kk = idx / SBM; }
        int gm = mn0 + mm, gk = k0 + kk;
        float v = 0.f;
        if (gm < mn_max && gk < k_max) v = contig_k ? p[int64_t(gm) * ld + gk] : p[int64_t(gk) * ld + gm];
        s[kk][mm] = v;
    }
}

__global__ void __launch_bounds__(256)
gemm_simt_kernel(GemmParams p) {
    __shared__ float As[SBK][SBM + 4];
    __shared__ float Bs[SBK][SBN + 4];
    const int bz = blockIdx.z / p.ksplit, ks = blockIdx.z % p.ksplit;
    const float* A = p.a + bz * p.stride_a;
    const float* B = p.b + bz * p.stride_b;
    float* C = p.c + bz * p.stride_c;
    float* G = p.gelu_out ? p.gelu_out + bz * p.stride_c : nullptr;
    const int m0 = blockIdx.y * SBM, n0 = blockIdx.x * SBN;
    const int tx = threadIdx.x % 16, ty = threadIdx.x / 16;
    float acc[4][4];
#pragma unroll
    for (int i = 0; i < 4; ++i)
#pragma unroll
        for (int j = 0; j < 4; ++j) acc[i][j] = 0.f;

    for (int seg = 0; seg < 2; ++seg) {
        const float* a = seg == 0 ? A : p.a2;
        const float* b = seg == 0 ? B : p.b2;
        const int K = seg == 0 ? p.k : p.k2;
        if (K == 0 || !a || !b) continue;
        const int64_t lda = seg == 0 ? p.lda : p.lda2, ldb = seg == 0 ? p.ldb : p.ldb2;
        const bool a_ck = seg == 0 ? (p.trans_a == 0) : true;
        const bool b_ck = seg == 0 ? (p.trans_b != 0) : true;
        // split-K: this CTA owns k-blocks [kb0, kb1) of segment 1 (segment 2 is never split)
        int k_begin = 0, k_end = K;
        if (p.ksplit > 1) {
            const int kblocks = (K + SBK - 1) / SBK, per = (kblocks + p.ksplit - 1) / p.ksplit;
            k_begin = ks * per * SBK;
            k_end = min(K, (ks + 1) * per * SBK);
        }
        for (int k0 = k_begin; k0 < k_end; k0 += SBK) {
            load_tile(As, a, lda, a_ck, m0, k0, p.m, K);
            load_tile(Bs, b, ldb, b_ck, n0, k0, p.n, K);
            __syncthreads();
#pragma unroll
            for (int kk = 0; kk < SBK; ++kk) {
                float4 av = *reinterpret_cast<const float4*>(&As[kk][ty * 4]);
                float4 bv = *reinterpret_cast<const float4*>(&Bs[kk][tx * 4]);
                float ar[4] = {av.x, av.y, av.z, av.w}, br[4] = {bv.x, bv.y, bv.z, bv.w};
#pragma unroll
                for (int i = 0; i < 4; ++i)
#pragma unroll
                    for (int j = 0; j < 4; ++j) acc[i][j] = fmaf(ar[i], br[j], acc[i][j]);
            }
            __syncthreads();
        }
    }
    if (p.ksplit > 1) {   // raw partial sums; alpha/bias/beta are applied by the reduction kernel
        float* P = p.partial + (size_t(bz) * p.ksplit + ks) * size_t(p.m) * p.n;
#pragma unroll
        for (int i = 0; i < 4; ++i) {
            int gm = m0 + ty * 4 + i;
            if (gm >= p.m) continue;
#pragma unroll
            for (int j = 0; j < 4; ++j) {
                int gn = n0 + tx * 4 + j;
                if (gn < p.n) P[size_t(gm) * p.n + gn] = acc[i][j];
            }
        }
        return;
    }
#pragma unroll
    for (int i = 0; i < 4; ++i) {
        int gm = m0 + ty * 4 + i;
        if (gm >= p.m) continue;
#pragma unroll
        for (int j = 0; j < 4; ++j) {
            int gn = n0 + tx * 4 + j;
            if (gn >= p.n) continue;
            float v = acc[i][j] * p.alpha;
            if (p.bias) v += p.bias[gn];
            int64_t off = int64_t(gm) * p.ldc + gn;
            if (p.beta != 0.f) v += p.beta * C[off];
            C[off] = v;
            if (G) G[off] = gelu_tanh_f(v);
        }
    }
}

// Fixed-order reduction of the split-K partials + epilogue (deterministic, unlike atomics).
__global__ void gemm_splitk_reduce_kernel(GemmParams p) {
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
        p.c[off] = v;
        if (p.gelu_out) p.gelu_out[off] = gelu_tanh_f(v);
    }
}

int launch_gemm_simt(const GemmParams& p, cudaStream_t st) {
    dim3 grid(unsigned(ceil_div(p.n, SBN)), unsigned(ceil_div(p.m, SBM)), unsigned(p.batch * p.ksplit));
    if (grid.y > 65535 || grid.z > 65535) return set_error(1, "mft_gemm(simt): grid too large");
    gemm_simt_kernel<<<grid, 256, 0, st>>>(p);
    add_launches(1);
    if (p.ksplit > 1) {
        gemm_splitk_reduce_kernel<<<stream_grid(int64_t(p.batch) * p.m * p.n, 256), 256, 0, st>>>(p);
        add_launches(1);
    }
    cudaError_t e = cudaGetLastError();
    if (e != cudaSuccess) return set_error(2, "mft_gemm(simt): %s", cudaGetErrorString(e));
    return 0;
}

}  // namespace mft
