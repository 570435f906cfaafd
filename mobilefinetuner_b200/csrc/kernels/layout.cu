// Layout / gather kernels: transpose, head permute, column sums, embedding gather, RoPE, GQA repeat.
#include <cuda_bf16.h>
#include "common.cuh"

namespace mft {

// 32x32 shared-memory tile transpose (coalesced on both sides, padded against bank conflicts).
__global__ void transpose2d_kernel(const float* __restrict__ in, int64_t ldi, float* __restrict__ out, int64_t ldo,
                                   int64_t rows, int64_t cols, int64_t stride_in, int64_t stride_out) {
    pdl_sync();
    __shared__ float tile[32][33];
    in += blockIdx.z * stride_in;
    out += blockIdx.z * stride_out;
    const int64_t c0 = int64_t(blockIdx.x) * 32, r0 = int64_t(blockIdx.y) * 32;
    for (int j = threadIdx.y; j < 32; j += blockDim.y) {
        int64_t r = r0 + j, c = c0 + threadIdx.x;
        if (r < rows && c < cols) tile[j][threadIdx.x] = in[r * ldi + c];
    }
    __syncthreads();
    for (int j = threadIdx.y; j < 32; j += blockDim.y) {
        int64_t c = c0 + j, r = r0 + threadIdx.x;
        if (r < rows && c < cols) out[c * ldo + r] = tile[threadIdx.x][j];
    }
}

// fp32 -> bf16 (round to nearest even), row pitches in elements; 4 elements per thread
__global__ void cast_bf16_kernel(const float* __restrict__ in, int64_t ldi, __nv_bfloat16* __restrict__ out, int64_t ldo,
                                 int64_t rows, int64_t cols) {
    pdl_sync();
    const int64_t c4 = (cols + 3) >> 2;
    const int64_t n = rows * c4;
    const bool vec = ((ldi & 3) == 0) && ((ldo & 3) == 0) && ((reinterpret_cast<uintptr_t>(in) & 15u) == 0) &&
                     ((reinterpret_cast<uintptr_t>(out) & 7u) == 0);
    for (int64_t idx = blockIdx.x * int64_t(blockDim.x) + threadIdx.x; idx < n; idx += int64_t(gridDim.x) * blockDim.x) {
        const int64_t r = idx / c4, c = (idx % c4) << 2;
        const float* src = in + r * ldi + c;
        __nv_bfloat16* dst = out + r * ldo + c;
        if (vec && c + 4 <= cols) {
            float4 v = ld_stream(reinterpret_cast<const float4*>(src));
            __nv_bfloat162 lo = __floats2bfloat162_rn(v.x, v.y), hi = __floats2bfloat162_rn(v.z, v.w);
            uint2 pk;
            pk.x = *reinterpret_cast<uint32_t*>(&lo);
            pk.y = *reinterpret_cast<uint32_t*>(&hi);
            *reinterpret_cast<uint2*>(dst) = pk;
        } else {
            for (int e = 0; e < 4 && c + e < cols; ++e) dst[e] = __float2bfloat16_rn(src[e]);
        }
    }
}

// out[c, r] (bf16, ldo) = in[r, c] (fp32, ldi): K-major bf16 shadow of a weight in one pass
__global__ void transpose_cast_bf16_kernel(const float* __restrict__ in, int64_t ldi, __nv_bfloat16* __restrict__ out, int64_t ldo,
                                           int64_t rows, int64_t cols) {
    pdl_sync();
    __shared__ float tile[32][33];
    const int64_t c0 = int64_t(blockIdx.x) * 32, r0 = int64_t(blockIdx.y) * 32;
    for (int j = threadIdx.y; j < 32; j += blockDim.y) {
        int64_t r = r0 + j, c = c0 + threadIdx.x;
        if (r < rows && c < cols) tile[j][threadIdx.x] = in[r * ldi + c];
    }
    __syncthreads();
    for (int j = threadIdx.y; j < 32; j += blockDim.y) {
        int64_t c = c0 + j, r = r0 + threadIdx.x;
        if (r < rows && c < cols) out[c * ldo + r] = __float2bfloat16_rn(tile[threadIdx.x][j]);
    }
}

// out[b, j, i, :] = in[b, i, j, :]  — rows of d3 floats are moved whole (coalesced both sides).
__global__ void permute0213_kernel(const float* __restrict__ in, float* __restrict__ out, int64_t d0, int64_t d1,
                                   int64_t d2, int64_t d3) {
    pdl_sync();
    const int64_t n = d0 * d1 * d2 * d3;
    for (int64_t idx = blockIdx.x * int64_t(blockDim.x) + threadIdx.x; idx < n; idx += int64_t(gridDim.x) * blockDim.x) {
        int64_t d = idx % d3;
        int64_t t = idx / d3;
        int64_t i = t % d1;  // output layout [d0, d2, d1, d3]
        t /= d1;
        int64_t j = t % d2;
        int64_t b = t / d2;
        out[idx] = in[((b * d1 + i) * d2 + j) * d3 + d];
    }
}

// Column sums, stage 1: each CTA owns a 32-column strip and a slab of rows, 32x8 threads.
constexpr int kColsumRowsPerCta = 256;
__global__ void colsum_partial_kernel(const float* __restrict__ g, int64_t ldg, float* __restrict__ partial,
                                      int64_t rows, int64_t n) {
    pdl_sync();
    __shared__ float red[8][33];
    const int64_t c = int64_t(blockIdx.x) * 32 + threadIdx.x;
    const int64_t r0 = int64_t(blockIdx.y) * kColsumRowsPerCta;
    int64_t r1 = r0 + kColsumRowsPerCta;
    if (r1 > rows) r1 = rows;
    float acc = 0.f;
    if (c < n)
        for (int64_t r = r0 + threadIdx.y; r < r1; r += 8) acc += g[r * ldg + c];
    red[threadIdx.y][threadIdx.x] = acc;
    __syncthreads();
    if (threadIdx.y == 0 && c < n) {
        float s = 0.f;
#pragma unroll
        for (int j = 0; j < 8; ++j) s += red[j][threadIdx.x];
        partial[int64_t(blockIdx.y) * n + c] = s;
    }
}
// stage 2: fixed-order sum of the slabs (deterministic).
__global__ void colsum_final_kernel(const float* __restrict__ partial, float* __restrict__ out, int64_t slabs, int64_t n) {
    pdl_sync();
    int64_t c = blockIdx.x * int64_t(blockDim.x) + threadIdx.x;
    if (c >= n) return;
    float s = 0.f;
    for (int64_t j = 0; j < slabs; ++j) s += partial[j * n + c];
    out[c] = s;
}

// Embedding gather: one warp per token row, float4 when C % 4 == 0.
__global__ void embedding_kernel(const int32_t* __restrict__ ids, const float* __restrict__ wte,
                                 const float* __restrict__ wpe, float* __restrict__ x, int64_t tokens, int64_t S,
                                 int64_t C, int64_t vocab, float scale, int32_t* status) {
    pdl_sync();
    const int lane = threadIdx.x & 31;
    const int64_t warps_per_grid = (int64_t(gridDim.x) * blockDim.x) >> 5;
    for (int64_t t = (blockIdx.x * int64_t(blockDim.x) + threadIdx.x) >> 5; t < tokens; t += warps_per_grid) {
        int32_t id = ids[t];
        float* dst = x + t * C;
        if (id < 0 || id >= vocab) {
            if (status && lane == 0) *status = 1;
            for (int64_t i = lane; i < C; i += 32) dst[i] = 0.f;
            continue;
        }
        const float* src = wte + int64_t(id) * C;
        const float* pos = wpe ? wpe + (t % S) * C : nullptr;
        if ((C & 3) == 0) {
            const float4* s4 = reinterpret_cast<const float4*>(src);
            const float4* p4 = reinterpret_cast<const float4*>(pos);
            float4* d4 = reinterpret_cast<float4*>(dst);
            for (int64_t i = lane; i < (C >> 2); i += 32) {
                float4 v = __ldg(s4 + i);
                if (pos) { float4 p = __ldg(p4 + i); v.x += p.x; v.y += p.y; v.z += p.z; v.w += p.w; }
                else { v.x *= scale; v.y *= scale; v.z *= scale; v.w *= scale; }
                d4[i] = v;
            }
        } else {
            for (int64_t i = lane; i < C; i += 32) dst[i] = pos ? src[i] + pos[i] : src[i] * scale;
        }
    }
}

// RoPE on interleaved pairs. One thread per pair; float2 accesses.
__global__ void rope_kernel(const float* __restrict__ x, float* __restrict__ y, int64_t BH, int64_t S, int64_t Hd,
                            float theta, int inverse) {
    pdl_sync();
    const int64_t half = Hd >> 1;
    const int64_t n = BH * S * half;
    for (int64_t idx = blockIdx.x * int64_t(blockDim.x) + threadIdx.x; idx < n; idx += int64_t(gridDim.x) * blockDim.x) {
        int64_t d = idx % half;
        int64_t pos = (idx / half) % S;
        // reference: freq = 1/pow(theta, 2d/Hd); angle = pos*freq (float) — core/ops.cpp:2183-2187
        float freq = 1.0f / powf(theta, 2.0f * float(d) / float(Hd));
        float angle = float(pos) * freq;
        float c = cosf(angle), s = sinf(angle);
        if (inverse) s = -s;
        float2 v = reinterpret_cast<const float2*>(x)[idx];
        float2 r;
        r.x = v.x * c - v.y * s;
        r.y = v.x * s + v.y * c;
        reinterpret_cast<float2*>(y)[idx] = r;
    }
}

__global__ void repeat_kv_fwd_kernel(const float* __restrict__ x, float* __restrict__ y, int64_t B, int64_t KV,
                                     int64_t rep, int64_t SHd) {
    pdl_sync();
    const int64_t n = B * KV * rep * SHd;
    for (int64_t idx = blockIdx.x * int64_t(blockDim.x) + threadIdx.x; idx < n; idx += int64_t(gridDim.x) * blockDim.x) {
        int64_t e = idx % SHd;
        int64_t h = (idx / SHd) % (KV * rep);
        int64_t b = idx / (SHd * KV * rep);
        y[idx] = x[(b * KV + h / rep) * SHd + e];
    }
}
__global__ void repeat_kv_bwd_kernel(const float* __restrict__ gy, float* __restrict__ gx, int64_t B, int64_t KV,
                                     int64_t rep, int64_t SHd) {
    pdl_sync();
    const int64_t n = B * KV * SHd;
    for (int64_t idx = blockIdx.x * int64_t(blockDim.x) + threadIdx.x; idx < n; idx += int64_t(gridDim.x) * blockDim.x) {
        int64_t e = idx % SHd;
        int64_t kv = (idx / SHd) % KV;
        int64_t b = idx / (SHd * KV);
        float s = 0.f;
        for (int64_t r = 0; r < rep; ++r) s += gy[((b * KV + kv) * rep + r) * SHd + e];
        gx[idx] = s;
    }
}

}  // namespace mft

using namespace mft;

extern "C" int mft_transpose2d(const float* in, int64_t ldi, float* out, int64_t ldo, int64_t rows, int64_t cols,
                               int64_t batch, int64_t stride_in, int64_t stride_out, mft_stream_t stream) {
    if (rows == 0 || cols == 0 || batch == 0) return 0;
    if (!in || !out) return set_error(1, "mft_transpose2d: null pointer");
    if (batch > 65535) return set_error(1, "mft_transpose2d: batch %lld too large", (long long)batch);
    dim3 grid(unsigned(ceil_div(cols, 32)), unsigned(ceil_div(rows, 32)), unsigned(batch));
    if (grid.y > 65535) return set_error(1, "mft_transpose2d: too many rows");
    pdl_launch(transpose2d_kernel, grid, dim3(32, 8), 0, as_stream(stream))(in, ldi, out, ldo, rows, cols, stride_in, stride_out);
    MFT_CHECK_LAUNCH("mft_transpose2d");
    return 0;
}

extern "C" int mft_permute0213(const float* in, float* out, int64_t d0, int64_t d1, int64_t d2, int64_t d3,
                               mft_stream_t stream) {
    int64_t n = d0 * d1 * d2 * d3;
    if (n == 0) return 0;
    if (!in || !out) return set_error(1, "mft_permute0213: null pointer");
    pdl_launch(permute0213_kernel, stream_grid(n, 256), 256, 0, as_stream(stream))(in, out, d0, d1, d2, d3);
    MFT_CHECK_LAUNCH("mft_permute0213");
    return 0;
}

extern "C" size_t mft_colsum_workspace_bytes(int64_t rows, int64_t n) {
    return size_t(ceil_div(rows, kColsumRowsPerCta)) * size_t(n) * sizeof(float);
}

extern "C" int mft_colsum(const float* g, int64_t ldg, float* out, int64_t rows, int64_t n, void* workspace,
                          size_t workspace_bytes, mft_stream_t stream) {
    if (n == 0) return 0;
    if (!g || !out) return set_error(1, "mft_colsum: null pointer");
    if (rows == 0) return mft_fill(out, n, 0.f, stream);
    const int64_t slabs = ceil_div(rows, kColsumRowsPerCta);
    if (workspace_bytes < mft_colsum_workspace_bytes(rows, n) || !workspace)
        return set_error(1, "mft_colsum: workspace too small (%zu < %zu)", workspace_bytes, mft_colsum_workspace_bytes(rows, n));
    if (slabs > 65535) return set_error(1, "mft_colsum: too many rows");
    float* partial = static_cast<float*>(workspace);
    pdl_launch(colsum_partial_kernel, dim3(unsigned(ceil_div(n, 32)), unsigned(slabs)), dim3(32, 8), 0, as_stream(stream))(g, ldg, partial, rows, n);
    pdl_launch(colsum_final_kernel, unsigned(ceil_div(n, 256)), 256, 0, as_stream(stream))(partial, out, slabs, n);
    add_launches(1);
    MFT_CHECK_LAUNCH("mft_colsum");
    return 0;
}

extern "C" int mft_embedding_fwd(const int32_t* ids, const float* wte, const float* wpe, float* x, int64_t tokens,
                                 int64_t S, int64_t C, int64_t vocab, float scale, int32_t* status, mft_stream_t stream) {
    if (tokens == 0 || C == 0) return 0;
    if (!ids || !wte || !x) return set_error(1, "mft_embedding_fwd: null pointer");
    if ((C & 3) == 0 && !(aligned16(wte) && aligned16(x) && (!wpe || aligned16(wpe))))
        return set_error(1, "mft_embedding_fwd: tables must be 16-byte aligned");
    int blocks = stream_grid(tokens * 32, 256);
    pdl_launch(embedding_kernel, blocks, 256, 0, as_stream(stream))(ids, wte, wpe, x, tokens, S, C, vocab, scale, status);
    MFT_CHECK_LAUNCH("mft_embedding_fwd");
    return 0;
}

extern "C" int mft_rope(const float* x, float* y, int64_t BH, int64_t S, int64_t Hd, float theta, int inverse,
                        mft_stream_t stream) {
    if (BH * S * Hd == 0) return 0;
    if (!x || !y) return set_error(1, "mft_rope: null pointer");
    if (Hd & 1) return set_error(1, "mft_rope: head_dim must be even");
    pdl_launch(rope_kernel, stream_grid(BH * S * (Hd / 2), 256), 256, 0, as_stream(stream))(x, y, BH, S, Hd, theta, inverse);
    MFT_CHECK_LAUNCH("mft_rope");
    return 0;
}

extern "C" int mft_repeat_kv_fwd(const float* x, float* y, int64_t B, int64_t KV, int64_t rep, int64_t S, int64_t Hd,
                                 mft_stream_t stream) {
    int64_t n = B * KV * rep * S * Hd;
    if (n == 0) return 0;
    if (!x || !y) return set_error(1, "mft_repeat_kv_fwd: null pointer");
    pdl_launch(repeat_kv_fwd_kernel, stream_grid(n, 256), 256, 0, as_stream(stream))(x, y, B, KV, rep, S * Hd);
    MFT_CHECK_LAUNCH("mft_repeat_kv_fwd");
    return 0;
}

extern "C" int mft_repeat_kv_bwd(const float* gy, float* gx, int64_t B, int64_t KV, int64_t rep, int64_t S, int64_t Hd,
                                 mft_stream_t stream) {
    int64_t n = B * KV * S * Hd;
    if (n == 0) return 0;
    if (!gy || !gx) return set_error(1, "mft_repeat_kv_bwd: null pointer");
    pdl_launch(repeat_kv_bwd_kernel, stream_grid(n, 256), 256, 0, as_stream(stream))(gy, gx, B, KV, rep, S * Hd);
    MFT_CHECK_LAUNCH("mft_repeat_kv_bwd");
    return 0;
}

extern "C" int mft_cast_bf16(const float* in, int64_t ldi, void* out, int64_t ldo, int64_t rows, int64_t cols, mft_stream_t stream) {
    if (rows == 0 || cols == 0) return 0;
    if (!in || !out) return set_error(1, "mft_cast_bf16: null pointer");
    pdl_launch(cast_bf16_kernel, stream_grid(rows * ((cols + 3) / 4), 256), 256, 0, as_stream(stream))(in, ldi, static_cast<__nv_bfloat16*>(out), ldo, rows, cols);
    MFT_CHECK_LAUNCH("mft_cast_bf16");
    return 0;
}

extern "C" int mft_transpose_cast_bf16(const float* in, int64_t ldi, void* out, int64_t ldo, int64_t rows, int64_t cols, mft_stream_t stream) {
    if (rows == 0 || cols == 0) return 0;
    if (!in || !out) return set_error(1, "mft_transpose_cast_bf16: null pointer");
    dim3 grid(unsigned(ceil_div(cols, 32)), unsigned(ceil_div(rows, 32)));
    if (grid.y > 65535) return set_error(1, "mft_transpose_cast_bf16: too many rows");
    pdl_launch(transpose_cast_bf16_kernel, grid, dim3(32, 8), 0, as_stream(stream))(in, ldi, static_cast<__nv_bfloat16*>(out), ldo, rows, cols);
    MFT_CHECK_LAUNCH("mft_transpose_cast_bf16");
    return 0;
}

// ---------------------------------------------------------------- rare layouts / data preparation (device-side, no host loops)
namespace mft {
struct PermuteNd { int nd; int64_t out_shape[8]; int64_t src_stride[8]; };   // src_stride[i] = input stride of output axis i
__global__ void permute_nd_kernel(const float* __restrict__ in, float* __restrict__ out, int64_t n, PermuteNd p) {
    pdl_sync();
    for (int64_t idx = blockIdx.x * int64_t(blockDim.x) + threadIdx.x; idx < n; idx += int64_t(gridDim.x) * blockDim.x) {
        int64_t rem = idx, off = 0;
        for (int i = p.nd - 1; i >= 0; --i) { off += (rem % p.out_shape[i]) * p.src_stride[i]; rem /= p.out_shape[i]; }
        out[idx] = in[off];
    }
}
__global__ void cast_f32_i32_kernel(const float* __restrict__ in, int32_t* __restrict__ out, int64_t n) {
    pdl_sync();
    for (int64_t i = blockIdx.x * int64_t(blockDim.x) + threadIdx.x; i < n; i += int64_t(gridDim.x) * blockDim.x) out[i] = int32_t(in[i]);
}
__global__ void cast_i32_f32_kernel(const int32_t* __restrict__ in, float* __restrict__ out, int64_t n) {
    pdl_sync();
    for (int64_t i = blockIdx.x * int64_t(blockDim.x) + threadIdx.x; i < n; i += int64_t(gridDim.x) * blockDim.x) out[i] = float(in[i]);
}
__global__ void tril_ones_kernel(float* __restrict__ out, int64_t S) {
    pdl_sync();
    const int64_t n = S * S;
    for (int64_t i = blockIdx.x * int64_t(blockDim.x) + threadIdx.x; i < n; i += int64_t(gridDim.x) * blockDim.x)
        out[i] = (i % S) <= (i / S) ? 1.0f : 0.0f;
}
// out[N, C] = 1 at (i, target[i]) when the class is in range, 0 elsewhere; targets int32 or float-coded
__global__ void one_hot_kernel(const void* __restrict__ target, int target_is_i32, float* __restrict__ out, int64_t N, int64_t C) {
    pdl_sync();
    const int64_t n = N * C;
    for (int64_t i = blockIdx.x * int64_t(blockDim.x) + threadIdx.x; i < n; i += int64_t(gridDim.x) * blockDim.x) {
        const int64_t r = i / C, c = i % C;
        const int64_t t = target_is_i32 ? int64_t(static_cast<const int32_t*>(target)[r]) : int64_t(static_cast<const float*>(target)[r]);
        out[i] = t == c ? 1.0f : 0.0f;
    }
}
}  // namespace mft

extern "C" int mft_permute_nd(const float* in, float* out, int32_t ndim, const int64_t* out_shape, const int64_t* src_stride, mft_stream_t stream) {
    if (ndim < 1 || ndim > 8) return set_error(1, "mft_permute_nd: 1..8 dimensions");
    if (!in || !out || !out_shape || !src_stride) return set_error(1, "mft_permute_nd: null pointer");
    mft::PermuteNd p;
    p.nd = ndim;
    int64_t n = 1;
    for (int i = 0; i < ndim; ++i) { p.out_shape[i] = out_shape[i]; p.src_stride[i] = src_stride[i]; n *= out_shape[i]; }
    if (n == 0) return 0;
    pdl_launch(mft::permute_nd_kernel, stream_grid(n, 256), 256, 0, as_stream(stream))(in, out, n, p);
    MFT_CHECK_LAUNCH("mft_permute_nd");
    return 0;
}
extern "C" int mft_cast_i32(const void* in, void* out, int64_t n, int to_i32, mft_stream_t stream) {
    if (n == 0) return 0;
    if (!in || !out) return set_error(1, "mft_cast_i32: null pointer");
    if (to_i32) pdl_launch(mft::cast_f32_i32_kernel, stream_grid(n, 256), 256, 0, as_stream(stream))(static_cast<const float*>(in), static_cast<int32_t*>(out), n);
    else pdl_launch(mft::cast_i32_f32_kernel, stream_grid(n, 256), 256, 0, as_stream(stream))(static_cast<const int32_t*>(in), static_cast<float*>(out), n);
    MFT_CHECK_LAUNCH("mft_cast_i32");
    return 0;
}
extern "C" int mft_tril_ones(float* out, int64_t S, mft_stream_t stream) {
    if (S == 0) return 0;
    if (!out) return set_error(1, "mft_tril_ones: null pointer");
    pdl_launch(mft::tril_ones_kernel, stream_grid(S * S, 256), 256, 0, as_stream(stream))(out, S);
    MFT_CHECK_LAUNCH("mft_tril_ones");
    return 0;
}
extern "C" int mft_one_hot(const void* target, int target_is_i32, float* out, int64_t N, int64_t C, mft_stream_t stream) {
    if (N * C == 0) return 0;
    if (!target || !out) return set_error(1, "mft_one_hot: null pointer");
    pdl_launch(mft::one_hot_kernel, stream_grid(N * C, 256), 256, 0, as_stream(stream))(target, target_is_i32, out, N, C);
    MFT_CHECK_LAUNCH("mft_one_hot");
    return 0;
}
