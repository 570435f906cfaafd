// Causal self-attention forward on the 5th-generation tensor cores (tcgen05.mma, accumulators in TMEM, operands by TMA):
// GPT-2 semantics (core/memory_efficient_attention.cpp:92-162), head_dim 64, bf16 operands, fp32 scores / softmax / output.
//
// One CTA = 128 query rows of one (batch, head); keys stream through shared memory in tiles of 64.
//   warps 0-7 : softmax, TWO threads per query row. Warps w and w + 4 own TMEM lanes 32 (w & 3) .. +31 (a warp reaches the
//               lane quarter warp_id % 4); the first takes keys / output columns 0-31 of every tile, the second 32-63. Each
//               thread reads its 32 scores with tcgen05.ld, the pair agrees on the tile's row max through shared memory and
//               a 64-thread named barrier, and each keeps its own partial row sum and its 32 output accumulators in
//               registers, writing its half of the P row (bf16) into the 128B-swizzled K-major tile of the second product.
//               Half the registers and half the dependent chain per thread: twice the warps per scheduler hide the
//               tcgen05.ld / MUFU / barrier latencies that bounded the one-thread-per-row form (52 % issue-active).
//   warp 8    : TMA producer (lane 0): Q once, then K and V tiles into their two-stage rings (cp.async.bulk.tensor, SWIZZLE_128B).
//   warp 9    : TMEM allocator + tcgen05.mma issuer (lane 0):  S = Q K^T  (M 128, N 64, K 64: both operands K-major)  and
//               O_t = P V (M 128, N 64, K 64: V is read MN-major straight from its [key][head_dim] tile — no transpose).
// The score product runs two tiles ahead of the softmax (S is double-buffered in TMEM, K and V have separate two-stage rings so
// that K(t+2) can be fetched as soon as S(t) has read its stage), and P is double-buffered in shared memory, so the softmax chain
// per tile is S ready -> exp -> P stored with both products and the TMA off it; two CTAs per SM (84 KB of shared memory, 256
// TMEM columns each).
// O accumulates in TMEM over all key tiles (the tensor core adds P(t) V to it) and is read once, at the end: TMEM reads run at
// 64 B/clk per SM, and copying O out every tile doubled them. It is scaled by a max that follows the running row max lazily
// (see kTaRescaleStep); the rare rescale is a TMEM load / multiply / store by the owning threads between two products.
// q/k/v come from ONE bf16 [B, S, 3C] buffer (the qkv projection's bf16 output) through one tensor map.
#include <cuda.h>
#include <cstdlib>
#include <cstring>
#include <unordered_map>
#include "common.cuh"
#include "tc_common.cuh"

namespace mft {

constexpr int kTaBQ = 128, kTaBK = 64, kTaHD = 64;
constexpr int kTaThreads = 320;
constexpr int kTaSoftmaxWarps = 8, kTaProducerWarp = 8, kTaMmaWarp = 9;
constexpr int kTaTileBytes = kTaBK * kTaHD * 2;         // 8 KB: one K or V tile
constexpr int kTaQBytes = kTaBQ * kTaHD * 2;            // 16 KB
constexpr int kTaPBytes = kTaBQ * kTaBK * 2;            // 16 KB: one P tile (two of them: tile t+1 is written while P(t) V is read)
constexpr int kTaStages = 2;
constexpr int kTaXchBytes = 2 * 2 * kTaBQ * 4;          // row-max exchange between the two threads of a row: [buffer][half][row]
constexpr int kTaSmem = kTaQBytes + kTaStages * 2 * kTaTileBytes + 2 * kTaPBytes + 1024 /*align*/ + 128 /*barriers*/ + kTaXchBytes;
constexpr int kTaTmemCols = 256;                        // S of even / odd tiles: columns [0,64) / [64,128); O (accumulated over the key tiles): [128,192)
constexpr float kTaRescaleStep = 8.0f;                  // log2 units the row max may run ahead of the scaling max before O is rescaled

struct TaArgs {
    float* ctx; __nv_bfloat16* ctx16;
    int H, S;
    int64_t osb, oss;          // ctx element strides (batch, seq); heads are 64 apart
    float scale_log2;          // scale * log2(e)
};

// MN-major, 128B-swizzled B operand: rows = contraction index (keys), 128 bytes (64 bf16 of N) per row, 8-row groups
// 1024 bytes apart (SBO). One UMMA (K = 16) spans two groups.
__device__ __forceinline__ uint64_t make_mnmajor_sw128_desc(uint32_t smem_addr) {
    uint64_t d = 0;
    d |= uint64_t((smem_addr >> 4) & 0x3FFF);
    d |= uint64_t((1024 >> 4) & 0x3FFF) << 16;         // LBO: stride between 64-element blocks along N (one block here)
    d |= uint64_t((1024 >> 4) & 0x3FFF) << 32;         // SBO: stride between 8-row groups along K
    d |= uint64_t(1) << 46;
    d |= uint64_t(2) << 61;                            // SWIZZLE_128B
    return d;
}

__device__ __forceinline__ float ex2_approx(float x) {          // one MUFU.EX2; 2^(-inf) = 0
    float y;
    asm("ex2.approx.ftz.f32 %0, %1;" : "=f"(y) : "f"(x));
    return y;
}

// named barrier 1 + q for the 64 threads (two warps) that share TMEM lane quarter q
__device__ __forceinline__ void pair_sync(int q) {
    switch (q) {
        case 0: asm volatile("bar.sync 1, 64;" ::: "memory"); break;
        case 1: asm volatile("bar.sync 2, 64;" ::: "memory"); break;
        case 2: asm volatile("bar.sync 3, 64;" ::: "memory"); break;
        default: asm volatile("bar.sync 4, 64;" ::: "memory"); break;
    }
}

__global__ void __launch_bounds__(kTaThreads, 2)
attention_tc_fwd_kernel(const __grid_constant__ CUtensorMap map_qkv, TaArgs p, int C) {
    pdl_sync();
    extern __shared__ uint8_t smem_raw[];
    uint8_t* base = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));
    uint8_t* sQ = base;
    uint8_t* sK = sQ + kTaQBytes;                            // [stages][8 KB]
    uint8_t* sV = sK + kTaStages * kTaTileBytes;             // [stages][8 KB]
    uint8_t* sP = sV + kTaStages * kTaTileBytes;             // [2][16 KB]
    uint64_t* bars = reinterpret_cast<uint64_t*>(sP + 2 * kTaPBytes);
    uint64_t* bar_q = bars;                // Q landed
    uint64_t* bar_kfull = bars + 1;        // [2] K tile landed          (K and V have their own rings: K(t) is free again after
    uint64_t* bar_vfull = bars + 3;        // [2] V tile landed           S(t), long before P(t) V releases V(t))
    uint64_t* bar_kempty = bars + 5;       // [2] S(t) has read K stage t & 1
    uint64_t* bar_vempty = bars + 7;       // [2] P(t) V has read V stage t & 1
    uint64_t* bar_s = bars + 9;            // [2] S(t) ready in TMEM buffer t & 1
    uint64_t* bar_p = bars + 11;           // [2] P(t) written to shared memory buffer t & 1 (S(t) copied out, O rescaled if needed)
    uint64_t* bar_pv = bars + 13;          // [2] P(t) V retired
    uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(bars + 15);
    float* xch = reinterpret_cast<float*>(reinterpret_cast<uint8_t*>(bars) + 128);      // [2][2][128]

    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const int qt = gridDim.x - 1 - blockIdx.x;       // heavy query tiles first
    const int b = blockIdx.y / p.H, h = blockIdx.y % p.H;
    const int q0 = qt * kTaBQ;
    const int S = p.S;
    const int last_row = min(q0 + kTaBQ, S) - 1;
    const int n_tiles = last_row / kTaBK + 1;        // causal: keys 0 .. last_row

    if (threadIdx.x == 0) {
        mbar_init(bar_q, 1);
        for (int i = 0; i < 2; ++i) {
            mbar_init(&bar_kfull[i], 1); mbar_init(&bar_vfull[i], 1); mbar_init(&bar_kempty[i], 1); mbar_init(&bar_vempty[i], 1);
            mbar_init(&bar_s[i], 1); mbar_init(&bar_p[i], kTaSoftmaxWarps); mbar_init(&bar_pv[i], 1);
        }
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    }
    if (warp == kTaMmaWarp) {
        asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" :: "r"(smem_u32(tmem_slot)), "r"(uint32_t(kTaTmemCols)) : "memory");
        asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
    }
    tc_fence_before();
    __syncthreads();
    tc_fence_after();
    const uint32_t tmem_base = *tmem_slot;

    if (warp == kTaProducerWarp) {
        // ===================== TMA producer =====================
        if (lane == 0) {
            asm volatile("prefetch.tensormap [%0];" :: "l"(reinterpret_cast<uint64_t>(&map_qkv)) : "memory");
            mbar_arrive_expect_tx(bar_q, kTaQBytes);
            tma_load_3d(sQ, &map_qkv, bar_q, h * kTaHD, q0, b);                       // 64 rows per box: two boxes
            tma_load_3d(sQ + kTaQBytes / 2, &map_qkv, bar_q, h * kTaHD, q0 + 64, b);
            auto load_k = [&](int t) {
                const int st = t & 1;
                mbar_wait(&bar_kempty[st], ((t >> 1) & 1) ^ 1);
                mbar_arrive_expect_tx(&bar_kfull[st], kTaTileBytes);
                tma_load_3d(sK + st * kTaTileBytes, &map_qkv, &bar_kfull[st], C + h * kTaHD, t * kTaBK, b);
            };
            load_k(0);
            if (n_tiles > 1) load_k(1);
            for (int it = 0; it < n_tiles; ++it) {   // K runs two tiles ahead of V: it is asked for as soon as S(it) has released its stage
                const int st = it & 1;
                mbar_wait(&bar_vempty[st], ((it >> 1) & 1) ^ 1);
                mbar_arrive_expect_tx(&bar_vfull[st], kTaTileBytes);
                tma_load_3d(sV + st * kTaTileBytes, &map_qkv, &bar_vfull[st], 2 * C + h * kTaHD, it * kTaBK, b);
                if (it + 2 < n_tiles) load_k(it + 2);
            }
        }
    } else if (warp == kTaMmaWarp) {
        // ===================== MMA issuer =====================
        if (lane == 0) {
            // D = f32, A = B = bf16; S: both K-major, N = 64; O_t: A K-major (P), B MN-major (V), N = 64; M = 128
            constexpr uint32_t idesc_s = (1u << 4) | (1u << 7) | (1u << 10) | (uint32_t(kTaBK >> 3) << 17) | (uint32_t(kTaBQ >> 4) << 24);
            constexpr uint32_t idesc_o = idesc_s | (1u << 16);
            const uint64_t qdesc = make_kmajor_sw128_desc(smem_u32(sQ));
            const uint32_t tO = tmem_base + 128;
            // S(t) = Q K(t)^T into TMEM buffer t & 1; K stage t & 1 is released by the same commit
            auto issue_s = [&](int t) {
                const int sb = t & 1;
                mbar_wait(&bar_kfull[sb], (t >> 1) & 1);
                tc_fence_after();
                const uint64_t kdesc = make_kmajor_sw128_desc(smem_u32(sK + sb * kTaTileBytes));
#pragma unroll
                for (int k = 0; k < kTaHD / 16; ++k) tc_mma_bf16(tmem_base + uint32_t(sb) * 64, qdesc + uint64_t(k * 2), kdesc + uint64_t(k * 2), idesc_s, k ? 1u : 0u);
                tc_commit(&bar_kempty[sb]);
                tc_commit(&bar_s[sb]);
            };
            mbar_wait(bar_q, 0);
            issue_s(0);                              // two score tiles ahead of the softmax: its chain never contains a product
            if (n_tiles > 1) issue_s(1);
            for (int it = 0; it < n_tiles; ++it) {
                const int st = it & 1;
                mbar_wait(&bar_p[st], (it >> 1) & 1);   // P(it) is in its buffer, S(it) has been copied out, O rescaled if this tile asked for it
                mbar_wait(&bar_vfull[st], (it >> 1) & 1);
                tc_fence_after();
                const uint64_t pdesc = make_kmajor_sw128_desc(smem_u32(sP + st * kTaPBytes));
                const uint64_t vdesc = make_mnmajor_sw128_desc(smem_u32(sV + st * kTaTileBytes));
#pragma unroll
                for (int k = 0; k < kTaBK / 16; ++k)     // 16 keys per instruction: A advances 32 bytes, B advances 16 rows of 128 bytes
                    tc_mma_bf16(tO, pdesc + uint64_t(k * 2), vdesc + uint64_t(k * 128), idesc_o, (it | k) ? 1u : 0u);   // O accumulates over the tiles
                tc_commit(&bar_vempty[st]);
                tc_commit(&bar_pv[st]);
                if (it + 2 < n_tiles) issue_s(it + 2);  // its TMEM buffer is free: bar_p(it) means every thread has read S(it)
            }
        }
    } else {
        // ===================== softmax warps: two threads per query row (= TMEM lane) =====================
        const int wq = warp & 3, half = warp >> 2;   // lane quarter; key / output-column half
        const int row = wq * 32 + lane;              // 0..127
        const int qi = q0 + row;
        const uint32_t lane_addr = tmem_base + (uint32_t(wq * 32) << 16);
        float m_use = 0.f, l_run = 0.f;              // scaling max of this row (log2 domain); l_run: this thread's 32 keys of every tile only
        const int lim0 = min(qi, S - 1) - half * 32; // last visible key of this row, relative to this thread's first column
        const uint32_t o_addr = lane_addr + 128 + uint32_t(half * 32);   // this thread's 32 columns of O
        for (int it = 0; it < n_tiles; ++it) {
            const int key0 = it * kTaBK;
            mbar_wait(&bar_s[it & 1], (it >> 1) & 1);
            tc_fence_after();
            float s[32];
            tmem_ld32(lane_addr + uint32_t((it & 1) * 64 + half * 32), s);
            // causal mask and sequence tail by index, only on tiles that touch the diagonal or the end of the sequence
            if (key0 + kTaBK - 1 > q0 || key0 + kTaBK > S) {
                const int lim = lim0 - key0;
#pragma unroll
                for (int j = 0; j < 32; ++j) s[j] = j > lim ? -INFINITY : s[j];
            }
            // the scale is positive, so the row max can be taken on the raw scores; the log2-domain shift then folds into one
            // FMA per element: p = 2^(s * c - m)
            float mx0 = fmaxf(s[0], s[1]), mx1 = fmaxf(s[2], s[3]), mx2 = fmaxf(s[4], s[5]), mx3 = fmaxf(s[6], s[7]);
#pragma unroll
            for (int j = 8; j < 32; j += 8) {
                mx0 = fmaxf(mx0, fmaxf(s[j], s[j + 1])); mx1 = fmaxf(mx1, fmaxf(s[j + 2], s[j + 3]));
                mx2 = fmaxf(mx2, fmaxf(s[j + 4], s[j + 5])); mx3 = fmaxf(mx3, fmaxf(s[j + 6], s[j + 7]));
            }
            float mx = fmaxf(fmaxf(mx0, mx1), fmaxf(mx2, mx3));
            float* xb = xch + (it & 1) * 2 * kTaBQ;  // double-buffered: the partner reads buffer t before it can pass barrier t+1
            xb[half * kTaBQ + row] = mx;
            pair_sync(wq);
            mx = fmaxf(mx, xb[(half ^ 1) * kTaBQ + row]);
            const float m_tile = mx * p.scale_log2;  // -inf: every key of this tile is masked for the row (rows past the sequence end)
            // O accumulates in TMEM over the tiles, scaled by 2^-m_use. m_use follows the running max lazily: P stays below
            // 2^kTaRescaleStep (exact in bf16 / fp32 at any scale), and O and the row sum are rescaled only when a row's max has
            // run ahead by more than that step, which after the first tiles is rare. The first tile sets m_use (O is overwritten).
            if (it == 0) {
                m_use = m_tile == -INFINITY ? 0.f : m_tile;
            } else if (__any_sync(0xffffffffu, m_tile > m_use + kTaRescaleStep)) {      // both warps of a row group vote alike
                const float m_new = fmaxf(m_use, m_tile);
                const float f = ex2_approx(m_use - m_new);
                m_use = m_new;
                l_run *= f;
                mbar_wait(&bar_pv[(it - 1) & 1], ((it - 1) >> 1) & 1);      // every P V issued so far has retired
                tc_fence_after();
                float v[32];
                tmem_ld32(o_addr, v);
#pragma unroll
                for (int i = 0; i < 32; ++i) v[i] *= f;
                tmem_st32(o_addr, v);
                tc_fence_before();
            }
            float sum0 = 0.f, sum1 = 0.f;
            uint32_t pk[16];
#pragma unroll
            for (int j = 0; j < 32; j += 2) {
                const float p0 = ex2_approx(fmaf(s[j], p.scale_log2, -m_use)), p1 = ex2_approx(fmaf(s[j + 1], p.scale_log2, -m_use));
                sum0 += p0; sum1 += p1;
                const __nv_bfloat162 h2 = __floats2bfloat162_rn(p0, p1);
                pk[j >> 1] = *reinterpret_cast<const uint32_t*>(&h2);
            }
            l_run += sum0 + sum1;
            // P row -> 128B-swizzled K-major tile, buffer it & 1 (free once P(it-2) V has retired): 16-byte piece c of row r sits at
            // r*128 + ((c ^ (r & 7)) << 4); this thread owns pieces 4 half .. 4 half + 3
            if (it >= 2) mbar_wait(&bar_pv[it & 1], ((it - 2) >> 1) & 1);
            uint8_t* prow = sP + (it & 1) * kTaPBytes + row * 128;
#pragma unroll
            for (int c = 0; c < 4; ++c)
                *reinterpret_cast<uint4*>(prow + (((half * 4 + c) ^ (row & 7)) << 4)) = make_uint4(pk[4 * c], pk[4 * c + 1], pk[4 * c + 2], pk[4 * c + 3]);
            fence_proxy_async_smem();                 // generic-proxy writes visible to the tensor core's async proxy
            tc_fence_before();
            __syncwarp();
            if (lane == 0) mbar_arrive(&bar_p[it & 1]);
        }
        mbar_wait(&bar_pv[(n_tiles - 1) & 1], ((n_tiles - 1) >> 1) & 1);
        tc_fence_after();
        float acc[32];
        tmem_ld32(o_addr, acc);
        // row sum = the two threads' partial sums (both are relative to the same running max)
        float* xb = xch + (n_tiles & 1) * 2 * kTaBQ;
        xb[half * kTaBQ + row] = l_run;
        pair_sync(wq);
        const float l_row = l_run + xb[(half ^ 1) * kTaBQ + row];
        const float inv = l_row > 0.f ? 1.0f / l_row : 0.f;
        // every product has retired: the K/V/P region is free and stages this warp's 32 rows x 32 columns for coalesced stores
        float* stage = reinterpret_cast<float*>(sK) + warp * 32 * 36;        // 36 floats per row; 8 warps x 4.5 KB (fits K+V+P = 64 KB)
#pragma unroll
        for (int c = 0; c < 8; ++c)
            *reinterpret_cast<float4*>(stage + lane * 36 + c * 4) = make_float4(acc[4 * c] * inv, acc[4 * c + 1] * inv, acc[4 * c + 2] * inv, acc[4 * c + 3] * inv);
        __syncwarp();
        const int sub = lane >> 3, c4 = lane & 7;     // four rows per instruction, 8 lanes x float4 per row (128 bytes)
        float* ob = p.ctx ? p.ctx + int64_t(b) * p.osb + h * kTaHD + half * 32 : nullptr;
        __nv_bfloat16* ob16 = p.ctx16 ? p.ctx16 + int64_t(b) * p.osb + h * kTaHD + half * 32 : nullptr;
#pragma unroll 4
        for (int r = 0; r < 32; r += 4) {
            const int lr = r + sub, gr = q0 + wq * 32 + lr;
            if (gr >= S) continue;
            const float4 o4 = *reinterpret_cast<const float4*>(stage + lr * 36 + c4 * 4);
            if (ob) *reinterpret_cast<float4*>(ob + int64_t(gr) * p.oss + c4 * 4) = o4;
            if (ob16) st_bf16x4(ob16 + int64_t(gr) * p.oss + c4 * 4, o4);
        }
    }

    tc_fence_before();
    __syncthreads();
    if (warp == kTaMmaWarp) {
        tc_fence_after();
        asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" :: "r"(tmem_base), "r"(uint32_t(kTaTmemCols)) : "memory");
    }
}

// One tensor map over the bf16 [B][S][3C] projection output: box = 64 columns (one head's 128 bytes) x 64 rows, 128B swizzle.
static bool make_qkv_map(CUtensorMap* out, const void* ptr, uint64_t width, uint64_t S, uint64_t B, uint64_t ld, uint64_t batch_stride) {
    struct Key { const void* p; uint64_t w, s, b, ld, bs; bool operator==(const Key& o) const { return p == o.p && w == o.w && s == o.s && b == o.b && ld == o.ld && bs == o.bs; } };
    struct KeyHash { size_t operator()(const Key& k) const { return std::hash<const void*>()(k.p) ^ (k.w * 31 + k.s * 131 + k.b * 1031 + k.ld * 7 + k.bs); } };
    static std::unordered_map<Key, CUtensorMap, KeyHash> cache;
    static std::mutex mu;
    Key key{ptr, width, S, B, ld, batch_stride};
    {
        std::lock_guard<std::mutex> lk(mu);
        auto it = cache.find(key);
        if (it != cache.end()) { *out = it->second; return true; }
    }
    EncodeTiledFn enc = get_encode_fn();
    if (!enc) return false;
    alignas(64) CUtensorMap m;
    cuuint64_t dims[3] = {width, S, B};
    cuuint64_t strides[2] = {ld * 2, batch_stride * 2};
    cuuint32_t box[3] = {64, 64, 1};
    cuuint32_t estr[3] = {1, 1, 1};
    CUresult r = enc(&m, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 3, const_cast<void*>(ptr), dims, strides, box, estr, CU_TENSOR_MAP_INTERLEAVE_NONE,
                     CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_L2_256B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
    if (r != CUDA_SUCCESS) return false;
    {
        std::lock_guard<std::mutex> lk(mu);
        if (cache.size() > 1024) cache.clear();
        cache.emplace(key, m);
    }
    *out = m;
    return true;
}

}  // namespace mft

using namespace mft;

/* qkv16: bf16 [B, S, 3C] (row pitch ld elements, batch stride S*ld); ctx / ctx16 [B, S, C] (either may be NULL). */
extern "C" int mft_causal_attention_tc_fwd(const void* qkv_bf16, int64_t ld, float* ctx, void* ctx_bf16, int32_t B, int32_t H, int32_t S,
                                           int32_t Hd, float scale, mft_stream_t stream) {
    if (B == 0 || H == 0 || S == 0) return 0;
    if (!qkv_bf16 || (!ctx && !ctx_bf16)) return set_error(1, "mft_causal_attention_tc_fwd: null pointer");
    if (Hd != kTaHD) return set_error(1, "mft_causal_attention_tc_fwd: head_dim must be 64 (got %d)", Hd);
    const int64_t C = int64_t(H) * Hd;
    if (ld < 3 * C || (ld & 7) || !aligned16(qkv_bf16)) return set_error(1, "mft_causal_attention_tc_fwd: qkv must be 16-byte aligned with a pitch that is a multiple of 8 and >= 3C");
    if ((ctx && !aligned16(ctx)) || (ctx_bf16 && (reinterpret_cast<uintptr_t>(ctx_bf16) & 7u))) return set_error(1, "mft_causal_attention_tc_fwd: misaligned output");
    if (int64_t(B) * H > 65535) return set_error(1, "mft_causal_attention_tc_fwd: B*H too large");
    CUtensorMap map;
    if (!make_qkv_map(&map, qkv_bf16, uint64_t(3 * C), uint64_t(S), uint64_t(B), uint64_t(ld), uint64_t(S) * uint64_t(ld)))
        return set_error(3, "mft_causal_attention_tc_fwd: cuTensorMapEncodeTiled failed");
    static bool attr_set[kMaxDevices] = {};
    if (first_use_on_device(attr_set)) {
        cudaError_t e = cudaFuncSetAttribute(attention_tc_fwd_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, kTaSmem);
        if (e != cudaSuccess) return set_error(2, "mft_causal_attention_tc_fwd: smem attribute: %s", cudaGetErrorString(e));
    }
    TaArgs a;
    a.ctx = ctx; a.ctx16 = static_cast<__nv_bfloat16*>(ctx_bf16); a.H = H; a.S = S; a.osb = int64_t(S) * C; a.oss = C;
    a.scale_log2 = scale * 1.4426950408889634f;
    dim3 grid(unsigned(ceil_div(S, kTaBQ)), unsigned(B * H));
    pdl_launch(attention_tc_fwd_kernel, grid, kTaThreads, kTaSmem, as_stream(stream))(map, a, int(C));
    MFT_CHECK_LAUNCH("mft_causal_attention_tc_fwd");
    return 0;
}
