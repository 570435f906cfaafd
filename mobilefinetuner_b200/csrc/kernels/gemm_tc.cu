// tcgen05 TF32 GEMM for sm_100a:  C = alpha*(A*B^T [+ A2*B2^T]) (+bias) (+beta*C) [, gelu aux]
//   A [m,k] and B [n,k] both K-major fp32 in HBM.
// Persistent, warp-specialised CTA (one per SM):
//   warp 0    : TMA producer  (cp.async.bulk.tensor 3-D boxes, 128B swizzle, TFLOAT32 rounding in the TMA)
//   warp 1    : TMEM allocator + single-thread tcgen05.mma issuer (kind::tf32, M=128, N=BN, K=8 per instr)
//   warps 2-5 : epilogue (tcgen05.ld 32x32b -> registers -> alpha/bias/beta/GELU -> st.global.v4)
// Pipelines: NS-deep smem ring (full/empty mbarriers) between TMA and MMA; double-buffered TMEM accumulator
// (tmem_full/tmem_empty) between MMA and epilogue so tile i's epilogue overlaps tile i+1's main loop.
// The optional second K-segment (A2 [m,k2], B2 [n,k2]) feeds extra k-blocks into the same accumulator:
// this is how LoRA's rank-r update s*(xA)*B is fused with the base x*W product (nn/lora_linear.cpp:47-108).
#include <cuda.h>
#include <cstdlib>
#include <cstring>
#include <mutex>
#include <unordered_map>
#include <vector>
#include "common.cuh"
#include "gemm_internal.cuh"
#include "tc_common.cuh"

namespace mft {

constexpr int TBM = 128;          // tile rows  (UMMA M)
constexpr int TBK = 32;           // fp32 elements per k-block = one 128-byte swizzle span
constexpr int TUK = 8;            // UMMA K for tf32 (32 bytes)
#ifndef MFT_TC_EPI_WARPS
#define MFT_TC_EPI_WARPS 4
#endif
#ifndef MFT_TC_EPI_BOXES
#define MFT_TC_EPI_BOXES 1
#endif
constexpr int kEpiWarpsDefault = MFT_TC_EPI_WARPS;   // 4 or 8 epilogue warps: one or two per TMEM lane group (two: each takes half of the tile's columns); a template parameter of the kernel (EW)
constexpr int kEpiBoxes = MFT_TC_EPI_BOXES;   // staging boxes per epilogue warp (2: the next box is filled while the TMA drains one)
static_assert(kEpiWarpsDefault == 4 || kEpiWarpsDefault == 8, "epilogue warps");
constexpr int kEpiPitch = 36;     // floats per staged row (fallback path): 16-byte aligned and conflict-free for both access patterns
constexpr int kEpiBufBytes = 32 * 128;                  // one 32 x 32 fp32 box in the TMA (128B-swizzled) layout
constexpr int kEpiStageBytes = (kEpiBoxes * kEpiBufBytes > 32 * kEpiPitch * 4) ? kEpiBoxes * kEpiBufBytes : 5120;   // fallback path needs 4608; keep 1024-byte alignment
// staging per epilogue warp: the eight-warp form (GELU aux output) keeps two boxes, pre-activation and activation
constexpr int epi_stage_bytes(int ew) { return ew == 8 ? 2 * kEpiBufBytes : kEpiStageBytes; }
constexpr int epi_warp_bytes(int ew) { return epi_stage_bytes(ew) + 1024; }    // boxes + this warp's slice of the bias (<= 256 floats)
constexpr int kSmemTotal = 227 * 1024;

// CG = CTAs cooperating on one UMMA (cta_group). CG = 2: a cluster of two CTAs computes a 256 x BN tile pair — each CTA
// stages its own 128 rows of A and HALF of the B tile, the leader issues tcgen05.mma.cta_group::2 (M = 256) which reads
// both halves, and each CTA keeps its 128 x BN accumulator in its own TMEM. Halving the B traffic per SM is what lets
// the tensor pipe run above ~50 % (single-CTA UMMA is bounded by shared-memory bandwidth).
template <int BN, int CG = 1, int EW = kEpiWarpsDefault> struct TcCfg {
    static constexpr int b_rows = BN / CG;                 // B rows staged by one CTA
    static constexpr int stage_bytes = (TBM + b_rows) * TBK * 4;
    static constexpr int epi_bytes = EW * epi_warp_bytes(EW);   // per-warp staging for the output boxes
    static constexpr int budget = kSmemTotal - epi_bytes - 1024 /*align slack*/ - 256 /*barriers*/;
    static constexpr int stages = (budget / stage_bytes) > 8 ? 8 : (budget / stage_bytes);
    static constexpr int tmem_cols = 2 * BN;             // two accumulator buffers
    static constexpr int smem_bytes = stages * stage_bytes + epi_bytes + 1024 + 256;
    static_assert(stages >= 3, "smem ring too shallow");
    static_assert(stage_bytes % 1024 == 0, "stage size must keep the 1024-byte swizzle alignment");
};

// tanh-GELU for the epilogue: 0.5 x (1 + tanh(u)) = x * sigmoid(2u), with sigmoid through ex2.approx / rcp.approx — about ten
// instructions per element and ~1e-6 absolute accuracy, where tanhf costs ~50 with a single epilogue warp per scheduler
// (the GELU epilogue of fc_in took 93 us against 39 us for the plain product). u = 0.7978845608 (x + 0.044715 x^3).
__device__ __forceinline__ float gelu_tanh_epi(float x) {
    const float u2 = 1.5957691216f * (x + 0.044715f * x * x * x);       // 2u
    return __fdividef(x, 1.0f + __expf(-u2));
}

struct TcArgs {
    int m, n, batch;
    int kb1, kb2;              // k-blocks of segment 1 / segment 2
    float* c; int64_t ldc, stride_c;
    float* gelu_out;
    const float* bias;
    const float* addend; int64_t ld_addend;
    const float* gelu_grad_x; int64_t ld_gelu_grad_x;
    __nv_bfloat16* gelu16; int64_t ld_gelu16;   // bf16 copy of the GELU aux output (gelu_out may be NULL: bf16 only)
    __nv_bfloat16* c16; int64_t ld_c16;         // bf16 copy of C (c may be NULL on the TMA path: bf16 only)
    int tma_c16;               // 1: c is NULL and the bf16 result leaves through 64-column TMA boxes (map_c)
    int tma_g16;               // 1: with tma_c16, the GELU aux output (bf16 only) leaves through boxes too (map_g; EW = 8 form)
    float alpha, beta;
    int tiles_m, tiles_n;
    int group_m;               // raster: m-units per group (see tile_coords)
    int tma_c;                 // 1: the epilogue writes C (and gelu_out) with TMA box stores through map_c / map_g
};

// Raster order of the persistent tile loop. Tiles that run at the same time should share operand rows, so that each slab of
// A and B comes from HBM once and from L2 afterwards: m-units are taken in groups of `gm`; inside a group the index runs over
// the group's m-units first, then over ALL n-tiles, before the next group starts. gm >= tiles_mu is the plain m-fastest order
// (fine while A fits in L2); with 32 768 rows the plain order re-read A once per n-tile (twice the algorithmic DRAM bytes on
// fc_in, four times on the K = 4096 dX product).
__device__ __forceinline__ void tile_coords(int tile, int tiles_mu, int tiles_n, int gm, int& mu, int& nt, int& bz) {
    const int per_b = tiles_mu * tiles_n;
    bz = tile / per_b;
    int t = tile - bz * per_b;
    const int grp = t / (gm * tiles_n);
    const int first = grp * gm;
    const int rows = min(gm, tiles_mu - first);
    t -= grp * gm * tiles_n;
    nt = t / rows;
    mu = first + (t - nt * rows);
}

// BF16 = false: fp32 operands in HBM, TF32 MMA (32 elements per 128-byte k-block, K = 8 per instruction).
// BF16 = true : bf16 operands in HBM, kind::f16 MMA (64 elements per k-block, K = 16 per instruction): same bytes per
//               k-block, twice the FLOPs per byte staged.
// FUSED = true adds the residual / GELU-gradient operands to the epilogue; a separate instantiation so that the plain
// epilogue's instruction schedule (which bounds the kernel on short-K shapes) is not disturbed by code it never runs.
// FORM specialises the epilogue at compile time (the box epilogues are latency-critical: a runtime flag tested inside their
// unrolled chunk loop costs several percent of the kernel): 0 = every output form decided at run time; 1 = fp32 result through
// TMA boxes, nothing else; 2 = bf16-only result through boxes; 3 = bf16-only result and bf16-only GELU aux through boxes (EW = 8).
template <int BN, int CG, bool BF16, bool FUSED, int EW, int FORM>
__global__ void __launch_bounds__(64 + 32 * EW, 1)
gemm_tc_kernel(const __grid_constant__ CUtensorMap map_a, const __grid_constant__ CUtensorMap map_b,
               const __grid_constant__ CUtensorMap map_a2, const __grid_constant__ CUtensorMap map_b2,
               const __grid_constant__ CUtensorMap map_c, const __grid_constant__ CUtensorMap map_g, TcArgs p) {
    pdl_sync();
    using Cfg = TcCfg<BN, CG, EW>;
    constexpr int kEpiWarps = EW;
    constexpr int KE = BF16 ? 64 : 32;                     // elements per k-block (one 128-byte swizzle span)
    const uint32_t cta_rank = CG == 2 ? cluster_ctarank() : 0u;   // 0 = leader (issues the MMAs)
    const int unit = CG == 2 ? int(blockIdx.x >> 1) : int(blockIdx.x);        // CTA (or CTA pair) index
    const int num_units = CG == 2 ? int(gridDim.x >> 1) : int(gridDim.x);
    if (CG == 2) cluster_sync_all();                               // both CTAs alive before the paired TMEM allocation
    constexpr int NS = Cfg::stages;
    extern __shared__ uint8_t smem_raw[];
    uint8_t* tiles = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));
    uint8_t* epi_stage = tiles + NS * Cfg::stage_bytes;      // 1024-byte aligned: stage sizes are multiples of 1024
    uint64_t* bars = reinterpret_cast<uint64_t*>(tiles + NS * Cfg::stage_bytes + Cfg::epi_bytes);
    uint64_t* full = bars;                 // [NS]
    uint64_t* empty = bars + NS;           // [NS]
    uint64_t* tmem_full = bars + 2 * NS;   // [2]
    uint64_t* tmem_empty = bars + 2 * NS + 2;  // [2]
    uint32_t* tmem_base_slot = reinterpret_cast<uint32_t*>(bars + 2 * NS + 4);

    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    // tiles are counted in units of CG m-tiles: with CG = 2 a "tile" is a pair of adjacent 128-row tiles
    const int tiles_mu = (p.tiles_m + CG - 1) / CG;
    const int num_tiles = tiles_mu * p.tiles_n * p.batch;
    const int num_kb = p.kb1 + p.kb2;

    if (warp == 0 && lane == 0) {
        asm volatile("prefetch.tensormap [%0];" :: "l"(reinterpret_cast<uint64_t>(&map_a)) : "memory");
        asm volatile("prefetch.tensormap [%0];" :: "l"(reinterpret_cast<uint64_t>(&map_b)) : "memory");
        for (int i = 0; i < NS; ++i) { mbar_init(&full[i], CG); mbar_init(&empty[i], 1); }      // full: one arrival per CTA, at the leader
        for (int i = 0; i < 2; ++i) { mbar_init(&tmem_full[i], 1); mbar_init(&tmem_empty[i], kEpiWarps * CG); }   // tmem_empty: leader only
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    }
    if (warp == 1) {
        if (CG == 2) {
            asm volatile("tcgen05.alloc.cta_group::2.sync.aligned.shared::cta.b32 [%0], %1;"
                         :: "r"(smem_u32(tmem_base_slot)), "r"(uint32_t(Cfg::tmem_cols)) : "memory");
            asm volatile("tcgen05.relinquish_alloc_permit.cta_group::2.sync.aligned;" ::: "memory");
        } else {
            asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;"
                         :: "r"(smem_u32(tmem_base_slot)), "r"(uint32_t(Cfg::tmem_cols)) : "memory");
            asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
        }
    }
    tc_fence_before();
    if (CG == 2) cluster_sync_all(); else __syncthreads();
    tc_fence_after();
    const uint32_t tmem_base = *tmem_base_slot;

    if (warp == 0) {
        // ===================== TMA producer =====================
        if (lane == 0) {
            int stage = 0; uint32_t phase = 0;
            for (int tile = unit; tile < num_tiles; tile += num_units) {
                int mu, nt, bz;
                tile_coords(tile, tiles_mu, p.tiles_n, p.group_m, mu, nt, bz);
                const int mt = mu * CG + int(cta_rank);                         // this CTA's own 128-row tile
                const int brow = nt * BN + int(cta_rank) * Cfg::b_rows;         // this CTA's share of the B tile
                for (int kb = 0; kb < num_kb; ++kb) {
                    mbar_wait(&empty[stage], phase ^ 1);
                    uint8_t* sa = tiles + stage * Cfg::stage_bytes;
                    uint8_t* sb = sa + TBM * TBK * 4;
                    if (CG == 1) {
                        mbar_arrive_expect_tx(&full[stage], uint32_t(Cfg::stage_bytes));
                        if (kb < p.kb1) {
                            tma_load_3d(sa, &map_a, &full[stage], kb * KE, mt * TBM, bz);
                            tma_load_3d(sb, &map_b, &full[stage], kb * KE, brow, bz);
                        } else {
                            tma_load_3d(sa, &map_a2, &full[stage], (kb - p.kb1) * KE, mt * TBM, 0);
                            tma_load_3d(sb, &map_b2, &full[stage], (kb - p.kb1) * KE, brow, 0);
                        }
                    } else {
                        if (kb < p.kb1) {
                            tma_load_3d_2sm(sa, &map_a, &full[stage], kb * KE, mt * TBM, bz);
                            tma_load_3d_2sm(sb, &map_b, &full[stage], kb * KE, brow, bz);
                        } else {
                            tma_load_3d_2sm(sa, &map_a2, &full[stage], (kb - p.kb1) * KE, mt * TBM, 0);
                            tma_load_3d_2sm(sb, &map_b2, &full[stage], (kb - p.kb1) * KE, brow, 0);
                        }
                        // the leader's barrier collects both CTAs' bytes: it expects 2 x stage_bytes and one arrival per CTA
                        if (cta_rank == 0) mbar_arrive_expect_tx(&full[stage], uint32_t(Cfg::stage_bytes) * 2u);
                        else mbar_arrive_remote(&full[stage], 0);
                    }
                    if (++stage == NS) { stage = 0; phase ^= 1; }
                }
            }
        }
    } else if (warp == 1) {
        // ===================== MMA issuer (one thread) =====================
        if (lane == 0 && cta_rank == 0) {
            // instruction descriptor: D=f32, A=B=tf32, both K-major, N=BN, M=128 per CTA (256 across the pair)
            constexpr uint32_t fmt = BF16 ? 1u : 2u;       // a/b format: 1 = BF16, 2 = TF32
            constexpr uint32_t idesc = (1u << 4) | (fmt << 7) | (fmt << 10) | (uint32_t(BN >> 3) << 17) | (uint32_t((TBM * CG) >> 4) << 24);
            int stage = 0; uint32_t phase = 0;
            int it = 0;
            for (int tile = unit; tile < num_tiles; tile += num_units, ++it) {
                const int as = it & 1;
                const uint32_t aphase = (it >> 1) & 1;
                mbar_wait(&tmem_empty[as], aphase ^ 1);
                tc_fence_after();
                const uint32_t d_tmem = tmem_base + uint32_t(as * BN);
                for (int kb = 0; kb < num_kb; ++kb) {
                    mbar_wait(&full[stage], phase);
                    tc_fence_after();
                    const uint32_t sa = smem_u32(tiles + stage * Cfg::stage_bytes);
                    const uint32_t sb = sa + TBM * TBK * 4;
                    const uint64_t adesc = make_kmajor_sw128_desc(sa);
                    const uint64_t bdesc = make_kmajor_sw128_desc(sb);
#pragma unroll
                    for (int k = 0; k < TBK / TUK; ++k) {
                        // advance 32 bytes (one UMMA K) inside the 128B swizzle span: +2 in 16-byte units
                        const uint32_t acc = (kb | k) != 0 ? 1u : 0u;
                        if (BF16) {
                            if (CG == 2) tc_mma_bf16_2sm(d_tmem, adesc + uint64_t(k * 2), bdesc + uint64_t(k * 2), idesc, acc);
                            else tc_mma_bf16(d_tmem, adesc + uint64_t(k * 2), bdesc + uint64_t(k * 2), idesc, acc);
                        } else {
                            if (CG == 2) tc_mma_tf32_2sm(d_tmem, adesc + uint64_t(k * 2), bdesc + uint64_t(k * 2), idesc, acc);
                            else tc_mma_tf32(d_tmem, adesc + uint64_t(k * 2), bdesc + uint64_t(k * 2), idesc, acc);
                        }
                    }
                    // frees the smem slot (in both CTAs when paired) once these MMAs retire
                    if (CG == 2) tc_commit_2sm(&empty[stage]); else tc_commit(&empty[stage]);
                    if (++stage == NS) { stage = 0; phase ^= 1; }
                }
                if (CG == 2) tc_commit_2sm(&tmem_full[as]); else tc_commit(&tmem_full[as]);   // accumulator(s) ready
            }
            if (CG == 2 && it > 0) {   // do not run ahead of the peer's last epilogue arrivals
                const int last = it - 1;
                mbar_wait(&tmem_empty[last & 1], (last >> 1) & 1);
            }
        }
    } else {
        // ===================== epilogue warps (TMEM -> regs -> smem transpose -> HBM) =====================
        // tcgen05.ld hands each thread one accumulator ROW; storing that directly would hit 32 different 128-byte lines
        // per warp instruction. Each 32x32 chunk is therefore bounced through a per-warp smem buffer so that a warp
        // instruction writes 4 full 128-byte row segments (8 lanes x float4 per row).
        const int lg = warp & 3;                          // TMEM lane group this warp may touch (hardware: warp id % 4)
        const int half = (warp - 2) >> 2;                 // which part of the tile's columns this warp drains
        constexpr int HC = BN / (kEpiWarps / 4);          // columns per epilogue warp
        // all boxes first (each 1024-byte aligned: the 128B swizzle is a function of the absolute smem address), bias slices after
        uint8_t* wbuf = epi_stage + (warp - 2) * epi_stage_bytes(EW);
        float* stg = reinterpret_cast<float*>(wbuf);      // fallback path: 32 x kEpiPitch floats (fits in the two boxes)
        float* sbias = reinterpret_cast<float*>(epi_stage + kEpiWarps * epi_stage_bytes(EW) + (warp - 2) * 1024);   // [HC] (HC <= 256)
        const int r4 = lane >> 3, c4 = lane & 7;          // read-back role: row within a group of 4, float4 column
        const bool vec_ok = ((p.ldc & 3) == 0) && ((reinterpret_cast<uintptr_t>(p.c) & 15u) == 0) &&
                            ((p.stride_c & 3) == 0) &&
                            (!p.gelu_out || (reinterpret_cast<uintptr_t>(p.gelu_out) & 15u) == 0) &&
                            (!p.addend || (((reinterpret_cast<uintptr_t>(p.addend) & 15u) | (p.ld_addend & 3)) == 0)) &&
                            (!p.gelu_grad_x || (((reinterpret_cast<uintptr_t>(p.gelu_grad_x) & 15u) | (p.ld_gelu_grad_x & 3)) == 0)) &&
                            (!p.bias || (reinterpret_cast<uintptr_t>(p.bias) & 15u) == 0) &&
                            (!p.c16 || (((reinterpret_cast<uintptr_t>(p.c16) & 7u) | (p.ld_c16 & 3)) == 0));
        const bool use_tma = FORM != 0 ? true : (!FUSED && p.tma_c != 0);
        const bool f_c16box = FORM != 0 ? (FORM >= 2) : (p.tma_c16 != 0);
        const bool f_g16box = FORM != 0 ? (FORM == 3) : (EW == 8 && p.tma_g16 != 0);
        const bool f_c16reg = FORM != 0 ? false : (p.c16 != nullptr);
        uint32_t nstore = 0;                              // boxes this warp has handed to the TMA (ring of two buffers)
        int it = 0;
        for (int tile = unit; tile < num_tiles; tile += num_units, ++it) {
            int mu, nt, bz;
            tile_coords(tile, tiles_mu, p.tiles_n, p.group_m, mu, nt, bz);
            const int mt = mu * CG + int(cta_rank);
            const int as = it & 1;
            const uint32_t aphase = (it >> 1) & 1;
            mbar_wait(&tmem_full[as], aphase);
            tc_fence_after();
            const int row_base = mt * TBM + lg * 32;
            const int n0 = nt * BN;
            float* cbase = p.c + int64_t(bz) * p.stride_c;
            float* gbase = (FORM == 0 && p.gelu_out) ? p.gelu_out + int64_t(bz) * p.stride_c : nullptr;
            const float* abase = (FUSED && p.addend) ? p.addend + int64_t(bz) * p.stride_c : nullptr;
            const float* xbase = (FUSED && p.gelu_grad_x) ? p.gelu_grad_x + int64_t(bz) * p.stride_c : nullptr;
            if (use_tma) {
                // ---- TMA path: thread = accumulator row; the row's 32 values of a chunk go to the swizzled box as eight
                // 16-byte pieces (piece j of row r sits at r*128 + ((j ^ (r & 7)) << 4): conflict-free for a quarter warp),
                // then one lane hands the box to the TMA, which writes full lines and clips at the matrix edges.
                const int colw = n0 + half * HC;          // first column of this warp
                if (p.bias && colw < p.n) {               // this warp's bias slice -> smem (coalesced)
                    const bool al = (reinterpret_cast<uintptr_t>(p.bias) & 15u) == 0 && (colw & 3) == 0;
                    for (int i = lane * 4; i < HC; i += 128) {
                        const int c = colw + i;
                        float4 b4 = make_float4(0.f, 0.f, 0.f, 0.f);
                        if (c + 4 <= p.n && al) b4 = __ldg(reinterpret_cast<const float4*>(p.bias + c));
                        else {
                            if (c < p.n) b4.x = __ldg(p.bias + c);
                            if (c + 1 < p.n) b4.y = __ldg(p.bias + c + 1);
                            if (c + 2 < p.n) b4.z = __ldg(p.bias + c + 2);
                            if (c + 3 < p.n) b4.w = __ldg(p.bias + c + 3);
                        }
                        *reinterpret_cast<float4*>(sbias + i) = b4;
                    }
                    __syncwarp();
                }
#pragma unroll
                for (int j = 0; j < HC / 32; ++j) {
                    const int cc = half * HC + j * 32;    // column offset inside the tile
                    if (n0 + cc >= p.n) break;            // warp-uniform
                    float v[32];
                    tmem_ld32(tmem_base + (uint32_t(lg * 32) << 16) + uint32_t(as * BN + cc), v);
                    if (p.bias) {
#pragma unroll
                        for (int q = 0; q < 8; ++q) {
                            const float4 b4 = *reinterpret_cast<const float4*>(sbias + j * 32 + q * 4);   // same address for all lanes
                            v[4 * q] = v[4 * q] * p.alpha + b4.x; v[4 * q + 1] = v[4 * q + 1] * p.alpha + b4.y;
                            v[4 * q + 2] = v[4 * q + 2] * p.alpha + b4.z; v[4 * q + 3] = v[4 * q + 3] * p.alpha + b4.w;
                        }
                    } else if (p.alpha != 1.0f) {
#pragma unroll
                        for (int q = 0; q < 32; ++q) v[q] *= p.alpha;
                    }
                    if (f_c16box) {
                        // bf16-only result: two 32-column chunks share one 64-column box (128 bytes of bf16 per row, same
                        // 128B swizzle); the box leaves after its second half, or after the last chunk inside the matrix
                        uint8_t* box = wbuf;
                        if ((j & 1) == 0) {
                            if (lane == 0) tma_store_wait_read<0>();
                            __syncwarp();
                        }
                        uint8_t* rowp = box + lane * 128;
#pragma unroll
                        for (int q = 0; q < 4; ++q) {
                            const __nv_bfloat162 h0 = __floats2bfloat162_rn(v[8 * q], v[8 * q + 1]), h1 = __floats2bfloat162_rn(v[8 * q + 2], v[8 * q + 3]);
                            const __nv_bfloat162 h2 = __floats2bfloat162_rn(v[8 * q + 4], v[8 * q + 5]), h3 = __floats2bfloat162_rn(v[8 * q + 6], v[8 * q + 7]);
                            *reinterpret_cast<uint4*>(rowp + ((((j & 1) * 4 + q) ^ (lane & 7)) << 4)) =
                                make_uint4(*reinterpret_cast<const uint32_t*>(&h0), *reinterpret_cast<const uint32_t*>(&h1),
                                           *reinterpret_cast<const uint32_t*>(&h2), *reinterpret_cast<const uint32_t*>(&h3));
                        }
                        if (f_g16box) {                                  // GELU of the same values into the second box
                            uint8_t* rowg = box + kEpiBufBytes + lane * 128;
#pragma unroll
                            for (int q = 0; q < 4; ++q) {
                                const __nv_bfloat162 h0 = __floats2bfloat162_rn(gelu_tanh_epi(v[8 * q]), gelu_tanh_epi(v[8 * q + 1]));
                                const __nv_bfloat162 h1 = __floats2bfloat162_rn(gelu_tanh_epi(v[8 * q + 2]), gelu_tanh_epi(v[8 * q + 3]));
                                const __nv_bfloat162 h2 = __floats2bfloat162_rn(gelu_tanh_epi(v[8 * q + 4]), gelu_tanh_epi(v[8 * q + 5]));
                                const __nv_bfloat162 h3 = __floats2bfloat162_rn(gelu_tanh_epi(v[8 * q + 6]), gelu_tanh_epi(v[8 * q + 7]));
                                *reinterpret_cast<uint4*>(rowg + ((((j & 1) * 4 + q) ^ (lane & 7)) << 4)) =
                                    make_uint4(*reinterpret_cast<const uint32_t*>(&h0), *reinterpret_cast<const uint32_t*>(&h1),
                                               *reinterpret_cast<const uint32_t*>(&h2), *reinterpret_cast<const uint32_t*>(&h3));
                            }
                        }
                        const bool last_chunk = (j == HC / 32 - 1) || (n0 + cc + 32 >= p.n);
                        if ((j & 1) || last_chunk) {
                            fence_proxy_async_smem();
                            __syncwarp();
                            if (lane == 0) {
                                tma_store_3d(&map_c, box, n0 + (cc & ~63), row_base, bz);
                                if (f_g16box) tma_store_3d(&map_g, box + kEpiBufBytes, n0 + (cc & ~63), row_base, bz);
                                tma_store_commit();
                            }
                        }
                    } else {
                        uint8_t* box = wbuf + (nstore % kEpiBoxes) * kEpiBufBytes;
                        if (lane == 0) tma_store_wait_read<kEpiBoxes - 1>();      // the store that last used this box has read it
                        __syncwarp();
                        uint8_t* rowp = box + lane * 128;
#pragma unroll
                        for (int q = 0; q < 8; ++q)
                            *reinterpret_cast<float4*>(rowp + ((q ^ (lane & 7)) << 4)) = make_float4(v[4 * q], v[4 * q + 1], v[4 * q + 2], v[4 * q + 3]);
                        fence_proxy_async_smem();
                        __syncwarp();
                        if (lane == 0) { tma_store_3d(&map_c, box, n0 + cc, row_base, bz); tma_store_commit(); }
                        ++nstore;
                        if (f_c16reg) {                               // fp32 through the TMA, bf16 copy straight from registers
                            const int grow = row_base + lane;
                            if (grow < p.m) {
                                __nv_bfloat16* d16 = p.c16 + int64_t(grow) * p.ld_c16 + n0 + cc;
                                if (n0 + cc + 32 <= p.n) {
#pragma unroll
                                    for (int q = 0; q < 4; ++q) {
                                        const __nv_bfloat162 h0 = __floats2bfloat162_rn(v[8 * q], v[8 * q + 1]), h1 = __floats2bfloat162_rn(v[8 * q + 2], v[8 * q + 3]);
                                        const __nv_bfloat162 h2 = __floats2bfloat162_rn(v[8 * q + 4], v[8 * q + 5]), h3 = __floats2bfloat162_rn(v[8 * q + 6], v[8 * q + 7]);
                                        *reinterpret_cast<uint4*>(d16 + 8 * q) = make_uint4(*reinterpret_cast<const uint32_t*>(&h0), *reinterpret_cast<const uint32_t*>(&h1),
                                                                                           *reinterpret_cast<const uint32_t*>(&h2), *reinterpret_cast<const uint32_t*>(&h3));
                                    }
                                } else {
#pragma unroll
                                    for (int e = 0; e < 32; ++e)          // compile-time indices: v[] must stay in registers
                                        if (n0 + cc + e < p.n) d16[e] = __float2bfloat16_rn(v[e]);
                                }
                            }
                        }
                    }
                    if (gbase) {
                        uint8_t* box = wbuf + (nstore % kEpiBoxes) * kEpiBufBytes;
                        if (lane == 0) tma_store_wait_read<kEpiBoxes - 1>();
                        __syncwarp();
                        uint8_t* rowp = box + lane * 128;
#pragma unroll
                        for (int q = 0; q < 8; ++q) {
                            float4 gv;
                            gv = make_float4(gelu_tanh_epi(v[4 * q]), gelu_tanh_epi(v[4 * q + 1]), gelu_tanh_epi(v[4 * q + 2]), gelu_tanh_epi(v[4 * q + 3]));
                            *reinterpret_cast<float4*>(rowp + ((q ^ (lane & 7)) << 4)) = gv;
                        }
                        fence_proxy_async_smem();
                        __syncwarp();
                        if (lane == 0) { tma_store_3d(&map_g, box, n0 + cc, row_base, bz); tma_store_commit(); }
                        ++nstore;
                    }
                }
            } else {
#pragma unroll 1
            for (int c0 = half * HC; c0 < (half + 1) * HC; c0 += 32) {
                if (n0 + c0 >= p.n) break;               // warp-uniform
                float v[32];
                tmem_ld32(tmem_base + (uint32_t(lg * 32) << 16) + uint32_t(as * BN + c0), v);
                float4* srow = reinterpret_cast<float4*>(stg + lane * kEpiPitch);
#pragma unroll
                for (int j = 0; j < 8; ++j) srow[j] = make_float4(v[4 * j], v[4 * j + 1], v[4 * j + 2], v[4 * j + 3]);
                __syncwarp();
                const int col = n0 + c0 + c4 * 4;
                if (vec_ok && col + 4 <= p.n && (!FUSED || (!abase && !xbase))) {
                    float4 bb = make_float4(0.f, 0.f, 0.f, 0.f);
                    if (p.bias) bb = __ldg(reinterpret_cast<const float4*>(p.bias + col));
#pragma unroll
                    for (int q = 0; q < 8; ++q) {
                        const int r = q * 4 + r4;
                        const int grow = row_base + r;
                        if (grow < p.m) {
                            float4 a = *reinterpret_cast<const float4*>(stg + r * kEpiPitch + c4 * 4);
                            float4 o = make_float4(a.x * p.alpha + bb.x, a.y * p.alpha + bb.y, a.z * p.alpha + bb.z, a.w * p.alpha + bb.w);
                            float* dst = cbase + int64_t(grow) * p.ldc + col;
                            if (p.beta != 0.f) {
                                float4 old = *reinterpret_cast<const float4*>(dst);
                                o.x += p.beta * old.x; o.y += p.beta * old.y; o.z += p.beta * old.z; o.w += p.beta * old.w;
                            }
                            if (p.c) *reinterpret_cast<float4*>(dst) = o;          // c == NULL: pre-activation kept as bf16 only
                            if (p.c16) st_bf16x4(p.c16 + int64_t(grow) * p.ld_c16 + col, o);
                            if (gbase || p.gelu16) {
                                const float4 gv = make_float4(gelu_tanh_epi(o.x), gelu_tanh_epi(o.y), gelu_tanh_epi(o.z), gelu_tanh_epi(o.w));
                                if (gbase) *reinterpret_cast<float4*>(gbase + int64_t(grow) * p.ldc + col) = gv;
                                if (p.gelu16) st_bf16x4(p.gelu16 + int64_t(grow) * p.ld_gelu16 + col, gv);
                            }
                        }
                    }
                } else if (FUSED && vec_ok && col + 4 <= p.n) {   // fused residual / GELU-gradient operands
                    float4 bb = make_float4(0.f, 0.f, 0.f, 0.f);
                    if (p.bias) bb = __ldg(reinterpret_cast<const float4*>(p.bias + col));
                    // every global operand of the chunk is requested before the first use, the staged accumulators are
                    // read back in one batch, and only then the math and the eight row stores run
                    float4 ad[8], gx[8], old[8], acc4[8];
                    if (abase) {
#pragma unroll
                        for (int q = 0; q < 8; ++q) {
                            const int grow = row_base + q * 4 + r4;
                            if (grow < p.m) ad[q] = __ldcs(reinterpret_cast<const float4*>(abase + int64_t(grow) * p.ld_addend + col));
                        }
                    }
                    if (xbase) {
#pragma unroll
                        for (int q = 0; q < 8; ++q) {
                            const int grow = row_base + q * 4 + r4;
                            if (grow < p.m) gx[q] = __ldcs(reinterpret_cast<const float4*>(xbase + int64_t(grow) * p.ld_gelu_grad_x + col));
                        }
                    }
                    if (p.beta != 0.f) {
#pragma unroll
                        for (int q = 0; q < 8; ++q) {
                            const int grow = row_base + q * 4 + r4;
                            if (grow < p.m) old[q] = *reinterpret_cast<const float4*>(cbase + int64_t(grow) * p.ldc + col);
                        }
                    }
#pragma unroll
                    for (int q = 0; q < 8; ++q) acc4[q] = *reinterpret_cast<const float4*>(stg + (q * 4 + r4) * kEpiPitch + c4 * 4);
#pragma unroll
                    for (int q = 0; q < 8; ++q) {
                        const int grow = row_base + q * 4 + r4;
                        if (grow < p.m) {
                            const float4 a = acc4[q];
                            float4 o = make_float4(a.x * p.alpha + bb.x, a.y * p.alpha + bb.y, a.z * p.alpha + bb.z, a.w * p.alpha + bb.w);
                            if (p.beta != 0.f) { o.x += p.beta * old[q].x; o.y += p.beta * old[q].y; o.z += p.beta * old[q].z; o.w += p.beta * old[q].w; }
                            if (abase) { o.x += ad[q].x; o.y += ad[q].y; o.z += ad[q].z; o.w += ad[q].w; }
                            if (xbase) {
                                o.x *= gelu_tanh_grad_f(gx[q].x); o.y *= gelu_tanh_grad_f(gx[q].y);
                                o.z *= gelu_tanh_grad_f(gx[q].z); o.w *= gelu_tanh_grad_f(gx[q].w);
                            }
                            *reinterpret_cast<float4*>(cbase + int64_t(grow) * p.ldc + col) = o;
                            if (p.c16) st_bf16x4(p.c16 + int64_t(grow) * p.ld_c16 + col, o);
                            if (gbase || p.gelu16) {
                                const float4 gv = make_float4(gelu_tanh_epi(o.x), gelu_tanh_epi(o.y), gelu_tanh_epi(o.z), gelu_tanh_epi(o.w));
                                if (gbase) *reinterpret_cast<float4*>(gbase + int64_t(grow) * p.ldc + col) = gv;
                                if (p.gelu16) st_bf16x4(p.gelu16 + int64_t(grow) * p.ld_gelu16 + col, gv);
                            }
                        }
                    }
                } else {
#pragma unroll 1
                    for (int q = 0; q < 8; ++q) {
                        const int r = q * 4 + r4;
                        const int grow = row_base + r;
                        if (grow >= p.m) continue;
                        for (int e = 0; e < 4; ++e) {
                            const int cc = col + e;
                            if (cc >= p.n) break;
                            float o = stg[r * kEpiPitch + c4 * 4 + e] * p.alpha;
                            if (p.bias) o += __ldg(p.bias + cc);
                            float* dst = cbase + int64_t(grow) * p.ldc + cc;
                            if (p.beta != 0.f) o += p.beta * *dst;
                            if (FUSED && abase) o += abase[int64_t(grow) * p.ld_addend + cc];
                            if (FUSED && xbase) o *= gelu_tanh_grad_f(xbase[int64_t(grow) * p.ld_gelu_grad_x + cc]);
                            if (p.c) *dst = o;
                            if (p.c16) p.c16[int64_t(grow) * p.ld_c16 + cc] = __float2bfloat16_rn(o);
                            if (gbase || p.gelu16) {
                                const float gv = gelu_tanh_epi(o);
                                if (gbase) gbase[int64_t(grow) * p.ldc + cc] = gv;
                                if (p.gelu16) p.gelu16[int64_t(grow) * p.ld_gelu16 + cc] = __float2bfloat16_rn(gv);
                            }
                        }
                    }
                }
                __syncwarp();                             // buffer is rewritten by the next chunk
            }
            }
            tc_fence_before();
            __syncwarp();
            if (lane == 0) {                              // one arrival per epilogue warp (of both CTAs when paired), at the leader
                if (CG == 2) mbar_arrive_remote(&tmem_empty[as], 0); else mbar_arrive(&tmem_empty[as]);
            }
        }
    }

    if (warp >= 2 && lane == 0) tma_store_wait_all();     // shared memory stays valid until the TMA has drained every box
    tc_fence_before();
    if (CG == 2) cluster_sync_all(); else __syncthreads();
    if (warp == 1) {
        tc_fence_after();
        if (CG == 2) asm volatile("tcgen05.dealloc.cta_group::2.sync.aligned.b32 %0, %1;" :: "r"(tmem_base), "r"(uint32_t(Cfg::tmem_cols)) : "memory");
        else asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" :: "r"(tmem_base), "r"(uint32_t(Cfg::tmem_cols)) : "memory");
    }
}

}  // namespace mft

// ---------------------------------------------------------------- host side
namespace mft {

struct MapKey {
    const void* ptr; uint64_t inner, rows, batch, ld, stride; uint32_t box_rows; uint32_t bf16;
    bool operator==(const MapKey& o) const {
        return ptr == o.ptr && inner == o.inner && rows == o.rows && batch == o.batch && ld == o.ld &&
               stride == o.stride && box_rows == o.box_rows && bf16 == o.bf16;
    }
};
struct MapKeyHash {
    size_t operator()(const MapKey& k) const {
        uint64_t h = reinterpret_cast<uint64_t>(k.ptr);
        auto mix = [&h](uint64_t v) { h ^= v + 0x9e3779b97f4a7c15ull + (h << 6) + (h >> 2); };
        mix(k.inner); mix(k.rows); mix(k.batch); mix(k.ld); mix(k.stride); mix(k.box_rows); mix(k.bf16);
        return size_t(h);
    }
};

// K-major operand [batch][rows][inner] (ld elements between rows): box = {32, box_rows, 1}, 128B swizzle.
// Encoding costs ~1 us on the host; the step re-uses the same buffers so descriptors are cached.
static bool make_operand_map(CUtensorMap* out, const void* ptr, uint64_t inner, uint64_t rows, uint64_t batch,
                             uint64_t ld, uint64_t stride, uint32_t box_rows, bool bf16) {
    static std::unordered_map<MapKey, CUtensorMap, MapKeyHash> cache;
    static std::mutex mu;
    MapKey key{ptr, inner, rows, batch, ld, stride, box_rows, bf16 ? 1u : 0u};
    {
        std::lock_guard<std::mutex> lk(mu);
        auto it = cache.find(key);
        if (it != cache.end()) { *out = it->second; return true; }
    }
    EncodeTiledFn enc = get_encode_fn();
    if (!enc) return false;
    alignas(64) CUtensorMap m;
    cuuint64_t dims[3] = {inner, rows, batch};
    if (batch == 1 || stride == 0) stride = rows * ld;
    const uint64_t es = bf16 ? 2 : 4;
    cuuint64_t strides[2] = {ld * es, stride * es};
    cuuint32_t box[3] = {uint32_t(bf16 ? 64 : 32), box_rows, 1};     // 128 bytes of K per row
    cuuint32_t estr[3] = {1, 1, 1};
    CUresult r = enc(&m, bf16 ? CU_TENSOR_MAP_DATA_TYPE_BFLOAT16 : CU_TENSOR_MAP_DATA_TYPE_TFLOAT32, 3, const_cast<void*>(ptr), dims, strides, box, estr,
                     CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_L2_256B,
                     CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
    if (r != CUDA_SUCCESS) return false;
    {
        std::lock_guard<std::mutex> lk(mu);
        if (cache.size() > 4096) cache.clear();
        cache.emplace(key, m);
    }
    *out = m;
    return true;
}

// Output tensor [batch][rows][cols] fp32 (ld elements between rows): box = {32 cols, 32 rows, 1}, 128B swizzle, used by
// the epilogue's cp.async.bulk.tensor stores (clipped at the edges, so ragged m / n need no predicates).
static bool make_store_map(CUtensorMap* out, const void* ptr, uint64_t cols, uint64_t rows, uint64_t batch, uint64_t ld, uint64_t stride, bool bf16 = false) {
    static std::unordered_map<MapKey, CUtensorMap, MapKeyHash> cache;
    static std::mutex mu;
    MapKey key{ptr, cols, rows, batch, ld, stride, 32u, bf16 ? 3u : 2u};
    {
        std::lock_guard<std::mutex> lk(mu);
        auto it = cache.find(key);
        if (it != cache.end()) { *out = it->second; return true; }
    }
    EncodeTiledFn enc = get_encode_fn();
    if (!enc) return false;
    alignas(64) CUtensorMap m;
    cuuint64_t dims[3] = {cols, rows, batch};
    if (batch == 1 || stride == 0) stride = rows * ld;
    const uint64_t es = bf16 ? 2 : 4;
    cuuint64_t strides[2] = {ld * es, stride * es};
    cuuint32_t box[3] = {bf16 ? 64u : 32u, 32, 1};                  // 128 bytes per box row either way
    cuuint32_t estr[3] = {1, 1, 1};
    CUresult r = enc(&m, bf16 ? CU_TENSOR_MAP_DATA_TYPE_BFLOAT16 : CU_TENSOR_MAP_DATA_TYPE_FLOAT32, 3, const_cast<void*>(ptr), dims, strides, box, estr, CU_TENSOR_MAP_INTERLEAVE_NONE,
                     CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_NONE, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
    if (r != CUDA_SUCCESS) return false;
    {
        std::lock_guard<std::mutex> lk(mu);
        if (cache.size() > 4096) cache.clear();
        cache.emplace(key, m);
    }
    *out = m;
    return true;
}

bool gemm_tc_eligible(const GemmParams& p) {
    if (p.trans_a != 0 || p.trans_b == 0) return false;                 // both operands must be K-major
    if (p.m < 1 || p.n < 1 || p.k < 1) return false;
    const int64_t am = p.bf16 ? 7 : 3;                                  // row pitch must be a multiple of 16 bytes
    if ((p.lda & am) || (p.ldb & am) || !aligned16(p.a) || !aligned16(p.b)) return false;
    if (p.batch > 1 && ((p.stride_a & am) || (p.stride_b & am))) return false;
    if (p.k2 > 0 && ((p.lda2 & am) || (p.ldb2 & am) || !aligned16(p.a2) || !aligned16(p.b2) || p.batch != 1)) return false;
    if (p.bf16) return true;
    // tiny problems are launch-bound either way; keep them on the SIMT kernel (exact fp32)
    if (int64_t(p.m) * p.n * p.batch < 64 * 64 || p.k < 32) return false;
    return true;
}

// Optional in-library timing of every tcgen05 GEMM launch (CUDA events on the launching stream): feeds the
// roofline figure bench.py reports (sum of algorithmic FLOPs / sum of launch durations over profiled steps).
namespace {
struct TcProfile {
    bool on = false;
    std::vector<std::pair<cudaEvent_t, cudaEvent_t>> pool;
    std::vector<double> flops;
    size_t used = 0;
    double total_flops = 0, total_ms = 0;
    long long launches = 0;
} g_prof;
void prof_drain() {
    for (size_t i = 0; i < g_prof.used; ++i) {
        cudaEventSynchronize(g_prof.pool[i].second);
        float ms = 0.f;
        if (cudaEventElapsedTime(&ms, g_prof.pool[i].first, g_prof.pool[i].second) == cudaSuccess) {
            g_prof.total_ms += ms;
            g_prof.total_flops += g_prof.flops[i];
            ++g_prof.launches;
        }
    }
    g_prof.used = 0;
}
}  // namespace

void gemm_tc_profile_enable(bool on) {
    if (!on && g_prof.on) prof_drain();
    if (on) { g_prof.total_flops = g_prof.total_ms = 0; g_prof.launches = 0; g_prof.used = 0; }
    g_prof.on = on;
}
void gemm_tc_profile_read(double* flops, double* ms, long long* launches) {
    prof_drain();
    if (flops) *flops = g_prof.total_flops;
    if (ms) *ms = g_prof.total_ms;
    if (launches) *launches = g_prof.launches;
}

template <int BN, int CG, bool BF16, bool FUSED, int EW, int FORM>
static int launch_tc_form(const GemmParams& p, cudaStream_t st, const CUtensorMap& ma, const CUtensorMap& mb, const CUtensorMap& ma2,
                          const CUtensorMap& mb2, const CUtensorMap& mc, const CUtensorMap& mg, const TcArgs& a);

template <int BN, int CG, bool BF16, bool FUSED, int EW>
static int launch_tc_impl(const GemmParams& p, cudaStream_t st) {
    using Cfg = TcCfg<BN, CG, EW>;
    constexpr int KE = BF16 ? 64 : 32;
    CUtensorMap ma, mb, ma2, mb2;
    if (!make_operand_map(&ma, p.a, p.k, p.m, p.batch, p.lda, p.stride_a, TBM, BF16) ||
        !make_operand_map(&mb, p.b, p.k, p.n, p.batch, p.ldb, p.stride_b, Cfg::b_rows, BF16))
        return set_error(3, "mft_gemm(tc): cuTensorMapEncodeTiled failed");
    if (p.k2 > 0) {
        if (!make_operand_map(&ma2, p.a2, p.k2, p.m, 1, p.lda2, 0, TBM, BF16) ||
            !make_operand_map(&mb2, p.b2, p.k2, p.n, 1, p.ldb2, 0, Cfg::b_rows, BF16))
            return set_error(3, "mft_gemm(tc): cuTensorMapEncodeTiled failed (segment 2)");
    } else { ma2 = ma; mb2 = mb; }
    // output boxes: C (and the GELU aux) leave through the TMA when the layout allows it and nothing has to be read back
    CUtensorMap mc, mg;
    // (the GELU aux output stays on the register/STG path: with one epilogue warp per scheduler the tanh chains of a whole
    //  32-value row per thread run with no other warp to hide their latency, measured 125 us vs 81 us on fc_in)
    static const bool tma_gelu = [] { const char* e = getenv("MFT_GEMM_TMA_GELU"); return e && e[0] == '1'; }();
    const bool bf16_only = p.c == nullptr;                          // the result leaves as bf16 alone (64-column TMA boxes)
    const bool gelu_aux = p.gelu_out || p.gelu_out_bf16;
    if (bf16_only) {
        if (!p.c_bf16 || FUSED || p.beta != 0.f || !aligned16(p.c_bf16) || (p.ldc_bf16 & 7) || p.batch != 1)
            return set_error(1, "mft_gemm(tc): a bf16-only result needs c_bf16 (16-byte aligned, pitch %% 8 == 0), beta = 0, batch 1 and no fused operands");
    }
    const bool c_ok = !bf16_only && !FUSED && p.beta == 0.f && aligned16(p.c) && (p.ldc & 3) == 0 && (p.batch == 1 || (p.stride_c & 3) == 0) &&
                      !p.gelu_out_bf16 && (!p.gelu_out || (tma_gelu && aligned16(p.gelu_out))) &&   // the box path writes no bf16 GELU copy
                      (!p.c_bf16 || (p.batch == 1 && (p.ldc_bf16 & 7) == 0 && aligned16(p.c_bf16)));
    static const bool tma_store_on = [] { const char* e = getenv("MFT_GEMM_TMA_STORE"); return !(e && e[0] == '0'); }();
    bool tma_c = c_ok && tma_store_on;
    if (tma_c) {
        tma_c = make_store_map(&mc, p.c, p.n, p.m, p.batch, p.ldc, p.stride_c) &&
                (!p.gelu_out || make_store_map(&mg, p.gelu_out, p.n, p.m, p.batch, p.ldc, p.stride_c));
    }
    // bf16-only result through boxes: plain, or (eight-warp form) together with a bf16-only GELU aux output
    const bool box_gelu = bf16_only && EW == 8 && !p.gelu_out && p.gelu_out_bf16 && aligned16(p.gelu_out_bf16) && (p.ld_gelu_out_bf16 & 7) == 0;
    const bool box16 = bf16_only && (!gelu_aux || box_gelu);      // otherwise the register path writes the bf16 arrays
    if (box16) {
        if (!make_store_map(&mc, p.c_bf16, p.n, p.m, 1, p.ldc_bf16, 0, true)) return set_error(3, "mft_gemm(tc): cuTensorMapEncodeTiled failed (bf16 result)");
        mg = mc;
        if (box_gelu && !make_store_map(&mg, p.gelu_out_bf16, p.n, p.m, 1, p.ld_gelu_out_bf16, 0, true))
            return set_error(3, "mft_gemm(tc): cuTensorMapEncodeTiled failed (bf16 GELU result)");
        tma_c = true;
    } else if (!tma_c) { mc = ma; mg = ma; } else if (!p.gelu_out) mg = mc;
    TcArgs a;
    a.tma_c = tma_c ? 1 : 0;
    a.tma_c16 = box16 ? 1 : 0;
    a.tma_g16 = (box16 && box_gelu) ? 1 : 0;
    a.c16 = box16 ? nullptr : static_cast<__nv_bfloat16*>(p.c_bf16); a.ld_c16 = p.ldc_bf16;
    a.m = p.m; a.n = p.n; a.batch = p.batch;
    a.kb1 = int(ceil_div(p.k, KE)); a.kb2 = p.k2 > 0 ? int(ceil_div(p.k2, KE)) : 0;
    a.c = p.c; a.ldc = p.ldc; a.stride_c = p.stride_c; a.gelu_out = p.gelu_out; a.bias = p.bias;
    a.alpha = p.alpha; a.beta = p.beta;
    a.addend = p.addend; a.ld_addend = p.ld_addend; a.gelu_grad_x = p.gelu_grad_x; a.ld_gelu_grad_x = p.ld_gelu_grad_x;
    a.gelu16 = static_cast<__nv_bfloat16*>(p.gelu_out_bf16); a.ld_gelu16 = p.ld_gelu_out_bf16;
    a.tiles_m = int(ceil_div(p.m, TBM)); a.tiles_n = int(ceil_div(p.n, BN));
    {
        // raster groups (tile_coords): 16 m-units (4096 rows with CTA pairs) keep a group's A slab and one sweep of B inside
        // L2; a very wide B (the vocabulary) is swept fewer times with 32. MFT_TC_GROUP_M overrides (experiments).
        static const int forced = [] { const char* e = std::getenv("MFT_TC_GROUP_M"); return e && *e ? std::atoi(e) : 0; }();
        const int tiles_mu = int(ceil_div(a.tiles_m, CG));
        const int gm = forced > 0 ? forced : (a.tiles_n >= 64 ? 32 : 16);
        a.group_m = std::max(1, std::min(gm, tiles_mu));
    }
    // compile-time epilogue form (see gemm_tc_kernel): the plain box forms get their own instantiation
    if (!FUSED) {
        if (a.tma_c16 && a.tma_g16) { if (EW == 8) return launch_tc_form<BN, CG, BF16, false, EW, (EW == 8 ? 3 : 0)>(p, st, ma, mb, ma2, mb2, mc, mg, a); }
        else if (a.tma_c16) return launch_tc_form<BN, CG, BF16, false, EW, 2>(p, st, ma, mb, ma2, mb2, mc, mg, a);
        else if (a.tma_c && !a.c16 && !p.gelu_out && !p.gelu_out_bf16) return launch_tc_form<BN, CG, BF16, false, EW, 1>(p, st, ma, mb, ma2, mb2, mc, mg, a);
    }
    return launch_tc_form<BN, CG, BF16, FUSED, EW, 0>(p, st, ma, mb, ma2, mb2, mc, mg, a);
}

template <int BN, int CG, bool BF16, bool FUSED, int EW, int FORM>
static int launch_tc_form(const GemmParams& p, cudaStream_t st, const CUtensorMap& ma, const CUtensorMap& mb, const CUtensorMap& ma2,
                          const CUtensorMap& mb2, const CUtensorMap& mc, const CUtensorMap& mg, const TcArgs& a) {
    using Cfg = TcCfg<BN, CG, EW>;
    static bool attr_set[kMaxDevices] = {};
    if (first_use_on_device(attr_set)) {
        cudaError_t e = cudaFuncSetAttribute(gemm_tc_kernel<BN, CG, BF16, FUSED, EW, FORM>, cudaFuncAttributeMaxDynamicSharedMemorySize, Cfg::smem_bytes);
        if (e != cudaSuccess) return set_error(2, "mft_gemm(tc): smem attribute: %s", cudaGetErrorString(e));
    }
    const int64_t units = ceil_div(a.tiles_m, CG) * a.tiles_n * a.batch;    // CTAs, or CTA pairs
    const int max_units = kNumSMs / CG;
    const int grid = int(units < max_units ? units : max_units) * CG;
    size_t slot = 0;
    if (g_prof.on) {
        if (g_prof.used == g_prof.pool.size()) {
            if (g_prof.pool.size() >= 8192) prof_drain();
            else {
                cudaEvent_t e0, e1;
                cudaEventCreate(&e0); cudaEventCreate(&e1);
                g_prof.pool.emplace_back(e0, e1);
                g_prof.flops.push_back(0.0);
            }
        }
        slot = g_prof.used++;
        g_prof.flops[slot] = 2.0 * double(p.m) * double(p.n) * double(p.k + p.k2) * double(p.batch);
        cudaEventRecord(g_prof.pool[slot].first, st);
    }
    cudaLaunchConfig_t cfg;
    memset(&cfg, 0, sizeof(cfg));
    cfg.gridDim = dim3(unsigned(grid));
    cfg.blockDim = dim3(64 + 32 * EW);
    cfg.dynamicSmemBytes = Cfg::smem_bytes;
    cfg.stream = st;
    cudaLaunchAttribute attr[2];
    attr[0].id = cudaLaunchAttributeClusterDimension;
    attr[0].val.clusterDim.x = CG; attr[0].val.clusterDim.y = 1; attr[0].val.clusterDim.z = 1;
    attr[1].id = cudaLaunchAttributeProgrammaticStreamSerialization;      // the kernel starts with pdl_sync() (common.cuh)
    attr[1].val.programmaticStreamSerializationAllowed = 1;
    cfg.attrs = attr;
    cfg.numAttrs = pdl_enabled() ? 2 : 1;
    cudaError_t e = cudaLaunchKernelEx(&cfg, gemm_tc_kernel<BN, CG, BF16, FUSED, EW, FORM>, ma, mb, ma2, mb2, mc, mg, a);
    if (g_prof.on) cudaEventRecord(g_prof.pool[slot].second, st);
    add_launches(1);
    if (e == cudaSuccess) e = cudaGetLastError();
    if (e != cudaSuccess) return set_error(2, "mft_gemm(tc): %s", cudaGetErrorString(e));
    return 0;
}

template <int BN, int CG, bool BF16>
static int launch_tc(const GemmParams& p, cudaStream_t st) {
    if (p.addend || p.gelu_grad_x) return launch_tc_impl<BN, CG, BF16, true, kEpiWarpsDefault>(p, st);
    // The GELU aux output leaves through registers and st.global (tanh chains, two bf16 arrays): with one epilogue warp per
    // scheduler that epilogue, not the main loop, sets the tile time (fc_in: 82.7 us against 34 us for the plain product);
    // eight epilogue warps, two per TMEM lane group, halve it (57.8 us). The box (TMA) epilogues stay at four.
    if ((p.gelu_out || p.gelu_out_bf16) && BN >= 64) return launch_tc_impl<BN, CG, BF16, false, 8>(p, st);
    return launch_tc_impl<BN, CG, BF16, false, kEpiWarpsDefault>(p, st);
}

static bool env_two_cta() {
    const char* e = getenv("MFT_GEMM_2CTA");
    return !(e && e[0] == '0');
}

template <bool BF16>
static int dispatch_tc(const GemmParams& p, cudaStream_t st) {
    static const bool two_cta = env_two_cta();
    const int64_t tiles_m = ceil_div(p.m, TBM);
    const int64_t tm = tiles_m * p.batch;
    // Paired CTAs (cta_group::2) whenever there are at least two row tiles to pair and the output is wide enough.
    if (two_cta && tiles_m >= 2 && p.n > 64) {
        const int64_t pairs = ceil_div(tiles_m, 2) * p.batch;
#ifdef MFT_TC_FORCE_BN          /* variant builds only (tools/build_variants.sh): one tile width for every paired launch */
        return launch_tc<MFT_TC_FORCE_BN, 2, BF16>(p, st);
#endif
        if (p.n <= 128 || pairs * ceil_div(p.n, 256) < kNumSMs / 4) return launch_tc<128, 2, BF16>(p, st);
        return launch_tc<256, 2, BF16>(p, st);
    }
    // Single-CTA tiles: widest tile that still gives every SM work; narrow N uses narrow tiles.
    if (p.n <= 64 || tm * ceil_div(p.n, 128) < kNumSMs / 2) return launch_tc<64, 1, BF16>(p, st);
    if (p.n <= 128 || tm * ceil_div(p.n, 256) < kNumSMs) return launch_tc<128, 1, BF16>(p, st);
    return launch_tc<256, 1, BF16>(p, st);
}

int launch_gemm_tc(const GemmParams& p, cudaStream_t st) {
    return p.bf16 ? dispatch_tc<true>(p, st) : dispatch_tc<false>(p, st);
}

}  // namespace mft
