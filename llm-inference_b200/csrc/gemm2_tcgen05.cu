// Two-CTA tcgen05 GEMM for the prefill linears (launchLinearGemm with many token rows, linear.cu:39-104):
//   D[256, 256] (fp32, TMEM of two SMs) += A[256, 64] * B[256, 64]^T per K step, fp16 operands, K-major.
// A cluster of two CTAs (one TPC) owns a 256 x 256 output tile: each CTA loads ITS 128 rows of A and ITS
// 128 rows of B (half of the N extent) and keeps 128 rows x 256 columns of the accumulator in its own
// TMEM; the leader CTA's single MMA thread issues tcgen05.mma.cta_group::2 (M = 256), which reads both
// CTAs' shared memory. Against the one-CTA kernel (gemm_tcgen05.cu: 48 KB of operands per 512 tensor
// clocks, 4 stages) a stage is 32 KB per SM for the same 512 clocks: the smem read rate per SM drops from
// 96 to 64 B/clk and the ring holds 6 stages of look-ahead instead of 4.
//   warp 0    TMA producer (both CTAs): its own A / B boxes, completion signalled on the LEADER's full barrier
//   warp 1    TMEM allocation (both CTAs); MMA issue (leader only); tcgen05.commit multicast to both CTAs
//   warps 2-5 epilogue (both CTAs): own 128 rows from own TMEM, then a remote arrive on the leader's
//             accumulator-empty barrier
// GLU: as in gemm_tcgen05.cu — the leader's B half is 128 gate rows, the peer's the matching 128 up rows.
#include <cstdio>
#include "common.cuh"
#include "launch_count.h"
#include "tc_common.cuh"

namespace b200 {

constexpr int k2Threads = 192;
constexpr int k2BK = 64;
constexpr int k2Stages = 6;
constexpr uint32_t k2ABytes = 128 * k2BK * 2, k2BBytes = 128 * k2BK * 2, k2StageBytes = k2ABytes + k2BBytes;
constexpr uint32_t k2OffBar = k2Stages * k2StageBytes;                  // full[6], empty[6], acc_full[2], acc_empty[2]
constexpr uint32_t k2OffSlot = k2OffBar + (2 * k2Stages + 4) * 8;
constexpr uint32_t kStgBytes = 32 * 128;                                // epilogue staging per warp: 32 rows x 64 fp16
constexpr uint32_t k2OffStg = (k2OffSlot + 16 + 127) & ~127u;
constexpr uint32_t k2Smem = k2OffStg + 4 * kStgBytes;
static_assert(k2Smem <= 232448, "pair GEMM exceeds the 227 KB shared-memory limit");

struct Gemm2Params {
    __half* C;          // [tokens, ldc]
    const __half* bias; // [N] or nullptr (not with GLU)
    const __half* residual;   // [tokens, ldc] or nullptr: C = half(half(acc + bias) + residual), launchAddResidual folded in
    int tokens, n_out;  // valid rows of A; valid output features (GLU: inter)
    int ldc, k_blocks;
    int m_tiles, n_tiles;   // tiles of 256 rows x (256 features | GLU: 128 features)
    int group_m;            // tile order: bands of group_m row tiles swept over all column tiles (0: column-major over all rows)
    int wide_full;          // wide kernel: units [0, wide_full) are 256 x 512, the rest is run as single 256 x 256 column tiles
};

// unit u -> (row tile, column tile). The clusters resident at one time work on ~74 consecutive units: with bands of
// group_m row tiles they share group_m A tiles (L2-resident for the whole sweep of the band) and ~74 / group_m B tiles,
// instead of ALL row tiles x 2 column tiles (A = 64 MB at 8192 tokens re-streamed for every pair of column tiles).
__device__ __forceinline__ void tile_of_unit(const Gemm2Params& p, int n_tiles, int u, int& mt, int& nt) {
    if (p.group_m <= 0) { mt = u % p.m_tiles; nt = u / p.m_tiles; return; }
    const int per_band = p.group_m * n_tiles;
    const int band = u / per_band, r = u - band * per_band;
    const int rows = min(p.group_m, p.m_tiles - band * p.group_m);
    nt = r / rows;
    mt = band * p.group_m + (r - nt * rows);
}

// One thread's row of one accumulator tile: TMEM (taddr: this warp's lane quarter, the tile's first column) -> fp16 -> C.
// nt: column tile (256 features, GLU: 128 features with gate in columns [0,128) and up in [128,256) of the accumulator).
template <bool GLU>
__device__ __forceinline__ void store_tile_row(const Gemm2Params& p, uint32_t taddr, int row, int nt) {
    if (GLU) {
#pragma unroll 1
        for (int c0 = 0; c0 < 128; c0 += 32) {
            uint32_t g[32], uu[32];
            tmem_ld_32x32(taddr + (uint32_t)c0, g);
            tmem_ld_32x32(taddr + (uint32_t)(128 + c0), uu);
            const int col0 = nt * 128 + c0;
            if (row < p.tokens && col0 < p.n_out) {
                __half* dst = p.C + (size_t)row * p.ldc + col0;
#pragma unroll
                for (int q = 0; q < 4; ++q) {
                    float f[8];
#pragma unroll
                    for (int e = 0; e < 8; ++e) {
                        const float gv = __half2float(__float2half_rn(__uint_as_float(g[q * 8 + e])));
                        const float uv = __half2float(__float2half_rn(__uint_as_float(uu[q * 8 + e])));
                        const float sv = __half2float(__float2half_rn(gv / (1.f + expf(-gv))));
                        f[e] = sv * uv;
                    }
                    if (col0 + q * 8 + 8 <= p.n_out) st16(dst + q * 8, Unpack<__half>::put(f));
                }
            }
        }
    } else {
#pragma unroll 1
        for (int c0 = 0; c0 < 256; c0 += 32) {
            uint32_t v[32];
            tmem_ld_32x32(taddr + (uint32_t)c0, v);
            const int col0 = nt * 256 + c0;
            if (row < p.tokens && col0 < p.n_out) {
                __half* dst = p.C + (size_t)row * p.ldc + col0;
                if (col0 + 32 <= p.n_out && (p.ldc & 7) == 0) {
#pragma unroll
                    for (int q = 0; q < 4; ++q) {
                        float f[8];
#pragma unroll
                        for (int e = 0; e < 8; ++e) {
                            f[e] = __uint_as_float(v[q * 8 + e]);
                            if (p.bias) f[e] += __half2float(p.bias[col0 + q * 8 + e]);
                        }
                        if (p.residual) {   // the linear's output is rounded to fp16 first, as the unfused pair does
                            float r[8];
                            Unpack<__half>::get(ld16(p.residual + (size_t)row * p.ldc + col0 + q * 8), r);
#pragma unroll
                            for (int e = 0; e < 8; ++e) f[e] = __half2float(__float2half_rn(f[e])) + r[e];
                        }
                        st16(dst + q * 8, Unpack<__half>::put(f));
                    }
                } else {
                    // (fully unrolled with a predicate: a run-time index would put v[] into local memory - with the
                    // 227 KB carve-out there is no L1 behind it, and the whole epilogue crawled at L2 latency: STACK 128 -> 0)
#pragma unroll
                    for (int j = 0; j < 32; ++j) {
                        if (col0 + j >= p.n_out) continue;
                        float f = __uint_as_float(v[j]);
                        if (p.bias) f += __half2float(p.bias[col0 + j]);
                        if (p.residual) f = __half2float(__float2half_rn(f)) + __half2float(p.residual[(size_t)row * p.ldc + col0 + j]);
                        dst[j] = __float2half_rn(f);
                    }
                }
            }
        }
    }
}

// The same tile part for a whole WARP (32 rows x the tile's columns), with the stores coalesced: every thread owns an
// accumulator ROW in TMEM, so storing straight from registers makes each store instruction touch 32 rows x 16 bytes - 8192
// sixteen-byte transactions per 128 x 512 CTA tile, ~20 k clocks that the 256 x 512 kernel cannot hide. Here 64 output
// columns of the warp's 32 rows are rounded to fp16, staged in 4 KB of shared memory (16-byte chunks XOR-swizzled by row:
// conflict-free both ways), then written as 4 rows x 128 contiguous bytes per instruction. Rounding points as above.
__device__ __forceinline__ void stg_st_v4(uint32_t addr, uint32_t a, uint32_t b, uint32_t c, uint32_t d) {
    asm volatile("st.shared.v4.b32 [%0], {%1, %2, %3, %4};" ::"r"(addr), "r"(a), "r"(b), "r"(c), "r"(d) : "memory");
}
__device__ __forceinline__ uint4 stg_ld_v4(uint32_t addr) {
    uint4 v;
    asm volatile("ld.shared.v4.b32 {%0, %1, %2, %3}, [%4];" : "=r"(v.x), "=r"(v.y), "=r"(v.z), "=r"(v.w) : "r"(addr) : "memory");
    return v;
}
__device__ __forceinline__ uint32_t pack_half2(float lo, float hi) {
    const __half2 h = __floats2half2_rn(lo, hi);
    return *reinterpret_cast<const uint32_t*>(&h);
}

template <bool GLU>
__device__ __forceinline__ void store_tile_warp(const Gemm2Params& p, uint32_t taddr, int row0, int nt, uint32_t stg, int lane) {
    constexpr int kOutCols = GLU ? 128 : 256;
    const int tile_col0 = nt * kOutCols;
    const bool vec_ok = (p.ldc & 7) == 0;
#pragma unroll 1
    for (int c0 = 0; c0 < kOutCols; c0 += 64) {
#pragma unroll
        for (int sub = 0; sub < 2; ++sub) {
            const int cc = c0 + sub * 32;
            uint32_t h[16];
            if (GLU) {
                uint32_t g[32], uu[32];
                tmem_ld_32x32(taddr + (uint32_t)cc, g);
                tmem_ld_32x32(taddr + (uint32_t)(128 + cc), uu);
#pragma unroll
                for (int e = 0; e < 32; e += 2) {
                    float o[2];
#pragma unroll
                    for (int t = 0; t < 2; ++t) {
                        const float gv = __half2float(__float2half_rn(__uint_as_float(g[e + t])));
                        const float uv = __half2float(__float2half_rn(__uint_as_float(uu[e + t])));
                        const float sv = __half2float(__float2half_rn(gv / (1.f + expf(-gv))));
                        o[t] = sv * uv;
                    }
                    h[e >> 1] = pack_half2(o[0], o[1]);
                }
            } else {
                uint32_t v[32];
                tmem_ld_32x32(taddr + (uint32_t)cc, v);
#pragma unroll
                for (int e = 0; e < 32; e += 2) {
                    float f0 = __uint_as_float(v[e]), f1 = __uint_as_float(v[e + 1]);
                    if (p.bias) {
                        const int col = tile_col0 + cc + e;
                        if (col < p.n_out) f0 += __half2float(p.bias[col]);
                        if (col + 1 < p.n_out) f1 += __half2float(p.bias[col + 1]);
                    }
                    h[e >> 1] = pack_half2(f0, f1);
                }
            }
#pragma unroll
            for (int q = 0; q < 4; ++q)
                stg_st_v4(stg + (uint32_t)lane * 128u + (uint32_t)((((sub * 4 + q) ^ (lane & 7))) << 4), h[4 * q], h[4 * q + 1],
                          h[4 * q + 2], h[4 * q + 3]);
        }
        __syncwarp();
#pragma unroll
        for (int j = 0; j < 8; ++j) {
            const int r = j * 4 + (lane >> 3), ch = lane & 7;
            const uint4 val = stg_ld_v4(stg + (uint32_t)r * 128u + (uint32_t)((ch ^ (r & 7)) << 4));
            const int row = row0 + r, col = tile_col0 + c0 + ch * 8;
            if (row < p.tokens && col < p.n_out) {
                __half* dst = p.C + (size_t)row * p.ldc + col;
                Pack16 pk;
                *reinterpret_cast<uint4*>(&pk) = val;
                if (vec_ok && col + 8 <= p.n_out) {
                    if (!GLU && p.residual) {   // the linear's output was rounded to fp16 first, as the unfused pair does
                        float a[8], rr[8];
                        Unpack<__half>::get(pk, a);
                        Unpack<__half>::get(ld16(p.residual + (size_t)row * p.ldc + col), rr);
#pragma unroll
                        for (int e = 0; e < 8; ++e) a[e] += rr[e];
                        pk = Unpack<__half>::put(a);
                    }
                    st16(dst, pk);
                } else {
                    float a[8];
                    Unpack<__half>::get(pk, a);
#pragma unroll
                    for (int e = 0; e < 8; ++e) {
                        if (col + e >= p.n_out) continue;
                        float f = a[e];
                        if (!GLU && p.residual) f += __half2float(p.residual[(size_t)row * p.ldc + col + e]);
                        dst[e] = __float2half_rn(f);
                    }
                }
            }
        }
        __syncwarp();
    }
}

// QUAD: a cluster of FOUR CTAs = two MMA pairs on the same 256 rows of A and two neighbouring column tiles. Every CTA
// loads a 64-row box of A and multicasts it to its counterpart in the other pair, so a cluster reads A from L2 once
// instead of twice: 48 KB instead of 64 KB per pair and K step (the cuBLAS kernel of this shape, nvjet ..._2x2_2cta, moves
// the same 4.83 GB against the pair kernel's 6.44 GB at 8192 x 12288 x 4096). On a power-capped B200 the bytes moved set
// the clock. A stage is free when BOTH pairs' MMAs have read it (empty barriers count two commits); each pair
// keeps its own full barriers, accumulators and epilogue.
template <bool GLU, bool QUAD>
__global__ void __launch_bounds__(k2Threads, 1)
gemm2_tcgen05_kernel(const __grid_constant__ CUtensorMap map_a, const __grid_constant__ CUtensorMap map_b, const Gemm2Params p) {
    extern __shared__ __align__(1024) unsigned char smem[];
    uint64_t* full_bar = reinterpret_cast<uint64_t*>(smem + k2OffBar);
    uint64_t* empty_bar = full_bar + k2Stages;
    uint64_t* acc_full = empty_bar + k2Stages;   // [2]
    uint64_t* acc_empty = acc_full + 2;          // [2], used in the leader only
    uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(smem + k2OffSlot);
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const uint32_t crank = cluster_ctarank();
    const uint32_t rank = crank & 1u;            // position in the MMA pair, 0 = pair leader
    const uint32_t pair = QUAD ? crank >> 1 : 0u;
    const uint32_t leader = crank & ~1u;         // cluster rank of this pair's leader
    constexpr int CS = QUAD ? 4 : 2;
    if (threadIdx.x == 0) {
        if (smem_u32(smem) & 1023u) __trap();
        for (int s = 0; s < k2Stages; ++s) {
            mbar_init(&full_bar[s], 1);
            mbar_init(&empty_bar[s], QUAD ? 2 : 1);   // one commit per pair that reads the stage
        }
        for (int a = 0; a < 2; ++a) {
            mbar_init(&acc_full[a], 1);
            mbar_init(&acc_empty[a], 8);   // 4 epilogue warps x 2 CTAs
        }
        mbar_fence_init();
        tma_prefetch_desc(&map_a);
        tma_prefetch_desc(&map_b);
    }
    if (warp == 1) tmem_alloc_pair(tmem_slot, 512);
    tc_fence_before();
    __syncthreads();
    cluster_sync_all();   // the peer's barriers exist before anything is signalled on them
    tc_fence_after();
    const uint32_t tmem_base = *tmem_slot;
    pdl_launch_dependents();
    pdl_wait();

    const int n_tiles = QUAD ? p.n_tiles >> 1 : p.n_tiles;   // QUAD: a unit is two neighbouring column tiles
    const int units = p.m_tiles * n_tiles;
    const int cluster_id = blockIdx.x / CS, n_clusters = gridDim.x / CS;

    if (warp == 0) {
        // ================= TMA producer (both CTAs) =================
        int stage = 0;
        uint32_t phase = 0;
        for (int u = cluster_id; u < units; u += n_clusters) {
            int mt, nt;
            tile_of_unit(p, n_tiles, u, mt, nt);
            if (QUAD) nt = 2 * nt + (int)pair;
            const int a_row = mt * 256 + (int)rank * 128 + (QUAD ? (int)pair * 64 : 0);
            const int b_row = GLU ? (rank == 0 ? nt * 128 : p.n_out + nt * 128) : nt * 256 + (int)rank * 128;
            for (int kb = 0; kb < p.k_blocks; ++kb) {
                mbar_wait_bounded(&empty_bar[stage], phase ^ 1);
                if (lane == 0) {
                    unsigned char* sa = smem + (size_t)stage * k2StageBytes;
                    const uint32_t leader_full = mapa_shared(smem_u32(&full_bar[stage]), leader);
                    if (rank == 0) mbar_expect_tx(&full_bar[stage], 2 * k2StageBytes);   // both CTAs' boxes
                    if (QUAD)   // my 64 rows of this position's A half, to me and to my counterpart in the other pair
                        tma_load_2d_pair_multicast(sa + pair * (k2ABytes / 2), &map_a, kb * k2BK, a_row, leader_full,
                                                   (uint16_t)((1u << rank) | (1u << (rank + 2))));
                    else
                        tma_load_2d_pair(sa, &map_a, kb * k2BK, a_row, leader_full);
                    tma_load_2d_pair(sa + k2ABytes, &map_b, kb * k2BK, b_row, leader_full);
                }
                __syncwarp();
                if (++stage == k2Stages) { stage = 0; phase ^= 1; }
            }
        }
    } else if (warp == 1) {
        // ================= MMA issuer (leader CTA only) =================
        if (rank == 0) {
            constexpr uint32_t idesc = make_idesc_f16(256, 256);
            int stage = 0, acc = 0;
            uint32_t phase = 0, acc_phase = 0;
            for (int u = cluster_id; u < units; u += n_clusters) {
                mbar_wait_bounded(&acc_empty[acc], acc_phase ^ 1);
                tc_fence_after();
                const uint32_t d_tmem = tmem_base + (uint32_t)(acc * 256);
                for (int kb = 0; kb < p.k_blocks; ++kb) {
                    mbar_wait_bounded(&full_bar[stage], phase);
                    tc_fence_after();
                    if (lane == 0) {
                        const uint32_t sa = smem_u32(smem + (size_t)stage * k2StageBytes);
                        const uint64_t da = make_kmajor_sw128_desc(sa);
                        const uint64_t db = make_kmajor_sw128_desc(sa + k2ABytes);
#pragma unroll
                        for (int k = 0; k < k2BK / kUmmaK; ++k)
                            umma_f16_pair(d_tmem, da + (uint64_t)(2 * k), db + (uint64_t)(2 * k), idesc, (kb > 0 || k > 0) ? 1u : 0u);
                        umma_commit_pair(&empty_bar[stage], QUAD ? 0xF : 3);         // the producers may refill
                        if (kb == p.k_blocks - 1) umma_commit_pair(&acc_full[acc], (uint16_t)(3u << (2 * pair)));   // both epilogues of the pair may read
                    }
                    __syncwarp();
                    if (++stage == k2Stages) { stage = 0; phase ^= 1; }
                }
                acc ^= 1;
                if (acc == 0) acc_phase ^= 1;
            }
        }
    } else {
        // ================= epilogue warps (both CTAs) =================
        const int quarter = warp & 3;
        int acc = 0;
        uint32_t acc_phase = 0;
        for (int u = cluster_id; u < units; u += n_clusters) {
            int mt, nt;
            tile_of_unit(p, n_tiles, u, mt, nt);
            if (QUAD) nt = 2 * nt + (int)pair;
            mbar_wait_bounded(&acc_full[acc], acc_phase);
            tc_fence_after();
            const int row = mt * 256 + (int)rank * 128 + quarter * 32 + lane;
            const uint32_t taddr = tmem_base + ((uint32_t)(quarter * 32) << 16) + (uint32_t)(acc * 256);
            (void)row;
            store_tile_warp<GLU>(p, taddr, mt * 256 + (int)rank * 128 + quarter * 32, nt,
                                 smem_u32(smem + k2OffStg + (uint32_t)(warp - 2) * kStgBytes), lane);
            tc_fence_before();
            __syncwarp();
            if (lane == 0) mbar_arrive_cluster(mapa_shared(smem_u32(&acc_empty[acc]), leader));
            acc ^= 1;
            if (acc == 0) acc_phase ^= 1;
        }
    }
    tc_fence_before();
    __syncthreads();
    cluster_sync_all();   // neither CTA leaves while the other may still signal its barriers / read its smem
    if (warp == 1) {
        tc_fence_after();
        tmem_dealloc_pair(tmem_base, 512);
    }
}

// ------------------------------------------------------------------------------------------
// WIDE: the pair owns a 256 x 512 output tile = TWO neighbouring column tiles on the same 256 rows of A. Per K step a CTA
// loads its 128 rows of A ONCE (16 KB) and its 128-row halves of both B tiles (2 x 16 KB) and the leader issues the MMAs
// of both accumulators from the same A descriptors. Against the 256 x 256 tile: 24 instead of 32 KB loaded per 128 x 256 x
// 64 MAC block (L2 -> SM and TMA smem writes -25 %), shared-memory traffic (TMA writes + MMA operand reads) 112 KB per
// 1024 tensor clocks instead of 128 KB - the 256 x 256 kernel asks for exactly the 128 B/clk an SM's shared memory
// delivers, which is where its 83 % tensor-pipe ceiling comes from (ncu: profiles/r02_gemm_vs_cublas_full.md; cuBLAS'
// kernel for this shape has the same 48 KB stages x 4). Price: both accumulators fill all 512 TMEM columns, so the
// epilogue of a tile is not hidden behind the next tile's MMAs any more; eight epilogue warps (four per accumulator)
// keep it short, and the first K steps of the next tile's accumulator 0 are issued as soon as ITS half is drained.
//   warp 0     TMA producer (both CTAs)        warp 1     TMEM allocation; MMA issue (leader)
//   warps 2-5  epilogue of accumulator 0       warps 6-9  epilogue of accumulator 1
// ------------------------------------------------------------------------------------------
constexpr int kwThreads = 320;
constexpr int kwStages = 4;
constexpr uint32_t kwStageBytes = k2ABytes + 2 * k2BBytes;            // 48 KB
constexpr uint32_t kwOffBar = kwStages * kwStageBytes;                // full[4], empty[4], acc_full[2], acc_empty[2]
constexpr uint32_t kwOffSlot = kwOffBar + (2 * kwStages + 4) * 8;
constexpr uint32_t kwOffStg = (kwOffSlot + 16 + 127) & ~127u;
constexpr uint32_t kwSmem = kwOffStg + 8 * kStgBytes;
static_assert(kwSmem <= 232448, "wide GEMM exceeds the 227 KB shared-memory limit");

template <bool GLU>
__global__ void __launch_bounds__(kwThreads, 1)
gemm2w_tcgen05_kernel(const __grid_constant__ CUtensorMap map_a, const __grid_constant__ CUtensorMap map_b, const Gemm2Params p) {
    extern __shared__ __align__(1024) unsigned char smem[];
    uint64_t* full_bar = reinterpret_cast<uint64_t*>(smem + kwOffBar);
    uint64_t* empty_bar = full_bar + kwStages;
    uint64_t* acc_full = empty_bar + kwStages;   // [2]: accumulator h of the current unit is complete
    uint64_t* acc_empty = acc_full + 2;          // [2], leader only: accumulator h has been drained by both CTAs
    uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(smem + kwOffSlot);
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const uint32_t rank = cluster_ctarank();     // 0 = leader
    if (threadIdx.x == 0) {
        if (smem_u32(smem) & 1023u) __trap();
        for (int s = 0; s < kwStages; ++s) {
            mbar_init(&full_bar[s], 1);
            mbar_init(&empty_bar[s], 1);
        }
        for (int a = 0; a < 2; ++a) {
            mbar_init(&acc_full[a], 1);
            mbar_init(&acc_empty[a], 8);   // 4 epilogue warps x 2 CTAs
        }
        mbar_fence_init();
        tma_prefetch_desc(&map_a);
        tma_prefetch_desc(&map_b);
    }
    if (warp == 1) tmem_alloc_pair(tmem_slot, 512);
    tc_fence_before();
    __syncthreads();
    cluster_sync_all();
    tc_fence_after();
    const uint32_t tmem_base = *tmem_slot;
    pdl_launch_dependents();
    pdl_wait();

    const int n_tiles = p.n_tiles >> 1;          // a unit is two neighbouring column tiles
    const int units = p.m_tiles * n_tiles;
    const int cluster_id = blockIdx.x >> 1, n_clusters = gridDim.x >> 1;
    // Work items v: the first wide_full are whole units; the units after them are split into their two column tiles so that
    // the last, partly filled wave costs half a unit time (256 units on 74 clusters: 3 waves + 68 single tiles = 3.5 unit
    // times instead of 4). A single-tile item runs the same pipeline with one B box, one accumulator's MMAs and one store.
    const int items = p.wide_full + 2 * (units - p.wide_full);
    auto item = [&](int v, int& u, int& hsel) {
        if (v < p.wide_full) { u = v; hsel = -1; }
        else { u = p.wide_full + ((v - p.wide_full) >> 1); hsel = (v - p.wide_full) & 1; }
    };

    if (warp == 0) {
        // ================= TMA producer (both CTAs) =================
        int stage = 0;
        uint32_t phase = 0;
        for (int v = cluster_id; v < items; v += n_clusters) {
            int u, hsel, mt, nt;
            item(v, u, hsel);
            tile_of_unit(p, n_tiles, u, mt, nt);
            nt *= 2;
            const int a_row = mt * 256 + (int)rank * 128;
            const int b_row0 = GLU ? (rank == 0 ? nt * 128 : p.n_out + nt * 128) : nt * 256 + (int)rank * 128;
            const int b_row1 = b_row0 + (GLU ? 128 : 256);
            const uint32_t bytes = hsel < 0 ? 2 * kwStageBytes : 2 * (k2ABytes + k2BBytes);   // both CTAs' boxes
            for (int kb = 0; kb < p.k_blocks; ++kb) {
                mbar_wait_bounded(&empty_bar[stage], phase ^ 1);
                if (lane == 0) {
                    unsigned char* sa = smem + (size_t)stage * kwStageBytes;
                    const uint32_t leader_full = mapa_shared(smem_u32(&full_bar[stage]), 0);
                    if (rank == 0) mbar_expect_tx(&full_bar[stage], bytes);
                    tma_load_2d_pair(sa, &map_a, kb * k2BK, a_row, leader_full);
                    if (hsel != 1) tma_load_2d_pair(sa + k2ABytes, &map_b, kb * k2BK, b_row0, leader_full);
                    if (hsel != 0) tma_load_2d_pair(sa + k2ABytes + k2BBytes, &map_b, kb * k2BK, b_row1, leader_full);
                }
                __syncwarp();
                if (++stage == kwStages) { stage = 0; phase ^= 1; }
            }
        }
    } else if (warp == 1) {
        // ================= MMA issuer (leader CTA only) =================
        if (rank == 0) {
            constexpr uint32_t idesc = make_idesc_f16(256, 256);
            int stage = 0;
            uint32_t phase = 0, acc_phase = 0;
            for (int v = cluster_id; v < items; v += n_clusters) {
                int u, hsel;
                item(v, u, hsel);
                for (int kb = 0; kb < p.k_blocks; ++kb) {
                    mbar_wait_bounded(&full_bar[stage], phase);
                    const uint32_t sa = smem_u32(smem + (size_t)stage * kwStageBytes);
                    const uint64_t da = make_kmajor_sw128_desc(sa);
#pragma unroll
                    for (int h = 0; h < 2; ++h) {
                        // the previous unit's accumulator h must have been drained before its first MMA overwrites it
                        if (kb == 0) mbar_wait_bounded(&acc_empty[h], acc_phase ^ 1);
                        tc_fence_after();
                        if (lane == 0 && (hsel < 0 || hsel == h)) {
                            const uint64_t db = make_kmajor_sw128_desc(sa + k2ABytes + (uint32_t)h * k2BBytes);
#pragma unroll
                            for (int k = 0; k < k2BK / kUmmaK; ++k)
                                umma_f16_pair(tmem_base + (uint32_t)(h * 256), da + (uint64_t)(2 * k), db + (uint64_t)(2 * k), idesc,
                                              (kb > 0 || k > 0) ? 1u : 0u);
                        }
                        __syncwarp();
                    }
                    if (lane == 0) {
                        umma_commit_pair(&empty_bar[stage], 3);                      // both producers may refill
                        if (kb == p.k_blocks - 1) {                                  // both epilogues may read
                            umma_commit_pair(&acc_full[0], 3);
                            umma_commit_pair(&acc_full[1], 3);
                        }
                    }
                    __syncwarp();
                    if (++stage == kwStages) { stage = 0; phase ^= 1; }
                }
                acc_phase ^= 1;
            }
        }
    } else {
        // ================= epilogue warps (both CTAs): warps 2-5 accumulator 0, warps 6-9 accumulator 1 =================
        const int quarter = warp & 3;
        const int h = warp >= 6 ? 1 : 0;
        uint32_t acc_phase = 0;
        for (int v = cluster_id; v < items; v += n_clusters) {
            int u, hsel, mt, nt;
            item(v, u, hsel);
            tile_of_unit(p, n_tiles, u, mt, nt);
            nt = 2 * nt + h;
            mbar_wait_bounded(&acc_full[h], acc_phase);
            tc_fence_after();
            const uint32_t taddr = tmem_base + ((uint32_t)(quarter * 32) << 16) + (uint32_t)(h * 256);
            if (hsel < 0 || hsel == h)   // (a single-tile item: the other accumulator's warps only keep the barriers in step)
                store_tile_warp<GLU>(p, taddr, mt * 256 + (int)rank * 128 + quarter * 32, nt,
                                     smem_u32(smem + kwOffStg + (uint32_t)(warp - 2) * kStgBytes), lane);
            tc_fence_before();
            __syncwarp();
            if (lane == 0) mbar_arrive_cluster(mapa_shared(smem_u32(&acc_empty[h]), 0));
            acc_phase ^= 1;
        }
    }
    tc_fence_before();
    __syncthreads();
    cluster_sync_all();
    if (warp == 1) {
        tc_fence_after();
        tmem_dealloc_pair(tmem_base, 512);
    }
}

static void fill_cluster_cfg(cudaLaunchConfig_t& cfg, cudaLaunchAttribute* attr, int grid, int cluster, cudaStream_t s, bool pdl,
                             int threads = k2Threads, uint32_t smem_bytes = k2Smem) {
    cfg = cudaLaunchConfig_t{};
    cfg.gridDim = dim3(grid);
    cfg.blockDim = dim3(threads);
    cfg.dynamicSmemBytes = smem_bytes;
    cfg.stream = s;
    int na = 0;
    attr[na].id = cudaLaunchAttributeClusterDimension;
    attr[na].val.clusterDim.x = cluster;
    attr[na].val.clusterDim.y = 1;
    attr[na].val.clusterDim.z = 1;
    ++na;
    if (pdl) {
        attr[na].id = cudaLaunchAttributeProgrammaticStreamSerialization;
        attr[na].val.programmaticStreamSerializationAllowed = 1;
        ++na;
    }
    cfg.attrs = attr;
    cfg.numAttrs = na;
}

// clusters of four CTAs that can be resident at once (a cluster lives inside one GPC: fewer than sm_count / 4 where a GPC's
// SM count is not a multiple of 4), asked from the runtime once per device and instantiation; 0 = not launchable
template <bool GLU>
static int quad_clusters() {
    static int cached[64];
    static bool have[64];
    const int dev = current_device();
    if (dev < 0 || dev >= 64) return 0;
    if (!have[dev]) {
        auto kern = gemm2_tcgen05_kernel<GLU, true>;
        static unsigned long long attr_done = 0;
        int n = 0;
        if (ensure_max_dynamic_smem(reinterpret_cast<const void*>(kern), (int)k2Smem, attr_done) == cudaSuccess) {
            prefer_max_smem_carveout(reinterpret_cast<const void*>(kern));
            cudaLaunchConfig_t cfg;
            cudaLaunchAttribute attr[2];
            fill_cluster_cfg(cfg, attr, 4 * (sm_count() / 4), 4, nullptr, false);
            if (cudaOccupancyMaxActiveClusters(&n, kern, &cfg) != cudaSuccess) { n = 0; cudaGetLastError(); }
        }
        if (getenv("B200_GEMM2_VERBOSE")) fprintf(stderr, "[b200] gemm2 quad: %d clusters of 4 CTAs resident (of %d SMs)\n", n, sm_count());
        cached[dev] = n;
        have[dev] = true;
    }
    return cached[dev];
}

template <bool GLU>
static int gemm2_launch_quad(const CUtensorMap& ma, const CUtensorMap& mb, const Gemm2Params& p, int clusters, cudaStream_t s) {
    auto kern = gemm2_tcgen05_kernel<GLU, true>;
    cudaLaunchConfig_t cfg;
    cudaLaunchAttribute attr[2];
    fill_cluster_cfg(cfg, attr, 4 * clusters, 4, s, pdl_enabled());
    cudaError_t e = cudaLaunchKernelEx(&cfg, kern, ma, mb, p);
    count_launch();
    return (int)e;
}

template <bool GLU>
static int gemm2_launch_wide(const CUtensorMap& ma, const CUtensorMap& mb, const Gemm2Params& p, int clusters, cudaStream_t s) {
    auto kern = gemm2w_tcgen05_kernel<GLU>;
    static unsigned long long attr_done = 0;   // per instantiation, one bit per device
    {
        cudaError_t e = ensure_max_dynamic_smem(reinterpret_cast<const void*>(kern), (int)kwSmem, attr_done);
        if (e != cudaSuccess) return (int)e;
    }
    prefer_max_smem_carveout(reinterpret_cast<const void*>(kern));
    cudaLaunchConfig_t cfg;
    cudaLaunchAttribute attr[2];
    fill_cluster_cfg(cfg, attr, 2 * clusters, 2, s, pdl_enabled(), kwThreads, kwSmem);
    cudaError_t e = cudaLaunchKernelEx(&cfg, kern, ma, mb, p);
    count_launch();
    return (int)e;
}

template <bool GLU>
static int gemm2_launch(const CUtensorMap& ma, const CUtensorMap& mb, const Gemm2Params& p, int grid, cudaStream_t s) {
    auto kern = gemm2_tcgen05_kernel<GLU, false>;
    static unsigned long long attr_done = 0;   // per instantiation, one bit per device
    {
        cudaError_t e = ensure_max_dynamic_smem(reinterpret_cast<const void*>(kern), (int)k2Smem, attr_done);
        if (e != cudaSuccess) return (int)e;
    }
    prefer_max_smem_carveout(reinterpret_cast<const void*>(kern));
    cudaLaunchConfig_t cfg{};
    cfg.gridDim = dim3(grid);
    cfg.blockDim = dim3(k2Threads);
    cfg.dynamicSmemBytes = k2Smem;
    cfg.stream = s;
    cudaLaunchAttribute attr[2];
    int na = 0;
    attr[na].id = cudaLaunchAttributeClusterDimension;
    attr[na].val.clusterDim.x = 2;
    attr[na].val.clusterDim.y = 1;
    attr[na].val.clusterDim.z = 1;
    ++na;
    if (pdl_enabled()) {
        attr[na].id = cudaLaunchAttributeProgrammaticStreamSerialization;
        attr[na].val.programmaticStreamSerializationAllowed = 1;
        ++na;
    }
    cfg.attrs = attr;
    cfg.numAttrs = na;
    cudaError_t e = cudaLaunchKernelEx(&cfg, kern, ma, mb, p);
    count_launch();
    return (int)e;
}

// y[M, N] = x[M, K] W[N, K]^T (+ bias) or, with glu, out[M, N] = silu(x Wg^T) * (x Wu^T) with W = [2N, K].
// Returns B200_EUNSUPPORTED for shapes the pair kernel does not cover (the one-CTA kernel takes them).
int gemm2_tcgen05_f16(const void* x, const void* W, const void* bias, void* y, int M, int N, int K, bool glu, cudaStream_t s,
                      const void* residual) {
    static int mode = -1;   // B200_GEMM_PAIR=0 keeps the one-CTA kernel
    if (mode < 0) { const char* e = getenv("B200_GEMM_PAIR"); mode = (e && e[0] == '0') ? 0 : 1; }
    if (!mode) return B200_EUNSUPPORTED;
    // below 512 token rows only the one case the caller has measured (gemm_tcgen05.cu: one row tile x >= 60 column tiles)
    const bool one_row_tile = !glu && M > 160 && M <= 256 && (N + 255) / 256 >= 60;
    if ((M < 512 && !one_row_tile) || N < 256 || K < 64 || (K & 7)) return B200_EUNSUPPORTED;
    if (glu && ((N & 127) || bias || residual)) return B200_EUNSUPPORTED;
    Gemm2Params p{};
    p.C = reinterpret_cast<__half*>(y);
    p.bias = reinterpret_cast<const __half*>(bias);
    p.residual = reinterpret_cast<const __half*>(residual);
    p.tokens = M; p.n_out = N; p.ldc = N;
    p.k_blocks = (K + k2BK - 1) / k2BK;
    p.m_tiles = (M + 255) / 256;
    p.n_tiles = glu ? N / 128 : (N + 255) / 256;
    {   // bands of 16 row tiles (measured at 8192 x 12288 x 4096, power-capped: 1108 -> 1191 TFLOP/s sustained; 8: 1150, 32 = the
        // old column-major order). B200_GEMM2_GROUP_M overrides (0 = column-major over all row tiles).
        static int group_env = -2;
        if (group_env == -2) { const char* g = getenv("B200_GEMM2_GROUP_M"); group_env = g ? atoi(g) : -1; }
        p.group_m = group_env >= 0 ? group_env : 16;
    }
    CUtensorMap ma, mb;
    // Opt-in (B200_GEMM2_QUAD=1, read per call so that tests can switch it): four-CTA clusters with the A tile multicast
    // between two MMA pairs. Needs an even number of column tiles and >= 2 units per cluster. Measured on the B200 at
    // 8192 x 12288 x 4096 (tools/gemm_power.py, 3 s back to back, board at its 1 kW cap): L2 -> SM bytes 6.44 -> 4.83 GB and
    // the SM clock rises 1335 -> 1455 MHz (cuBLAS' clock), but only 33 clusters of 4 are resident (132 of 148 SMs: a cluster
    // lives inside one GPC) and the throughput per clock drops 891 -> 782 TFLOP/s/GHz: 1137 vs 1190 TFLOP/s sustained, 1262
    // vs 1305 burst. The pair kernel stays the default.
    const char* quad_s = getenv("B200_GEMM2_QUAD");
    const int quad_env = (quad_s && quad_s[0] == '1') ? 1 : 0;
    // Opt-in (B200_GEMM2_WIDE=1, read per call so that tests can compare): 256 x 512 pair tiles (gemm2w_tcgen05_kernel); needs
    // an even number of column tiles and at least one unit per cluster. Measured at 8192 x 12288 x 4096 (tools/gemm_power.py,
    // three interleaved 2 s rounds per variant, board at its 1 kW cap): 1194 TFLOP/s at 1417 MHz against the 256 x 256
    // tile's 1210 at 1350 MHz (cuBLAS 1268 at 1430): the bytes saved do raise the clock, the un-hidden epilogue costs as
    // much per clock. An L2 prefetch 4 / 8 / 16 K steps ahead of the loads (cp.async.bulk.prefetch.tensor) lost 2-3 % on both
    // kernels (more L2 traffic = more energy) and was removed again.
    const char* wide_s = getenv("B200_GEMM2_WIDE");
    const int wide_env = (wide_s && wide_s[0] == '1') ? 1 : 0;
    if (wide_env && !quad_env && M >= 512 && (p.n_tiles & 1) == 0) {
        int clusters = sm_count() / 2;
        const int units_w = p.m_tiles * (p.n_tiles / 2);
        if (units_w >= clusters) {
            // the last wave: if at most half of the clusters would get a unit, run those units as single column tiles
            const int rem = units_w % clusters;
            p.wide_full = (rem > 0 && 2 * rem <= clusters) ? units_w - rem : units_w;
            if (!make_tma_map_2d(&ma, x, M, K, 128) || !make_tma_map_2d(&mb, W, glu ? 2ll * N : (long long)N, K, 128)) return B200_EUNSUPPORTED;
            return glu ? gemm2_launch_wide<true>(ma, mb, p, clusters, s) : gemm2_launch_wide<false>(ma, mb, p, clusters, s);
        }
    }
    if (quad_env && M >= 512 && (p.n_tiles & 1) == 0) {
        int clusters = glu ? quad_clusters<true>() : quad_clusters<false>();
        const int units4 = p.m_tiles * (p.n_tiles / 2);
        if (clusters > units4) clusters = units4;
        if (clusters >= 8 && units4 >= 2 * clusters) {
            if (!make_tma_map_2d(&ma, x, M, K, 64) || !make_tma_map_2d(&mb, W, glu ? 2ll * N : (long long)N, K, 128)) return B200_EUNSUPPORTED;
            return glu ? gemm2_launch_quad<true>(ma, mb, p, clusters, s) : gemm2_launch_quad<false>(ma, mb, p, clusters, s);
        }
    }
    if (!make_tma_map_2d(&ma, x, M, K, 128) || !make_tma_map_2d(&mb, W, glu ? 2ll * N : (long long)N, K, 128)) return B200_EUNSUPPORTED;
    const int units = p.m_tiles * p.n_tiles;
    int clusters = sm_count() / 2;
    if (clusters > units) clusters = units;
    return glu ? gemm2_launch<true>(ma, mb, p, 2 * clusters, s) : gemm2_launch<false>(ma, mb, p, 2 * clusters, s);
}

}  // namespace b200
