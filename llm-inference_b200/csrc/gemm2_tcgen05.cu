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
constexpr uint32_t k2Smem = k2OffSlot + 16;

struct Gemm2Params {
    __half* C;          // [tokens, ldc]
    const __half* bias; // [N] or nullptr (not with GLU)
    const __half* residual;   // [tokens, ldc] or nullptr: C = half(half(acc + bias) + residual), launchAddResidual folded in
    int tokens, n_out;  // valid rows of A; valid output features (GLU: inter)
    int ldc, k_blocks;
    int m_tiles, n_tiles;   // tiles of 256 rows x (256 features | GLU: 128 features)
    int group_m;            // tile order: bands of group_m row tiles swept over all column tiles (0: column-major over all rows)
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
                            for (int j = 0; j < 32 && col0 + j < p.n_out; ++j) {
                                float f = __uint_as_float(v[j]);
                                if (p.bias) f += __half2float(p.bias[col0 + j]);
                                if (p.residual) f = __half2float(__float2half_rn(f)) + __half2float(p.residual[(size_t)row * p.ldc + col0 + j]);
                                dst[j] = __float2half_rn(f);
                            }
                        }
                    }
                }
            }
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

static void fill_cluster_cfg(cudaLaunchConfig_t& cfg, cudaLaunchAttribute* attr, int grid, int cluster, cudaStream_t s, bool pdl) {
    cfg = cudaLaunchConfig_t{};
    cfg.gridDim = dim3(grid);
    cfg.blockDim = dim3(k2Threads);
    cfg.dynamicSmemBytes = k2Smem;
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
