// tcgen05 / TMEM GEMM for fp16 linears with more than B200_GEMV_MAX_TOKENS rows:
//   prefill QKV / o / gate_up / down / LM-head (launchLinearGemm, src/kernels/linear.cu:39-104 ->
//   cublasWrapper::Gemm, cublas_utils.cc:32-71, call sites context_attention.cpp:99,166, ffn.cpp:72,88)
//   and the batched decode step (8 < B <= 256 sequences), where the same kernel runs "swapped".
//
//   D[128, BN] (fp32, TMEM) += A_tile[128, 64] * B_tile[BN, 64]^T      (both operands K-major fp16)
//
//   normal  : A = x [Nt, K] (tokens on the MMA M axis), B = W [N, K]    -> y[tok, n]   tensor-bound
//   swapped : A = W [N, K] (weight rows on the M axis), B = x [Nt, K]   -> y[tok, n]   HBM-bound:
//             every weight byte is read once, Nt <= 256 tokens ride along as the MMA N extent;
//             optional split-K over gridDim-level units so that N = 4096 projections still fill
//             148 SMs (fp32 partials in a workspace, reduced by a small epilogue kernel).
//
// Structure (one persistent CTA per SM, 192 threads):
//   warp 0    TMA producer: cp.async.bulk.tensor.2d (UTMALDG) of 128B-swizzled [rows x 64] boxes
//             into a STAGES-deep smem ring, completion on mbarriers (expect_tx);
//   warp 1    allocates TMEM, then one elected lane issues tcgen05.mma.cta_group::1.kind::f16
//             (UTCHMMA), 4 x (K=16) per stage; tcgen05.commit releases the smem stage and, at the end
//             of a tile, publishes the accumulator;
//   warps 2-5 epilogue: tcgen05.ld (LDTM) 32 lanes x 32 columns per instruction, convert, store;
//             two accumulator stages in TMEM so the epilogue of tile i overlaps the MMAs of i+1.
#include <mutex>
#include <unordered_map>

#include "common.cuh"
#include "launch_count.h"
#include "tc_common.cuh"

namespace b200 {

constexpr int kBM = 128;           // MMA M (rows of A per tile)
constexpr int kBK = 64;            // K per stage: 64 fp16 = one 128-byte swizzle row
constexpr int kGemmThreads = 192;  // producer, MMA, 4 epilogue warps

struct GemmParams {
    void* C;            // fp16 [tokens, ldc]
    float* partial;     // split-K workspace [splits][tokens][ldc] (swapped mode), or nullptr
    const void* bias;   // fp16 [N_out] or nullptr
    int rows_a;         // valid rows of the A operand (tokens in normal mode, N_out in swapped mode)
    int rows_b;         // valid rows of the B operand
    int K;
    int ldc;            // = N_out
    int tokens;
    int m_tiles, n_tiles, splits, k_blocks;
};

template <int BN>
struct GemmSmem {
    static constexpr int kStageBytes = (kBM + BN) * kBK * 2;
    static constexpr int kStages = (BN >= 256) ? 4 : (BN >= 128) ? 6 : (BN >= 64) ? 8 : 10;
    static constexpr int kBarrierBytes = (2 * kStages + 4) * 8 + 16;
    static constexpr int kTotal = kStages * kStageBytes + kBarrierBytes + 1024;  // + alignment slack
    static constexpr int kTmemCols = (2 * BN < 32) ? 32 : 2 * BN;                // two accumulator stages
};

// GLU (normal mode, BN = 256): W = [2*inter, K] with gate rows first (act_kernel.cu:28-29). The B tile of
// feature tile nt is 128 gate rows (nt*128) followed by the matching 128 up rows (inter + nt*128), so the
// accumulator holds gate in columns [0,128) and up in [128,256) of the same 128 features, and the epilogue
// writes silu(gate)*up (launchAct, act_kernel.cu:57) straight to out [tokens, inter]: the [tokens, 2*inter]
// intermediate of ffn.cpp:72-80 is never stored.
template <int BN, bool SWAP, bool GLU = false>
__global__ void __launch_bounds__(kGemmThreads, 1)
gemm_tcgen05_kernel(const __grid_constant__ CUtensorMap map_a, const __grid_constant__ CUtensorMap map_b,
                    const GemmParams p) {
    using S = GemmSmem<BN>;
    constexpr int STAGES = S::kStages;
    constexpr uint32_t A_BYTES = kBM * kBK * 2, B_BYTES = BN * kBK * 2;
    extern __shared__ unsigned char smem_raw[];
    // 128B-swizzled tiles must start on a 1024-byte boundary
    unsigned char* tiles = reinterpret_cast<unsigned char*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~(uintptr_t)1023);
    uint64_t* full_bar = reinterpret_cast<uint64_t*>(tiles + (size_t)STAGES * S::kStageBytes);
    uint64_t* empty_bar = full_bar + STAGES;
    uint64_t* acc_full = empty_bar + STAGES;   // [2]
    uint64_t* acc_empty = acc_full + 2;        // [2]
    uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(acc_empty + 2);

    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    if (threadIdx.x == 0) {
        for (int s = 0; s < STAGES; ++s) {
            mbar_init(&full_bar[s], 1);
            mbar_init(&empty_bar[s], 1);
        }
        for (int a = 0; a < 2; ++a) {
            mbar_init(&acc_full[a], 1);
            mbar_init(&acc_empty[a], 4);
        }
        mbar_fence_init();
        tma_prefetch_desc(&map_a);
        tma_prefetch_desc(&map_b);
    }
    if (warp == 1) tmem_alloc(tmem_slot, S::kTmemCols);
    tc_fence_before();
    __syncthreads();
    tc_fence_after();
    const uint32_t tmem_base = *tmem_slot;
    pdl_launch_dependents();
    pdl_wait();   // x (and, for PDL chains, the previous kernel's output) must be complete

    const int units = p.m_tiles * p.n_tiles * p.splits;

    if (warp == 0) {
        // ================= TMA producer =================
        int stage = 0;
        uint32_t phase = 0;
        for (int u = blockIdx.x; u < units; u += gridDim.x) {
            const int mt = u % p.m_tiles, rest = u / p.m_tiles;
            const int nt = rest % p.n_tiles, sp = rest / p.n_tiles;
            const int kb0 = (int)((long long)p.k_blocks * sp / p.splits);
            const int kb1 = (int)((long long)p.k_blocks * (sp + 1) / p.splits);
            for (int kb = kb0; kb < kb1; ++kb) {
                mbar_wait(&empty_bar[stage], phase ^ 1);
                if (lane == 0) {
                    unsigned char* sa = tiles + (size_t)stage * S::kStageBytes;
                    mbar_expect_tx(&full_bar[stage], A_BYTES + B_BYTES);
                    tma_load_2d(sa, &map_a, kb * kBK, mt * kBM, &full_bar[stage]);
                    if (GLU) {   // map_b boxes are 128 rows: gate rows, then the up rows of the same features
                        tma_load_2d(sa + A_BYTES, &map_b, kb * kBK, nt * 128, &full_bar[stage]);
                        tma_load_2d(sa + A_BYTES + B_BYTES / 2, &map_b, kb * kBK, p.rows_b + nt * 128, &full_bar[stage]);
                    } else {
                        tma_load_2d(sa + A_BYTES, &map_b, kb * kBK, nt * BN, &full_bar[stage]);
                    }
                }
                __syncwarp();
                if (++stage == STAGES) { stage = 0; phase ^= 1; }
            }
        }
    } else if (warp == 1) {
        // ================= MMA issuer =================
        constexpr uint32_t idesc = make_idesc_f16(kBM, BN);
        int stage = 0;
        uint32_t phase = 0;
        int acc = 0;
        uint32_t acc_phase = 0;
        for (int u = blockIdx.x; u < units; u += gridDim.x) {
            const int sp = (u / p.m_tiles) / p.n_tiles;
            const int kb0 = (int)((long long)p.k_blocks * sp / p.splits);
            const int kb1 = (int)((long long)p.k_blocks * (sp + 1) / p.splits);
            mbar_wait(&acc_empty[acc], acc_phase ^ 1);   // epilogue has drained this accumulator
            tc_fence_after();
            const uint32_t d_tmem = tmem_base + (uint32_t)(acc * BN);
            for (int kb = kb0; kb < kb1; ++kb) {
                mbar_wait(&full_bar[stage], phase);
                tc_fence_after();
                if (lane == 0) {
                    const uint32_t sa = smem_u32(tiles + (size_t)stage * S::kStageBytes);
                    const uint64_t da = make_kmajor_sw128_desc(sa);
                    const uint64_t db = make_kmajor_sw128_desc(sa + A_BYTES);
#pragma unroll
                    for (int k = 0; k < kBK / kUmmaK; ++k)
                        umma_f16(d_tmem, da + (uint64_t)(2 * k), db + (uint64_t)(2 * k), idesc,
                                 (kb > kb0 || k > 0) ? 1u : 0u);
                    umma_commit(&empty_bar[stage]);                 // smem stage reusable when the MMAs retire
                    if (kb == kb1 - 1) umma_commit(&acc_full[acc]);  // accumulator complete
                }
                __syncwarp();
                if (++stage == STAGES) { stage = 0; phase ^= 1; }
            }
            if (kb1 == kb0 && lane == 0) umma_commit(&acc_full[acc]);  // empty K range (never with K > 0)
            acc ^= 1;
            if (acc == 0) acc_phase ^= 1;
        }
    } else {
        // ================= epilogue warps =================
        const int quarter = warp & 3;   // TMEM lanes [32*quarter, 32*quarter + 32) belong to this warp
        int acc = 0;
        uint32_t acc_phase = 0;
        __half* C = reinterpret_cast<__half*>(p.C);
        const __half* bias = reinterpret_cast<const __half*>(p.bias);
        for (int u = blockIdx.x; u < units; u += gridDim.x) {
            const int mt = u % p.m_tiles, rest = u / p.m_tiles;
            const int nt = rest % p.n_tiles, sp = rest / p.n_tiles;
            mbar_wait(&acc_full[acc], acc_phase);
            tc_fence_after();
            const int row = mt * kBM + quarter * 32 + lane;   // row of the A operand / of D
            const uint32_t taddr = tmem_base + ((uint32_t)(quarter * 32) << 16) + (uint32_t)(acc * BN);
            if (GLU) {
                // rows_b = inter; gate in columns [0,128), up in [128,256); same rounding points as the unfused
                // sequence: GEMM output rounded to fp16, silu(gate) rounded to fp16, product rounded to fp16
#pragma unroll 1
                for (int c0 = 0; c0 < 128; c0 += 32) {
                    uint32_t g[32], u[32];
                    tmem_ld_32x32(taddr + (uint32_t)c0, g);
                    tmem_ld_32x32(taddr + (uint32_t)(128 + c0), u);
                    const int col0 = nt * 128 + c0;
                    if (row < p.tokens && col0 < p.rows_b) {
                        __half* dst = C + (size_t)row * p.ldc + col0;
#pragma unroll
                        for (int q = 0; q < 4; ++q) {
                            float f[8];
#pragma unroll
                            for (int e = 0; e < 8; ++e) {
                                const float gv = __half2float(__float2half_rn(__uint_as_float(g[q * 8 + e])));
                                const float uv = __half2float(__float2half_rn(__uint_as_float(u[q * 8 + e])));
                                const float sv = __half2float(__float2half_rn(gv / (1.f + expf(-gv))));
                                f[e] = sv * uv;
                            }
                            if (col0 + q * 8 + 8 <= p.rows_b) st16(dst + q * 8, Unpack<__half>::put(f));
                        }
                    }
                }
            } else
#pragma unroll 1
            for (int c0 = 0; c0 < BN; c0 += 32) {
                uint32_t v[32];
                tmem_ld_32x32(taddr + (uint32_t)c0, v);
                const int col0 = nt * BN + c0;   // first row of the B operand covered by this chunk
                if (!SWAP) {
                    // row = token, col = output feature: 32 consecutive features per thread
                    if (row < p.tokens && col0 < p.rows_b) {
                        __half* dst = C + (size_t)row * p.ldc + col0;
                        if (col0 + 32 <= p.rows_b && (p.ldc & 7) == 0) {
#pragma unroll
                            for (int q = 0; q < 4; ++q) {
                                float f[8];
#pragma unroll
                                for (int e = 0; e < 8; ++e) {
                                    f[e] = __uint_as_float(v[q * 8 + e]);
                                    if (bias) f[e] += __half2float(bias[col0 + q * 8 + e]);
                                }
                                st16(dst + q * 8, Unpack<__half>::put(f));
                            }
                        } else {
                            // unrolled with a predicate: a run-time index would move v[] to local memory (no L1 behind
                            // it with the 227 KB carve-out) for EVERY path of the epilogue
#pragma unroll
                            for (int j = 0; j < 32; ++j) {
                                if (col0 + j >= p.rows_b) continue;
                                float f = __uint_as_float(v[j]);
                                if (bias) f += __half2float(bias[col0 + j]);
                                dst[j] = __float2half_rn(f);
                            }
                        }
                    }
                } else {
                    // row = output feature, col = token: lanes write consecutive features of one token
                    if (row < p.rows_a) {
                        const float bv = (bias && !p.partial) ? __half2float(bias[row]) : 0.f;
#pragma unroll
                        for (int j = 0; j < 32; ++j) {
                            const int tok = col0 + j;
                            if (tok < p.tokens) {
                                const float f = __uint_as_float(v[j]);
                                if (p.partial)
                                    p.partial[((size_t)sp * p.tokens + tok) * p.ldc + row] = f;
                                else
                                    C[(size_t)tok * p.ldc + row] = __float2half_rn(f + bv);
                            }
                        }
                    }
                }
            }
            tc_fence_before();
            __syncwarp();
            if (lane == 0) mbar_arrive(&acc_empty[acc]);
            acc ^= 1;
            if (acc == 0) acc_phase ^= 1;
        }
    }
    tc_fence_before();
    __syncthreads();
    if (warp == 1) {
        tc_fence_after();
        tmem_dealloc(tmem_base, S::kTmemCols);
    }
}

// split-K reduction: y[tok, n] = half(sum_s partial[s, tok, n] + bias[n])
__global__ void __launch_bounds__(256)
splitk_reduce_kernel(const float* __restrict__ partial, const __half* __restrict__ bias, __half* __restrict__ y,
                     int splits, size_t elems, int ldc) {
    pdl_wait();
    const size_t i = ((size_t)blockIdx.x * blockDim.x + threadIdx.x) * 4;
    if (i >= elems) return;
    float4 acc = *reinterpret_cast<const float4*>(partial + i);
    for (int s = 1; s < splits; ++s) {
        const float4 t = *reinterpret_cast<const float4*>(partial + (size_t)s * elems + i);
        acc.x += t.x; acc.y += t.y; acc.z += t.z; acc.w += t.w;
    }
    if (bias) {
        const int n = (int)(i % (size_t)ldc);
        acc.x += __half2float(bias[n]); acc.y += __half2float(bias[n + 1]);
        acc.z += __half2float(bias[n + 2]); acc.w += __half2float(bias[n + 3]);
    }
    pdl_launch_dependents();
    __half2* out = reinterpret_cast<__half2*>(y + i);
    out[0] = __floats2half2_rn(acc.x, acc.y);
    out[1] = __floats2half2_rn(acc.z, acc.w);
}

// ------------------------------------------------------------------------------------------
// host side
// ------------------------------------------------------------------------------------------
typedef CUresult (*EncodeTiledFn)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*,
                                  const cuuint64_t*, const cuuint32_t*, const cuuint32_t*, CUtensorMapInterleave,
                                  CUtensorMapSwizzle, CUtensorMapL2promotion, CUtensorMapFloatOOBfill);

// the driver entry point is resolved through the runtime, so the library has no link-time
// dependency on libcuda (it must load on the GPU-less build box)
static EncodeTiledFn encode_tiled_fn() {
    static EncodeTiledFn fn = nullptr;
    static std::once_flag once;
    std::call_once(once, [] {
        void* p = nullptr;
        cudaDriverEntryPointQueryResult q;
        if (cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &p, cudaEnableDefault, &q) == cudaSuccess &&
            q == cudaDriverEntryPointSuccess)
            fn = reinterpret_cast<EncodeTiledFn>(p);
    });
    return fn;
}

// 2-D map over a row-major fp16 matrix [rows, K]: box = [box_rows x 64], 128B swizzle, zero fill
// outside the tensor (ragged M / N / K tails need no special code in the kernel)
bool make_tma_map_2d(CUtensorMap* out, const void* ptr, long long rows, int K, int box_rows) {
    struct Key {
        const void* p; long long rows; int K, box;
        bool operator==(const Key& o) const { return p == o.p && rows == o.rows && K == o.K && box == o.box; }
    };
    struct Hash {
        size_t operator()(const Key& k) const {
            return std::hash<const void*>()(k.p) ^ (size_t)k.rows * 1000003u ^ (size_t)k.K * 7919u ^ (size_t)k.box;
        }
    };
    static std::unordered_map<Key, CUtensorMap, Hash> cache;
    static std::mutex mu;
    std::lock_guard<std::mutex> g(mu);
    const Key key{ptr, rows, K, box_rows};
    auto it = cache.find(key);
    if (it != cache.end()) { *out = it->second; return true; }
    EncodeTiledFn enc = encode_tiled_fn();
    if (!enc) return false;
    const cuuint64_t dims[2] = {(cuuint64_t)K, (cuuint64_t)rows};
    const cuuint64_t strides[1] = {(cuuint64_t)K * 2};
    const cuuint32_t box[2] = {(cuuint32_t)kBK, (cuuint32_t)box_rows};
    const cuuint32_t estr[2] = {1, 1};
    CUtensorMap m;
    const CUresult r = enc(&m, CU_TENSOR_MAP_DATA_TYPE_FLOAT16, 2, const_cast<void*>(ptr), dims, strides, box, estr,
                           CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_128B,
                           CU_TENSOR_MAP_L2_PROMOTION_L2_256B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
    if (r != CUDA_SUCCESS) return false;
    if (cache.size() > 4096) cache.clear();
    cache.emplace(key, m);
    *out = m;
    return true;
}

template <int BN, bool SWAP, bool GLU = false>
static int gemm_launch(const CUtensorMap& ma, const CUtensorMap& mb, const GemmParams& p, int grid, cudaStream_t s) {
    auto kern = gemm_tcgen05_kernel<BN, SWAP, GLU>;
    static unsigned long long attr_done = 0;   // per instantiation, one bit per device
    {
        cudaError_t e = ensure_max_dynamic_smem(reinterpret_cast<const void*>(kern), GemmSmem<BN>::kTotal, attr_done);
        if (e != cudaSuccess) return (int)e;
    }
    cudaError_t e = launch_k(kern, dim3(grid), dim3(kGemmThreads), (size_t)GemmSmem<BN>::kTotal, s, true, ma, mb, p);
    count_launch();
    return (int)e;
}

// split-K factor for the swapped (weight-streaming) mode: minimise rounds x K-blocks per unit
static int pick_splits(int m_tiles, int k_blocks, int sms) {
    int best = 1;
    long long best_cost = -1;
    for (int s = 1; s <= 16 && k_blocks / s >= 4; ++s) {
        const long long rounds = ((long long)m_tiles * s + sms - 1) / sms;
        const long long per = (k_blocks + s - 1) / s;
        const long long cost = rounds * (per + 2) + (s > 1 ? s : 0);   // +2: tile prologue/epilogue; +s: partial traffic
        if (best_cost < 0 || cost < best_cost) { best_cost = cost; best = s; }
    }
    return best;
}

int gemm2_tcgen05_f16(const void* x, const void* W, const void* bias, void* y, int M, int N, int K, bool glu, cudaStream_t s,
                      const void* residual = nullptr);

// fewest token rows b200_linear sends to the tcgen05 weight-streaming GEMM instead of the GEMV (B200_GEMM_MIN_TOKENS
// overrides, for sweeps)
int gemm_min_tokens() {
    static int v = -1;
    if (v < 0) { const char* e = getenv("B200_GEMM_MIN_TOKENS"); v = e ? atoi(e) : 5; if (v < 1) v = 1; }
    return v;
}

size_t gemm_tcgen05_workspace_bytes(int M, int N, int K) {
    if (M < gemm_min_tokens() || M > 256) return 0;
    const int splits = pick_splits((N + kBM - 1) / kBM, (K + kBK - 1) / kBK, sm_count());
    return splits > 1 ? (size_t)splits * M * N * sizeof(float) : 0;
}

// y[M, N] = x[M, K] * W[N, K]^T (+ bias), fp16 in / fp32 accumulate / fp16 out
int gemm_tcgen05_f16(const void* x, const void* W, const void* bias, void* y, int M, int N, int K, void* workspace,
                     size_t workspace_bytes, cudaStream_t s) {
    auto al = [](const void* q) { return (reinterpret_cast<uintptr_t>(q) & 15u) == 0; };
    if (M < 1 || N < 8 || K < 8 || (K & 7) || !al(x) || !al(W)) return B200_EUNSUPPORTED;
    static int disabled = -1;
    if (disabled < 0) { const char* e = getenv("B200_NO_TCGEN05"); disabled = (e && e[0] == '1') ? 1 : 0; }
    if (disabled) return B200_EUNSUPPORTED;
    const int sms = sm_count();
    GemmParams p{};
    p.C = y; p.bias = bias; p.K = K; p.ldc = N; p.tokens = M;
    p.k_blocks = (K + kBK - 1) / kBK;
    CUtensorMap ma, mb;
    // 161..256 token rows against a wide weight matrix (>= 60 column tiles of 256: most of the 74 cluster slots busy): one row
    // tile of the two-CTA kernel beats the swapped kernel's 2 waves of 128 x 256 tiles (measured, us: 22016 x 4096 at 192 /
    // 256 rows: 60.9 / 61.4 vs 75.8 / 87.6; 12288 x 4096 at 256: 35.5 vs 44.0; narrow outputs stay swapped: 4096 x 11008 at
    // 256: 38.4 vs 65.5)
    if (M > 160 && M <= 256 && (N + 255) / 256 >= 60) {
        const int rc2 = gemm2_tcgen05_f16(x, W, bias, y, M, N, K, false, s);
        if (rc2 != B200_EUNSUPPORTED) return rc2;
    }
    if (M <= 256) {
        // swapped: weights stream through the M axis, tokens are the N extent
        const int BN = M <= 32 ? 32 : M <= 64 ? 64 : M <= 128 ? 128 : 256;
        p.rows_a = N; p.rows_b = M;
        p.m_tiles = (N + kBM - 1) / kBM; p.n_tiles = 1;
        p.splits = pick_splits(p.m_tiles, p.k_blocks, sms);
        if (p.splits > 1) {
            const size_t need = (size_t)p.splits * M * N * sizeof(float);
            if (!workspace || workspace_bytes < need || (N & 3)) p.splits = 1;
            else p.partial = reinterpret_cast<float*>(workspace);
        }
        if (!make_tma_map_2d(&ma, W, N, K, kBM) || !make_tma_map_2d(&mb, x, M, K, BN)) return B200_EUNSUPPORTED;
        const int units = p.m_tiles * p.splits;
        const int grid = units < sms ? units : sms;
        int rc;
        switch (BN) {
            case 32: rc = gemm_launch<32, true>(ma, mb, p, grid, s); break;
            case 64: rc = gemm_launch<64, true>(ma, mb, p, grid, s); break;
            case 128: rc = gemm_launch<128, true>(ma, mb, p, grid, s); break;
            default: rc = gemm_launch<256, true>(ma, mb, p, grid, s); break;
        }
        if (rc != 0 || p.splits == 1) return rc;
        const size_t elems = (size_t)M * N;
        cudaError_t e = launch_k(splitk_reduce_kernel, dim3((unsigned)((elems / 4 + 255) / 256)), dim3(256), 0, s, true,
                                 (const float*)p.partial, (const __half*)bias, (__half*)y, p.splits, elems, N);
        count_launch();
        return (int)e;
    }
    // normal: tokens on the M axis; large shapes go to the two-CTA kernel (gemm2_tcgen05.cu)
    {
        const int rc2 = gemm2_tcgen05_f16(x, W, bias, y, M, N, K, false, s);
        if (rc2 != B200_EUNSUPPORTED) return rc2;
    }
    const int BN = N >= 256 ? 256 : N >= 128 ? 128 : N >= 64 ? 64 : 32;
    p.rows_a = M; p.rows_b = N;
    p.m_tiles = (M + kBM - 1) / kBM; p.n_tiles = (N + BN - 1) / BN; p.splits = 1;
    if (!make_tma_map_2d(&ma, x, M, K, kBM) || !make_tma_map_2d(&mb, W, N, K, BN)) return B200_EUNSUPPORTED;
    const int units = p.m_tiles * p.n_tiles;
    const int grid = units < sms ? units : sms;
    switch (BN) {
        case 32: return gemm_launch<32, false>(ma, mb, p, grid, s);
        case 64: return gemm_launch<64, false>(ma, mb, p, grid, s);
        case 128: return gemm_launch<128, false>(ma, mb, p, grid, s);
        default: return gemm_launch<256, false>(ma, mb, p, grid, s);
    }
}

// out[M, inter] = silu(x W_gate^T) * (x W_up^T), W = [2*inter, K] (gate rows, then up rows): the gate_up GEMM of
// ffn.cpp:72 with launchAct (ffn.cpp:80) as its epilogue. Prefill shapes only (M > 256, inter % 128 == 0).
int gemm_tcgen05_f16_swiglu(const void* x, const void* W, void* out, int M, int inter, int K, cudaStream_t s) {
    auto al = [](const void* q) { return (reinterpret_cast<uintptr_t>(q) & 15u) == 0; };
    if (M <= 256 || inter < 128 || (inter & 127) || K < 8 || (K & 7) || !al(x) || !al(W) || !al(out)) return B200_EUNSUPPORTED;
    const char* e = getenv("B200_NO_TCGEN05");
    if (e && e[0] == '1') return B200_EUNSUPPORTED;
    {
        const int rc2 = gemm2_tcgen05_f16(x, W, nullptr, out, M, inter, K, true, s);
        if (rc2 != B200_EUNSUPPORTED) return rc2;
    }
    GemmParams p{};
    p.C = out; p.bias = nullptr; p.K = K; p.ldc = inter; p.tokens = M;
    p.k_blocks = (K + kBK - 1) / kBK;
    p.rows_a = M; p.rows_b = inter;
    p.m_tiles = (M + kBM - 1) / kBM; p.n_tiles = inter / 128; p.splits = 1;
    CUtensorMap ma, mb;
    if (!make_tma_map_2d(&ma, x, M, K, kBM) || !make_tma_map_2d(&mb, W, 2ll * inter, K, 128)) return B200_EUNSUPPORTED;
    const int units = p.m_tiles * p.n_tiles, sms = sm_count();
    return gemm_launch<256, false, true>(ma, mb, p, units < sms ? units : sms, s);
}

// y[M, N] = residual[M, N] + half(x W^T): launchLinearGemm followed by launchAddResidual (ffn.cpp:88 +
// context_decoder.cpp:134) in the epilogue of the two-CTA GEMM; B200_EUNSUPPORTED outside its shapes
int gemm_tcgen05_f16_residual(const void* x, const void* W, const void* residual, void* y, int M, int N, int K, cudaStream_t s) {
    auto al = [](const void* q) { return (reinterpret_cast<uintptr_t>(q) & 15u) == 0; };
    if (!al(x) || !al(W) || !al(y) || !al(residual) || (N & 7)) return B200_EUNSUPPORTED;
    const char* e = getenv("B200_NO_TCGEN05");
    if (e && e[0] == '1') return B200_EUNSUPPORTED;
    return gemm2_tcgen05_f16(x, W, nullptr, y, M, N, K, false, s, residual);
}

}  // namespace b200
