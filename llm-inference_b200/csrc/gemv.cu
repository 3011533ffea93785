// Weight-streaming GEMV for the decode-time linears:  y[b, n] = sum_k x[b, k] * W[n, k].
// Replaces cuBLAS GemmEx with n = batch (src/kernels/linear.cu:39-104, cublas_utils.cc:32-71)
// at the call sites masked_self_attention.cpp:62,84 and ffn.cpp:72,88, and optionally fuses the
// neighbouring bandwidth kernels of the decode layer (rmsnorm_kernel.cu:63, act_kernel.cu:18,
// qkv_bias_and_RoPE.cu:323, the KV append of fused_decoder_self_attention.cu:170-172,274-276,
// and the residual adds of fused_addresidual_norm.cu:62 / add_residual.cu:9).
//
// Structure (HBM-bound; algorithmic bytes = N*K*sizeof(T) + small vectors):
//  * one persistent CTA per SM owns a contiguous, byte-balanced range of row-pair "units";
//  * a producer warp streams those rows through a ring of 32 KB stages with 1-D bulk async
//    copies (cp.async.bulk -> SASS UBLKCP, completion on an mbarrier); each stage holds
//    8 units x 2 rows x 2 KB K-segment;
//  * 8 consumer warps: warp w owns unit w of the stage, dots both rows against x held in
//    shared memory (one x packet read serves two weight packets), fp32 accumulation
//    (fma.rn.f32.f16 -> FHFMA for half), warp-shuffle reduce at the end of the row;
//    from 2 fp16 token rows on the consumers switch to mma.sync (kSegPitchMma below): measured at batch 2 / 3 / 4,
//    7B shapes, fused order, context 512: 3.50 / 4.45 / 4.71 -> 3.39 / 3.82 / 4.07 ms per step;
//  * under programmatic dependent launch the producer starts streaming WEIGHTS before
//    griddepcontrol.wait, so the HBM pipe stays busy across kernel boundaries; only the
//    x load waits for the predecessor.
#include "common.cuh"
#include "launch_count.h"
#include "p2p_internal.h"

namespace b200 {

constexpr int kGemvWarps = 8;                       // consumer warps == units per stage
constexpr int kGemvThreads = (kGemvWarps + 1) * 32; // + producer warp
#ifndef B200_GEMV_SEG_KB
#define B200_GEMV_SEG_KB 2
#endif
constexpr int kSegBytes = B200_GEMV_SEG_KB * 1024;  // K-segment per row per stage = the size of one bulk copy
constexpr int kStageBytes = kGemvWarps * 2 * kSegBytes;  // 32 KB at 2 KB segments
constexpr int kMaxStages = 4;
constexpr int kJ = kSegBytes / 16 / 32;  // 16B packets per lane per row-segment (=4)
// Tensor-core consumer (fp16, 2..8 token rows): one FHFMA per weight element PER TOKEN makes the scalar loop
// issue-bound from 2 rows on (2.8 TB/s at 4 rows). The MMA consumer keeps the same 2 KB contiguous row stream but
// reads the stage as a [16 rows x 1024 k] fp16 matrix: warp w takes k in [128 w, 128 w + 128) of ALL 16 rows with
// eight mma.sync.m16n8k16 (A = weights, B = the <= 8 token rows of x, fp32 accumulate), and the eight warps' partial
// [16 x 8] sums meet in shared memory once per row group. Row pitch 2 KB + 16 B and x pitch K + 8 elements keep every
// fragment load (one 32-bit word per lane) conflict-free.
constexpr int kSegPitchMma = kSegBytes + 16;
constexpr int kStageBytesMma = kGemvWarps * 2 * kSegPitchMma;
constexpr int kXPadMma = 8;   // elements

enum { kModePlain = 0, kModeResidual = 1, kModeSwiglu = 2, kModeRopeKV = 3 };

struct GemvParams {
    const void* x;
    const void* W;
    const void* bias;
    void* y;
    const void* gamma;
    float eps;
    int N;       // output features (for SWIGLU: inter size; W has 2N rows)
    int K;
    int units;   // row pairs to process
    int stages;  // ring depth actually allocated
    int nb;      // valid tokens (<= NB)
    unsigned flags;  // B200_GEMV_NORM / _NORM_GFIRST (prologue, run-time)
    // rope / kv
    void* k_cache;
    void* v_cache;
    int head_num, kv_head_num, head_size, max_seq_len, step, rotary_dim;
    float rotary_base;
    const int* step_dev;
    int step_stride;   // 1: step_dev[token] is each sequence's own step (ragged batch); 0: one shared *step_dev
    const float* rope_den;
    P2PWire wire;   // world > 0 (plain mode, fp16): outputs are pushed to the peers instead of written to y
    int token0;     // index of this launch's first token in the message (launches tiled over tokens)
};

// unit -> (row_a, row_b); row_b < 0 means "no second row"
template <int MODE>
__device__ __forceinline__ void unit_rows(const GemvParams& p, int u, int& ra, int& rb) {
    if (MODE == kModeSwiglu) {
        ra = u;
        rb = p.N + u;
    } else if (MODE == kModeRopeKV) {
        const int hd = p.head_size >> 1;
        const int head = u / hd, i = u - head * hd;
        ra = head * p.head_size + i;
        rb = ra + hd;
    } else {
        ra = 2 * u;
        rb = (2 * u + 1 < p.N) ? 2 * u + 1 : -1;
    }
}

__device__ __forceinline__ void consumer_bar() {
    asm volatile("bar.sync 1, %0;" ::"n"(kGemvWarps * 32) : "memory");
}

__device__ __forceinline__ void mma_f16_16816(float (&c)[4], uint32_t a0, uint32_t a1, uint32_t a2, uint32_t a3,
                                              uint32_t b0, uint32_t b1) {
    asm volatile("mma.sync.aligned.m16n8k16.row.col.f32.f16.f16.f32 {%0,%1,%2,%3}, {%4,%5,%6,%7}, {%8,%9}, {%0,%1,%2,%3};"
                 : "+f"(c[0]), "+f"(c[1]), "+f"(c[2]), "+f"(c[3])
                 : "r"(a0), "r"(a1), "r"(a2), "r"(a3), "r"(b0), "r"(b1));
}

template <typename T, int NB, int MODE, bool MMA = false>
__global__ void __launch_bounds__(kGemvThreads, 1) gemv_stream_kernel(const GemvParams p) {
    constexpr int E = DT<T>::kPer16B;
    constexpr int KSEG = kSegBytes / (int)sizeof(T);  // elements per segment
    constexpr int PITCH = MMA ? kSegPitchMma : kSegBytes;        // bytes between the rows of a stage
    constexpr int STAGE = MMA ? kStageBytesMma : kStageBytes;
    extern __shared__ __align__(128) unsigned char smem[];
    unsigned char* ring = smem;
    const int xp = MMA ? p.K + kXPadMma : p.K;                   // elements between the token rows of x
    T* xs = reinterpret_cast<T*>(smem + (size_t)p.stages * STAGE);
    uint64_t* full_bar = reinterpret_cast<uint64_t*>(
        smem + (size_t)p.stages * STAGE + (((size_t)NB * xp * sizeof(T) + 15) & ~(size_t)15));
    uint64_t* empty_bar = full_bar + kMaxStages;
    float* red = reinterpret_cast<float*>(empty_bar + kMaxStages);  // [NB][kGemvWarps] + [NB]
    float* red_mma = red + NB * kGemvWarps + NB;                    // MMA: [kGemvWarps][16][8]

    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const int nseg = (p.K + KSEG - 1) / KSEG;
    const int nstages = p.stages;
    // byte-balanced contiguous unit range of this CTA
    const int u0 = (int)(((long long)p.units * blockIdx.x) / gridDim.x);
    const int u1 = (int)(((long long)p.units * (blockIdx.x + 1)) / gridDim.x);
    const int ngroups = (u1 - u0 + kGemvWarps - 1) / kGemvWarps;

    if (threadIdx.x == 0) {
        for (int s = 0; s < nstages; ++s) {
            mbar_init(&full_bar[s], 1);
            mbar_init(&empty_bar[s], kGemvWarps);
        }
        mbar_fence_init();
    }
    __syncthreads();
    // Dependents only prefetch weights before their own wait, so they may start at once - except after the qkv GEMV
    // (RoPE/KV epilogue), whose dependent is the decode attention: launched this early its small CTAs all become resident
    // on the first ~50 SMs (13 fit next to one GEMV CTA) and later stream the KV cache from a third of the chip
    // (measured, batch 2 x 4096 keys: 5.06 ms per step against 4.42 in the reference launch order). There the trigger
    // comes when this CTA's rows are done.
    if (MODE != kModeRopeKV) pdl_launch_dependents();

    if (warp == kGemvWarps) {
        // ============ producer warp: lane l streams row (l&1) of unit (l>>1) of each group ============
        // 16 lanes issue their bulk copies in parallel; lane 0 also arms the barrier. The
        // expect_tx may land after some complete_tx of the same phase: the phase cannot finish
        // before the single arrival that expect_tx carries.
        const unsigned char* Wb = reinterpret_cast<const unsigned char*>(p.W);
        const size_t row_bytes = (size_t)p.K * sizeof(T);
        const int last_seg_bytes = (p.K - (nseg - 1) * KSEG) * (int)sizeof(T);
        int slot = 0;
        uint32_t phase = 0;  // parity to wait for on empty_bar once the ring has wrapped
        bool wrapped = false;
        for (int g = 0; g < ngroups; ++g) {
            const int ubase = u0 + g * kGemvWarps;
            const int nu = min(kGemvWarps, u1 - ubase);
            // this lane's source row for the whole group
            const unsigned char* src = nullptr;
            if (lane < 2 * nu) {
                int ra, rb;
                unit_rows<MODE>(p, ubase + (lane >> 1), ra, rb);
                const int r = (lane & 1) ? rb : ra;
                if (r >= 0) src = Wb + (size_t)r * row_bytes;
            }
            const uint32_t rows_here = __popc(__ballot_sync(0xffffffffu, src != nullptr));
            unsigned char* dst_lane = ring + (size_t)lane * PITCH;
            for (int seg = 0; seg < nseg; ++seg) {
                if (wrapped) mbar_wait(&empty_bar[slot], phase);
                const uint32_t seg_bytes = (seg == nseg - 1) ? (uint32_t)last_seg_bytes : (uint32_t)kSegBytes;
                if (lane == 0) mbar_expect_tx(&full_bar[slot], seg_bytes * rows_here);
                if (src)
                    bulk_g2s(dst_lane + (size_t)slot * STAGE, src + (size_t)seg * kSegBytes, seg_bytes,
                             &full_bar[slot]);
                if (++slot == nstages) {
                    slot = 0;
                    if (wrapped) phase ^= 1;
                    wrapped = true;
                }
            }
        }
        return;
    }

    // ========================= consumers =========================
    // x (and the optional RMSNorm prologue) depend on the predecessor kernel.
    pdl_wait();
    unsigned p2p_call = 0;   // number of the allreduce call these outputs belong to (p2p_allreduce.cu)
    if (MODE == kModePlain && p.wire.world > 0)
        p2p_call = *reinterpret_cast<volatile unsigned*>(p.wire.peer[p.wire.rank]) + 1u;
    {
        const int pk = p.K / E;  // packets per token
        const T* xg = reinterpret_cast<const T*>(p.x);
        float ss[NB];
#pragma unroll
        for (int b = 0; b < NB; ++b) ss[b] = 0.f;
        for (int i = threadIdx.x; i < pk; i += kGemvWarps * 32) {
#pragma unroll
            for (int b = 0; b < NB; ++b) {
                if (b >= p.nb) break;
                Pack16 v = ld16(xg + (size_t)b * p.K + (size_t)i * E);
                st16(xs + (size_t)b * xp + (size_t)i * E, v);
                if (p.flags & B200_GEMV_NORM) {
                    float f[E];
                    Unpack<T>::get(v, f);
#pragma unroll
                    for (int e = 0; e < E; ++e) ss[b] = fmaf(f[e], f[e], ss[b]);
                }
            }
        }
        if (p.flags & B200_GEMV_NORM) {
#pragma unroll
            for (int b = 0; b < NB; ++b) {
                const float s = warp_sum(ss[b]);
                if (lane == 0) red[b * kGemvWarps + warp] = s;
            }
            consumer_bar();
            if (threadIdx.x < NB) {
                float s = 0.f;
                for (int w = 0; w < kGemvWarps; ++w) s += red[threadIdx.x * kGemvWarps + w];
                red[NB * kGemvWarps + threadIdx.x] = rsqrtf(s / (float)p.K + p.eps);
            }
            consumer_bar();
            const T* gm = reinterpret_cast<const T*>(p.gamma);
            const bool gfirst = (p.flags & B200_GEMV_NORM_GFIRST) != 0;
            for (int i = threadIdx.x; i < pk; i += kGemvWarps * 32) {
                float g[E];
                Unpack<T>::get(ld16(gm + (size_t)i * E), g);
#pragma unroll
                for (int b = 0; b < NB; ++b) {
                    if (b >= p.nb) break;
                    const float inv = red[NB * kGemvWarps + b];
                    float f[E];
                    Unpack<T>::get(ld16(xs + (size_t)b * xp + (size_t)i * E), f);
#pragma unroll
                    for (int e = 0; e < E; ++e)
                        f[e] = gfirst ? (g[e] * f[e]) * inv
                                                               : (f[e] * inv) * g[e];
                    st16(xs + (size_t)b * xp + (size_t)i * E, Unpack<T>::put(f));  // rounds to T
                }
            }
        }
        consumer_bar();
    }

    int step = p.step;
    if (MODE == kModeRopeKV) {
        if (p.step_dev) step = *p.step_dev;
    }

    float acc_a[NB], acc_b[NB];
#pragma unroll
    for (int b = 0; b < NB; ++b) acc_a[b] = acc_b[b] = 0.f;

    const int last_seg_pk = (p.K - (nseg - 1) * KSEG) / E;
    int slot = 0;
    uint32_t full_phase = 0;
    for (int g = 0; g < ngroups; ++g) {
      const int unit = u0 + g * kGemvWarps + warp;
      const bool active = unit < u1;
      int ra = 0, rb = -1;
      if (active) unit_rows<MODE>(p, unit, ra, rb);
      float cm[4] = {0.f, 0.f, 0.f, 0.f};   // MMA: this warp's k-slice of rows (lane/4, lane/4 + 8) x tokens (2(lane%4), +1)
      for (int seg = 0; seg < nseg; ++seg) {
        mbar_wait(&full_bar[slot], full_phase);
        if constexpr (MMA) {
            const int gq = lane >> 2, tq = lane & 3;
            const int seg_len = (seg == nseg - 1) ? last_seg_pk * E : KSEG;   // elements; a multiple of 16 (host check)
            const unsigned char* st = ring + (size_t)slot * STAGE;
            const unsigned char* ar0 = st + (size_t)gq * PITCH;
            const unsigned char* ar1 = st + (size_t)(gq + 8) * PITCH;
            const T* xr = xs + (size_t)gq * xp + (size_t)seg * KSEG;   // token gq (zero fragment when gq >= nb)
            const bool tok = gq < p.nb;
            constexpr int KW = KSEG / kGemvWarps;   // this warp's k slice of the segment
#pragma unroll
            for (int i = 0; i < KW / 16; ++i) {
                const int kk = warp * KW + i * 16;
                if (kk < seg_len) {
                    const int o = (kk + 2 * tq) * 2;
                    const uint32_t a0 = *reinterpret_cast<const uint32_t*>(ar0 + o);
                    const uint32_t a1 = *reinterpret_cast<const uint32_t*>(ar1 + o);
                    const uint32_t a2 = *reinterpret_cast<const uint32_t*>(ar0 + o + 16);
                    const uint32_t a3 = *reinterpret_cast<const uint32_t*>(ar1 + o + 16);
                    uint32_t b0 = 0u, b1 = 0u;
                    if (tok) {
                        b0 = *reinterpret_cast<const uint32_t*>(xr + kk + 2 * tq);
                        b1 = *reinterpret_cast<const uint32_t*>(xr + kk + 2 * tq + 8);
                    }
                    mma_f16_16816(cm, a0, a1, a2, a3, b0, b1);
                }
            }
        } else if (active) {
            const int seg_pk = (seg == nseg - 1) ? last_seg_pk : KSEG / E;
            const unsigned char* sa = ring + (size_t)slot * kStageBytes + (2 * warp) * kSegBytes;
            const unsigned char* sb = sa + kSegBytes;
            const T* xseg = xs + (size_t)seg * KSEG;
#pragma unroll
            for (int j0 = 0; j0 < kJ; j0 += 4) {   // four packets per lane in flight (register budget)
                Pack16 wa[4], wb[4];
#pragma unroll
                for (int j = 0; j < 4; ++j) {
                    const int pkt = lane + 32 * (j0 + j);
                    if (pkt < seg_pk) {
                        wa[j] = ld16(sa + pkt * 16);
                        if (rb >= 0) wb[j] = ld16(sb + pkt * 16);
                    }
                }
#pragma unroll
                for (int j = 0; j < 4; ++j) {
                    const int pkt = lane + 32 * (j0 + j);
                    if (pkt < seg_pk) {
#pragma unroll
                        for (int b = 0; b < NB; ++b) {
                            const Pack16 xv = ld16(xseg + (size_t)b * xp + (size_t)pkt * E);
                            acc_a[b] = dot16<T>(wa[j], xv, acc_a[b]);
                            if (rb >= 0) acc_b[b] = dot16<T>(wb[j], xv, acc_b[b]);
                        }
                    }
                }
            }
        }
        __syncwarp();
        if (lane == 0) mbar_arrive(&empty_bar[slot]);
        if (++slot == nstages) {
            slot = 0;
            full_phase ^= 1;
        }

        if constexpr (MMA) {
            if (seg == nseg - 1) {
                // ---- end of the row group: the eight warps' k-slices meet; warp w then owns rows 2w, 2w+1 as before
                const int gq = lane >> 2, tq = lane & 3;
                float* part = red_mma + warp * 128;
                part[gq * 8 + 2 * tq] = cm[0];
                part[gq * 8 + 2 * tq + 1] = cm[1];
                part[(gq + 8) * 8 + 2 * tq] = cm[2];
                part[(gq + 8) * 8 + 2 * tq + 1] = cm[3];
                consumer_bar();
                float sum = 0.f;
                if (lane < 16) {
#pragma unroll
                    for (int w = 0; w < kGemvWarps; ++w) sum += red_mma[w * 128 + (2 * warp + (lane >> 3)) * 8 + (lane & 7)];
                }
                consumer_bar();   // red_mma is free for the next group
#pragma unroll
                for (int b = 0; b < NB; ++b) {
                    acc_a[b] = __shfl_sync(0xffffffffu, sum, b);
                    acc_b[b] = __shfl_sync(0xffffffffu, sum, 8 + b);
                }
            }
        }
        if (seg == nseg - 1 && active) {
            // ---- end of the row pair: reduce + epilogue (lane 0 writes) ----
            if constexpr (!MMA) {
#pragma unroll
                for (int b = 0; b < NB; ++b) {
                    acc_a[b] = warp_sum(acc_a[b]);
                    acc_b[b] = warp_sum(acc_b[b]);
                }
            }
            if (lane == 0) {
                const T* bias = reinterpret_cast<const T*>(p.bias);
                T* y = reinterpret_cast<T*>(p.y);
                if (MODE == kModeSwiglu) {
#pragma unroll
                    for (int b = 0; b < NB; ++b) {
                        if (b >= p.nb) break;
                        float gv = acc_a[b], uv = acc_b[b];
                        if (bias) { gv += DT<T>::to_f(bias[ra]); uv += DT<T>::to_f(bias[rb]); }
                        gv = DT<T>::to_f(DT<T>::from_f(gv));  // the unfused path stores gate/up in T
                        uv = DT<T>::to_f(DT<T>::from_f(uv));
                        const float s = DT<T>::to_f(DT<T>::from_f(gv / (1.f + expf(-gv))));
                        y[(size_t)b * p.N + ra] = DT<T>::from_f(s * uv);
                    }
                } else if (MODE == kModeRopeKV) {
                    const int d = p.head_size, hd = d >> 1;
                    const int head = ra / d, i = ra - head * d;
                    const int qkv_heads = p.head_num + 2 * p.kv_head_num;
                    const bool rotated = head < p.head_num + p.kv_head_num && i < (p.rotary_dim >> 1);
                    float c = 1.f, s = 0.f;
                    if (rotated) {
                        const float ang = p.rope_den ? (float)(step - 1) / p.rope_den[i]
                                                     : rope_angle(i, p.rotary_dim, p.rotary_base, (float)(step - 1));
                        c = cosf(ang);
                        s = sinf(ang);
                    }
#pragma unroll
                    for (int b = 0; b < NB; ++b) {
                        if (b >= p.nb) break;
                        if (p.step_stride) {   // ragged batch: token b sits at its own position
                            step = p.step_dev[b * p.step_stride];
                            if (rotated) {
                                const float ang = p.rope_den ? (float)(step - 1) / p.rope_den[i]
                                                             : rope_angle(i, p.rotary_dim, p.rotary_base, (float)(step - 1));
                                c = cosf(ang);
                                s = sinf(ang);
                            }
                        }
                        float av = acc_a[b], bv = acc_b[b];
                        if (bias) { av += DT<T>::to_f(bias[ra]); bv += DT<T>::to_f(bias[rb]); }
                        av = DT<T>::to_f(DT<T>::from_f(av));  // qkv_buf is stored in T before RoPE
                        bv = DT<T>::to_f(DT<T>::from_f(bv));
                        const T oa = DT<T>::from_f(c * av - s * bv);
                        const T ob = DT<T>::from_f(c * bv + s * av);
                        T* yt = y + (size_t)b * qkv_heads * d;
                        yt[ra] = oa;
                        yt[rb] = ob;
                        if (head >= p.head_num) {
                            const bool is_k = head < p.head_num + p.kv_head_num;
                            const int kvh = is_k ? head - p.head_num : head - p.head_num - p.kv_head_num;
                            T* cache = reinterpret_cast<T*>(is_k ? p.k_cache : p.v_cache);
                            T* row = cache + (((size_t)b * p.kv_head_num + kvh) * p.max_seq_len +
                                              (size_t)(step - 1)) * d;
                            row[i] = oa;
                            row[i + hd] = ob;
                        }
                    }
                } else {
#pragma unroll
                    for (int b = 0; b < NB; ++b) {
                        if (b >= p.nb) break;
                        float av = acc_a[b], bv = acc_b[b];
                        if (MODE == kModeResidual) {
                            // unfused: linear output rounded to T, then T(out + bias + residual)
                            av = DT<T>::to_f(DT<T>::from_f(av));
                            bv = DT<T>::to_f(DT<T>::from_f(bv));
                            if (bias) { av += DT<T>::to_f(bias[ra]); if (rb >= 0) bv += DT<T>::to_f(bias[rb]); }
                            av += DT<T>::to_f(y[(size_t)b * p.N + ra]);
                            if (rb >= 0) bv += DT<T>::to_f(y[(size_t)b * p.N + rb]);
                        } else if (bias) {
                            av += DT<T>::to_f(bias[ra]);
                            if (rb >= 0) bv += DT<T>::to_f(bias[rb]);
                        }
                        if (MODE == kModePlain && sizeof(T) == 2 && p.wire.world > 0) {
                            // tensor-parallel row-split linear: element pair (b, ra..ra+1) is half of a 16-byte
                            // low-latency line {data, call, data, call} in slot `rank` of EVERY rank's receive
                            // buffer (p2p_allreduce.cu); the 8-byte store is atomic, the tag makes it valid
                            const unsigned call = p2p_call;
                            const size_t e = (size_t)(b + p.token0) * p.N + ra;
                            const size_t off = kP2PHeader + (size_t)(call & 1u) * p.wire.world * p.wire.slot_bytes +
                                               (size_t)p.wire.rank * p.wire.slot_bytes + (e >> 2) * 16 + ((e >> 1) & 1) * 8;
                            const __half2 hv = __floats2half2_rn(av, bv);
                            const unsigned data = *reinterpret_cast<const unsigned*>(&hv);
                            for (int r = 0; r < p.wire.world; ++r)
                                asm volatile("st.volatile.global.v2.u32 [%0], {%1,%2};" ::"l"(p.wire.peer[r] + off), "r"(data),
                                             "r"(call)
                                             : "memory");
                            continue;
                        }
                        y[(size_t)b * p.N + ra] = DT<T>::from_f(av);
                        if (rb >= 0) y[(size_t)b * p.N + rb] = DT<T>::from_f(bv);
                    }
                }
            }
#pragma unroll
            for (int b = 0; b < NB; ++b) acc_a[b] = acc_b[b] = 0.f;
        }
      }
    }
    if (MODE == kModeRopeKV) pdl_launch_dependents();
}


// ------------------------------------------------------------------------------------------
// host side
// ------------------------------------------------------------------------------------------
static size_t gemv_smem_bytes(int stages, int NB, int K, int esz, bool mma = false) {
    size_t x = (((size_t)NB * (K + (mma ? kXPadMma : 0)) * esz) + 15) & ~(size_t)15;
    return (size_t)stages * (mma ? kStageBytesMma : kStageBytes) + x + 2 * kMaxStages * sizeof(uint64_t) +
           (size_t)(NB * kGemvWarps + NB) * sizeof(float) + (mma ? (size_t)kGemvWarps * 128 * sizeof(float) : 0) + 16;
}

// tensor-core consumer: fp16, 2..8 token rows, K a multiple of 16 (B200_GEMV_MMA=0 keeps the scalar loop, for A/B runs)
static bool gemv_use_mma(int NB, int K, int esz) {
    static int on = -1;
    if (on < 0) { const char* e = getenv("B200_GEMV_MMA"); on = (e && e[0] == '0') ? 0 : 1; }
    return on && esz == 2 && NB >= 2 && (K & 15) == 0;
}

static int gemv_max_stages() {
    static int v = 0;
    if (v == 0) {
        const char* e = getenv("B200_GEMV_STAGES");
        v = e ? atoi(e) : 3;  // 3 x 32 KB: two GEMV kernels fit one SM, so PDL prefetch can overlap
        if (v < 2) v = 2;
        if (v > kMaxStages) v = kMaxStages;
    }
    return v;
}

static int gemv_pick_stages(int NB, int K, int esz) {
    const bool mma = gemv_use_mma(NB, K, esz);
    for (int st = gemv_max_stages(); st >= 2; --st)
        if (gemv_smem_bytes(st, NB, K, esz, mma) <= 227 * 1024) return st;
    return 0;
}

template <typename T, int NB, int MODE, bool MMA = false>
static int gemv_launch_one(const GemvParams& p, size_t smem, int grid, cudaStream_t s) {
    auto kern = gemv_stream_kernel<T, NB, MODE, MMA>;
    static unsigned long long attr_done = 0;  // per instantiation, one bit per device
    {
        cudaError_t e = ensure_max_dynamic_smem(reinterpret_cast<const void*>(kern), 227 * 1024, attr_done);
        if (e != cudaSuccess) return (int)e;
    }
    cudaError_t e = launch_k(kern, dim3(grid), dim3(kGemvThreads), smem, s, true, p);
    count_launch();
    return (int)e;
}

template <typename T, int MODE>
static int gemv_dispatch_nb(GemvParams p, int n_tokens, cudaStream_t s) {
    const int esz = (int)sizeof(T);
    int NB = n_tokens <= 1 ? 1 : n_tokens <= 2 ? 2 : n_tokens <= 4 ? 4 : 8;
    int stages = gemv_pick_stages(NB, p.K, esz);
    if (stages == 0) {
        // x does not fit in shared memory next to the ring: run the tokens in smaller tiles
        // (the weights are streamed once per tile). Token tiles are independent in every mode.
        if (NB == 1) return B200_EUNSUPPORTED;
        const int half = NB / 2;
        const int y_stride = (MODE == kModeRopeKV) ? p.N : p.N;  // y is [tokens, N] in all modes
        for (int t0 = 0; t0 < n_tokens; t0 += half) {
            GemvParams q = p;
            q.x = (const char*)p.x + (size_t)t0 * p.K * esz;
            q.y = (char*)p.y + (size_t)t0 * y_stride * esz;
            q.token0 = p.token0 + t0;
            if (p.step_dev) q.step_dev = p.step_dev + (size_t)t0 * p.step_stride;
            if (MODE == kModeRopeKV) {
                const size_t per_seq = (size_t)p.kv_head_num * p.max_seq_len * p.head_size * esz;
                q.k_cache = (char*)p.k_cache + (size_t)t0 * per_seq;
                q.v_cache = (char*)p.v_cache + (size_t)t0 * per_seq;
            }
            const int nt = (n_tokens - t0 < half) ? n_tokens - t0 : half;
            int rc = gemv_dispatch_nb<T, MODE>(q, nt, s);
            if (rc != 0) return rc;
        }
        return B200_OK;
    }
    p.stages = stages;
    p.nb = n_tokens;
    const bool mma = gemv_use_mma(NB, p.K, esz);
    const size_t smem = gemv_smem_bytes(stages, NB, p.K, esz, mma);
    int grid = sm_count();
    const int groups = (p.units + kGemvWarps - 1) / kGemvWarps;
    if (grid > groups) grid = groups;
    if (grid < 1) grid = 1;
    if constexpr (sizeof(T) == 2) {
        if (mma) {
            switch (NB) {
                case 2: return gemv_launch_one<T, 2, MODE, true>(p, smem, grid, s);
                case 4: return gemv_launch_one<T, 4, MODE, true>(p, smem, grid, s);
                default: return gemv_launch_one<T, 8, MODE, true>(p, smem, grid, s);
            }
        }
    }
    switch (NB) {
        case 1: return gemv_launch_one<T, 1, MODE>(p, smem, grid, s);
        case 2: return gemv_launch_one<T, 2, MODE>(p, smem, grid, s);
        case 4: return gemv_launch_one<T, 4, MODE>(p, smem, grid, s);
        default: return gemv_launch_one<T, 8, MODE>(p, smem, grid, s);
    }
}

template <typename T>
static int gemv_fused_impl(const b200_gemv_args* a, cudaStream_t s) {
    constexpr int E = DT<T>::kPer16B;
    B200_RETURN_IF(a->K % E != 0, B200_EUNSUPPORTED);
    B200_RETURN_IF(a->n_tokens < 1 || a->n_tokens > B200_GEMV_MAX_TOKENS, B200_EUNSUPPORTED);
    const unsigned epi = a->flags & (B200_GEMV_RESIDUAL | B200_GEMV_SWIGLU | B200_GEMV_ROPE_KV);
    B200_RETURN_IF(epi & (epi - 1), B200_EINVAL);  // at most one epilogue mode
    B200_RETURN_IF((a->flags & B200_GEMV_NORM) && !a->gamma, B200_EINVAL);
    GemvParams p{};
    p.x = a->x; p.W = a->W; p.bias = a->bias; p.y = a->y; p.gamma = a->gamma; p.eps = a->eps;
    p.N = a->N; p.K = a->K;
    p.flags = a->flags & (B200_GEMV_NORM | B200_GEMV_NORM_GFIRST);
    p.k_cache = a->k_cache_layer; p.v_cache = a->v_cache_layer;
    p.head_num = a->head_num; p.kv_head_num = a->kv_head_num; p.head_size = a->head_size;
    p.max_seq_len = a->max_seq_len; p.step = a->step; p.rotary_dim = a->rotary_dim;
    p.rotary_base = a->rotary_base; p.step_dev = a->step_dev; p.rope_den = a->rope_den;
    p.step_stride = (a->step_dev && a->step_stride) ? 1 : 0;
    if (a->p2p) {
        const b200_p2p* c = reinterpret_cast<const b200_p2p*>(a->p2p);
        B200_RETURN_IF(epi != 0 || sizeof(T) != 2 || (a->N & 3) || a->bias, B200_EUNSUPPORTED);
        B200_RETURN_IF((size_t)a->n_tokens * a->N * 4 > c->slot_bytes || (!c->connected && c->world > 1), B200_EINVAL);
        p.wire = p2p_wire(c);
    }
    if (epi == B200_GEMV_SWIGLU) {
        p.units = a->N;
        return gemv_dispatch_nb<T, kModeSwiglu>(p, a->n_tokens, s);
    }
    if (epi == B200_GEMV_ROPE_KV) {
        B200_RETURN_IF(!a->k_cache_layer || !a->v_cache_layer, B200_EINVAL);
        B200_RETURN_IF(a->head_size <= 0 || (a->head_size & 1) || a->head_num <= 0 ||
                           a->kv_head_num <= 0 ||
                           a->N != (a->head_num + 2 * a->kv_head_num) * a->head_size,
                       B200_EINVAL);
        B200_RETURN_IF(a->rotary_dim <= 0 || a->rotary_dim > a->head_size || (a->rotary_dim & 1),
                       B200_EINVAL);
        B200_RETURN_IF(!a->step_dev && (a->step < 1 || a->step > a->max_seq_len), B200_EINVAL);
        p.units = a->N / 2;
        return gemv_dispatch_nb<T, kModeRopeKV>(p, a->n_tokens, s);
    }
    p.units = (a->N + 1) / 2;
    if (epi == B200_GEMV_RESIDUAL) return gemv_dispatch_nb<T, kModeResidual>(p, a->n_tokens, s);
    return gemv_dispatch_nb<T, kModePlain>(p, a->n_tokens, s);
}

// used by b200_linear (linear.cu): plain GEMV, any n_tokens (tiled by B200_GEMV_MAX_TOKENS)
int gemv_plain(const void* x, const void* W, const void* bias, void* y, int n_tokens, int N, int K,
               int dtype, cudaStream_t s) {
    const int esz = dtype == B200_F16 ? 2 : 4;
    for (int t0 = 0; t0 < n_tokens; t0 += B200_GEMV_MAX_TOKENS) {
        b200_gemv_args a{};
        a.x = (const char*)x + (size_t)t0 * K * esz;
        a.W = W;
        a.bias = bias;
        a.y = (char*)y + (size_t)t0 * N * esz;
        a.n_tokens = (n_tokens - t0 < B200_GEMV_MAX_TOKENS) ? n_tokens - t0 : B200_GEMV_MAX_TOKENS;
        a.N = N;
        a.K = K;
        int rc = dtype == B200_F16 ? gemv_fused_impl<__half>(&a, s) : gemv_fused_impl<float>(&a, s);
        if (rc != 0) return rc;
    }
    return B200_OK;
}

}  // namespace b200

__global__ void rope_table_kernel(float* den, int rot_dim, float base) {
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i < rot_dim / 2) den[i] = (float)pow((double)base, (double)((float)(2 * i) / (float)rot_dim));
}

extern "C" int b200_rope_table(float* den, int rotary_dim, float rotary_base, void* stream) {
    B200_RETURN_IF(!den || rotary_dim <= 0 || (rotary_dim & 1), B200_EINVAL);
    rope_table_kernel<<<(rotary_dim / 2 + 63) / 64, 64, 0, (cudaStream_t)stream>>>(den, rotary_dim, rotary_base);
    b200::count_launch();
    return (int)cudaGetLastError();
}

extern "C" int b200_gemv_fused(const b200_gemv_args* args, int dtype, void* stream) {
    B200_RETURN_IF(!args || !args->x || !args->W || !args->y, B200_EINVAL);
    B200_RETURN_IF(args->N <= 0 || args->K <= 0, B200_EINVAL);
    auto al = [](const void* p) { return (reinterpret_cast<uintptr_t>(p) & 15u) == 0; };
    B200_RETURN_IF(!al(args->x) || !al(args->W) || !al(args->gamma), B200_EINVAL);
    cudaStream_t s = (cudaStream_t)stream;
    if (dtype == B200_F32) return b200::gemv_fused_impl<float>(args, s);
    if (dtype == B200_F16) return b200::gemv_fused_impl<__half>(args, s);
    return B200_EINVAL;
}
