// Persistent whole-step decode kernel (fp16, batch <= 4, no tensor parallelism): ONE launch runs
// embedding -> 32 x (qkv GEMV+RMSNorm+RoPE+KV append -> split-KV attention -> o_proj+residual ->
// gate_up GEMV+RMSNorm+SwiGLU -> down+residual) -> final norm + LM head -> greedy pick -> advance,
// i.e. Llama<T>::continueTokenGen (src/models/llama/llama.cpp:319-361) with the layer loop of
// self_decoder.cpp:53-85, masked_self_attention.cpp:55-92, ffn.cpp:53-93 and the LM head + top-1
// of llama.cpp:215-271. Arithmetic and rounding points are those of the multi-kernel fused path
// (gemv.cu / decode_attention.cu), which stays as the reference for it in the tests.
//
// Why one kernel: at batch 1 a step is ~160 dependent bandwidth phases of 5-30 us each. Launched
// separately, every phase pays launch + ramp + tail and the weight prefetch of the next kernel is
// capped by what two co-resident CTAs can hold. Here one CTA per SM stays resident (cooperative
// launch), a producer warp streams the WHOLE step's weights through one deep smem ring with 1-D
// bulk copies (cp.async.bulk, UBLKCP) and never waits for a phase boundary, and the 8 consumer
// warps walk a small program of ops separated by grid barriers; while consumers wait at a
// barrier the ring (up to ~190 KB per SM, ~28 MB chip-wide) keeps HBM busy.
//
// HBM-bound: algorithmic bytes per step = all weights once + KV cache rows [0, step) once.
#include <cstdio>
#include <vector>

#include "common.cuh"
#include "decode_step.h"
#include "launch_count.h"

namespace b200 {

constexpr int kStepWarps = 8;                         // consumer warps
constexpr int kStepThreads = (kStepWarps + 1) * 32;   // + producer warp
constexpr int kStepConsumers = kStepWarps * 32;
// K-segment bytes per row per stage = the size of one bulk copy. 4 KB (round 2; was 2 KB): a stage is 64 KB, the ring
// 3 stages. The SM's copy engine costs ~27 ns + bytes / 124 GB/s per copy, so 2 KB copies cap an SM at 47.8 GB/s - just
// under its fair share of HBM - and 4 KB copies at 68 GB/s; the consumers pay their ~200-cycle hand-over once per 64 KB
// instead of once per 32 KB. Round 1 measured 4 KB copies SLOWER (HBM saturated -> uneven service -> every one of the 5
// grid barriers per layer waits for the slowest CTA); with 2 barriers per layer the balance flips: 2.70 -> 2.53 ms.
#ifndef B200_STEP_SEG_KB
#define B200_STEP_SEG_KB 4
#endif
constexpr int kSegB = B200_STEP_SEG_KB * 1024;
constexpr int kStageB = kStepWarps * 2 * kSegB;       // 64 KB: 8 row pairs x 4 KB
constexpr int kStepMaxStages = 6;
constexpr int kStepScratchBytes = 6 * 1024;            // split-KV merge weights: nb*h*(splits+1) floats
constexpr int kJS = kSegB / 16 / 32;                  // 16-byte packets per lane per segment

// ------------------------------------------------------------------------------------------
// small device helpers
// ------------------------------------------------------------------------------------------
// L2-coherent 16-byte load for activations another CTA of THIS kernel may have written
// (L1 is not coherent across SMs; .cg caches at L2 only)
__device__ __forceinline__ Pack16 ld16_cg(const void* p) {
    Pack16 r;
    asm volatile("ld.global.cg.v4.u32 {%0,%1,%2,%3}, [%4];"
                 : "=r"(r.w[0]), "=r"(r.w[1]), "=r"(r.w[2]), "=r"(r.w[3])
                 : "l"(p)
                 : "memory");
    return r;
}
// non-volatile variant: the compiler may hoist / batch these (barriers carry memory clobbers, so
// they never move across a grid barrier)
__device__ __forceinline__ Pack16 ld16_cg_nv(const void* p) {
    const uint4 v = __ldcg(reinterpret_cast<const uint4*>(p));
    Pack16 r;
    r.w[0] = v.x; r.w[1] = v.y; r.w[2] = v.z; r.w[3] = v.w;
    return r;
}
__device__ __forceinline__ float ldf_cg(const float* p) {
    float v;
    asm volatile("ld.global.cg.f32 %0, [%1];" : "=f"(v) : "l"(p) : "memory");
    return v;
}
__device__ __forceinline__ __half ldh_cg(const __half* p) {
    unsigned short v;
    asm volatile("ld.global.cg.u16 %0, [%1];" : "=h"(v) : "l"(p) : "memory");
    return __ushort_as_half(v);
}
__device__ __forceinline__ unsigned ld_acquire_u32(const unsigned* p) {
    unsigned v;
    asm volatile("ld.acquire.gpu.global.u32 %0, [%1];" : "=r"(v) : "l"(p) : "memory");
    return v;
}
__device__ __forceinline__ unsigned long long global_timer_ns() {
    unsigned long long t;
    asm volatile("mov.u64 %0, %globaltimer;" : "=l"(t));
    return t;
}
__device__ __forceinline__ void consumers_sync() {
    asm volatile("bar.sync 1, %0;" ::"n"(kStepConsumers) : "memory");
}

// grid-wide barrier among the consumer threads of all CTAs (all CTAs are co-resident: cooperative
// launch). `target` = arrivals expected so far. A 2 s timeout turns a would-be hang into a trap.
__device__ __forceinline__ void grid_barrier_arrive(unsigned* sync) {
    consumers_sync();   // every consumer's writes of this op are ordered before thread 0's release
    if (threadIdx.x == 0) asm volatile("red.release.gpu.global.add.u32 [%0], 1;" ::"l"(sync) : "memory");
}
__device__ __forceinline__ void grid_barrier_wait(unsigned* sync, unsigned target) {
    if (threadIdx.x == 0) {
        const unsigned long long t0 = global_timer_ns();
        while ((int)(ld_acquire_u32(&sync[0]) - target) < 0) {
            if (global_timer_ns() - t0 > 2000000000ull) {
                atomicExch(&sync[2], 1u);
                __trap();
            }
        }
    }
    consumers_sync();
}

// Dynamic shared memory of the step kernel: [ring: stages x kStageB][xs: x rows / attention merge][full / empty mbarriers]
// [red: NB*8 + NB + 32 floats][scratch: kStepScratchBytes]. The noinline consumers recompute these few pointers from the
// parameters (which live in shared memory) instead of receiving them as arguments: arguments beyond the register
// window travel through local memory, and local memory is an L2 round trip in this kernel.
template <int NB>
struct StepSmem {
    unsigned char* ring;
    __half* xs;
    uint64_t* full;
    uint64_t* empty;
    float* red;
    float* scratch;
    long long* starved;   // trace mode: cycles warp 0 waited for weights in the current op
    __device__ __forceinline__ StepSmem(unsigned char* smem, int stages, size_t xs_bytes) {
        ring = smem;
        xs = reinterpret_cast<__half*>(smem + (size_t)stages * kStageB);
        full = reinterpret_cast<uint64_t*>(smem + (size_t)stages * kStageB + xs_bytes);
        empty = full + kStepMaxStages;
        red = reinterpret_cast<float*>(empty + kStepMaxStages);
        scratch = red + (NB * kStepWarps + NB + 32);
        starved = reinterpret_cast<long long*>((reinterpret_cast<uintptr_t>(scratch) + kStepScratchBytes - 16) & ~(uintptr_t)7);
    }
};

// a copy of v the compiler cannot merge with other copies (see the call sites in decode_step_kernel)
__device__ __forceinline__ int opaque_i32(int v) {
    asm volatile("" : "+r"(v));
    return v;
}

struct Ring {
    unsigned char* base;
    uint64_t* full;
    uint64_t* empty;
    int stages;
    int slot;
    uint32_t phase;
};

template <int EPI>
__device__ __forceinline__ void op_unit_rows(const StepOp& op, int head_size, int u, int& ra, int& rb) {
    if (EPI == kEpiSwiglu || EPI == kEpiSwigluDown) {
        ra = u;
        rb = op.N + u;
    } else if (EPI == kEpiRopeKV) {
        const int hd = head_size >> 1;
        const int head = u / hd, i = u - head * hd;
        ra = head * head_size + i;
        rb = ra + hd;
    } else {
        ra = 2 * u;
        rb = (2 * u + 1 < op.N) ? 2 * u + 1 : -1;
    }
}

// ------------------------------------------------------------------------------------------
// producer warp: streams the weights of every GEMV op of the step, in program order
// ------------------------------------------------------------------------------------------
__device__ __forceinline__ void op_rows_rt(const StepOp& op, int head_size, int u, int& ra, int& rb) {
    if (op.epi == kEpiSwiglu || op.epi == kEpiSwigluDown) op_unit_rows<kEpiSwiglu>(op, head_size, u, ra, rb);
    else if (op.epi == kEpiRopeKV) op_unit_rows<kEpiRopeKV>(op, head_size, u, ra, rb);
    else op_unit_rows<kEpiPlain>(op, head_size, u, ra, rb);
}

// Row pairs of a CTA are walked in ngroups = ceil(n / 8) groups of EQUAL size ceil(n / ngroups) <= 8 (one row pair per
// consumer warp and stage): 42 pairs are 6 groups of 7, not 5 of 8 and a tail of 2 whose four nearly empty stages
// each still cost the consumers a full stage hand-over. Producer and consumers use the same rule.
__device__ __forceinline__ int group_size(int n_units, int ngroups) { return ngroups > 0 ? (n_units + ngroups - 1) / ngroups : 0; }

// this CTA's slice of one GEMV op, as the producer sees it
struct OpSlice {
    const unsigned char* W;
    const unsigned char* W2;   // kEpiSwigluDown: transposed down weights (rows [u0, u1) are this CTA's), else nullptr
    int epi, N, K, u0, u1, ngroups, gsz;
};
__device__ __forceinline__ bool load_slice(const StepOp* ops, int oi, OpSlice& sl) {
    if (ops[oi].type == kOpAttnO) {   // no row phase; W2 block of item blockIdx.x, K = bytes / 2 of the whole block
        sl.W = nullptr;
        sl.epi = -1; sl.N = ops[oi].N; sl.K = (ops[oi].N / ops[oi].act_stride) * (ops[oi].K / (ops[oi].units / ops[oi].act_stride));
        sl.W2 = (int)blockIdx.x < ops[oi].units ? reinterpret_cast<const unsigned char*>(ops[oi].W2) : nullptr;
        sl.u0 = blockIdx.x; sl.u1 = blockIdx.x + 1; sl.ngroups = 0; sl.gsz = 0;
        return sl.W2 != nullptr;
    }
    if (ops[oi].type != kOpGemv) return false;
    sl.W = reinterpret_cast<const unsigned char*>(ops[oi].W);
    sl.epi = ops[oi].epi; sl.N = ops[oi].N; sl.K = ops[oi].K;
    sl.W2 = sl.epi == kEpiSwigluDown ? reinterpret_cast<const unsigned char*>(ops[oi].W2) : nullptr;
    const int units = ops[oi].units;
    sl.u0 = (int)(((long long)units * blockIdx.x) / gridDim.x);
    sl.u1 = (int)(((long long)units * (blockIdx.x + 1)) / gridDim.x);
    sl.ngroups = (sl.u1 - sl.u0 + kStepWarps - 1) / kStepWarps;
    sl.gsz = group_size(sl.u1 - sl.u0, sl.ngroups);
    return true;
}
// source row of this lane (lane l -> row (l & 1) of unit (l >> 1)) in group g; nullptr if none
__device__ __forceinline__ const unsigned char* slice_lane_row(const OpSlice& sl, int g, int lane, int head_size) {
    const int ubase = sl.u0 + g * sl.gsz;
    const int nu = min(sl.gsz, sl.u1 - ubase);
    if (lane >= 2 * nu) return nullptr;
    const int u = ubase + (lane >> 1);
    int ra, rb;
    if (sl.epi == kEpiSwiglu || sl.epi == kEpiSwigluDown) { ra = u; rb = sl.N + u; }
    else if (sl.epi == kEpiRopeKV) {
        const int hd = head_size >> 1;
        const int head = u / hd, i = u - head * hd;
        ra = head * head_size + i;
        rb = ra + hd;
    } else { ra = 2 * u; rb = (2 * u + 1 < sl.N) ? 2 * u + 1 : -1; }
    const int r = (lane & 1) ? rb : ra;
    return r >= 0 ? sl.W + (size_t)r * ((size_t)sl.K * 2) : nullptr;
}

// position of the producer in the step's sequence of ring stages. Phase 1 of an op: (row group, K segment), one
// 2 KB row segment per lane. Phase 2 (kEpiSwigluDown only): the CTA's rows [u0, u1) of W2 are ONE contiguous byte
// range, streamed as stages of kStageB bytes in 2 KB pieces (lane l = piece l).
struct Cursor {
    OpSlice sl;
    const unsigned char* src;   // phase 1: this lane's source row in the current group (nullptr: lane idle)
    int oi, g, seg, nseg;
    uint32_t rows_here, last_seg_bytes;
    int s2, n2;                 // phase 2: stage index / number of stages (0: the op has no second phase)
    long long bytes2;           // phase 2: bytes of the CTA's W2 slice
    bool valid, phase2;
};
__device__ __forceinline__ void cursor_seek(Cursor& c, const StepOp* ops, int n_ops, int lane, int head_size) {
    // first GEMV op at or after c.oi that has rows for this CTA
    c.valid = false;
    c.phase2 = false;
    while (c.oi < n_ops) {
        if (load_slice(ops, c.oi, c.sl) && (c.sl.ngroups > 0 || c.sl.epi < 0)) {
            constexpr int KSEG = kSegB / 2;
            c.nseg = (c.sl.K + KSEG - 1) / KSEG;
            c.last_seg_bytes = (uint32_t)(c.sl.K - (c.nseg - 1) * KSEG) * 2u;
            c.g = 0;
            c.seg = 0;
            c.src = slice_lane_row(c.sl, 0, lane, head_size);
            c.rows_here = __popc(__ballot_sync(0xffffffffu, c.src != nullptr));
            c.s2 = 0;
            c.bytes2 = c.sl.W2 ? (long long)(c.sl.u1 - c.sl.u0) * c.sl.K * 2 : 0;
            c.n2 = (int)((c.bytes2 + kStageB - 1) / kStageB);
            c.phase2 = c.sl.ngroups == 0;   // kOpAttnO: straight to the contiguous phase
            c.valid = true;
            return;
        }
        ++c.oi;
    }
}
__device__ __forceinline__ void cursor_next(Cursor& c, const StepOp* ops, int n_ops, int lane, int head_size) {
    if (c.phase2) {
        if (++c.s2 < c.n2) return;
    } else {
        if (++c.seg < c.nseg) return;
        c.seg = 0;
        if (++c.g < c.sl.ngroups) {
            c.src = slice_lane_row(c.sl, c.g, lane, head_size);
            c.rows_here = __popc(__ballot_sync(0xffffffffu, c.src != nullptr));
            return;
        }
        if (c.n2 > 0) {
            c.phase2 = true;
            c.s2 = 0;
            return;
        }
    }
    ++c.oi;
    cursor_seek(c, ops, n_ops, lane, head_size);
}
__device__ __forceinline__ bool mbar_test(uint64_t* bar, uint32_t parity) {
    uint32_t done;
    asm volatile(
        "{\n"
        ".reg .pred p;\n"
        "mbarrier.test_wait.parity.shared::cta.b64 p, [%1], %2;\n"
        "selp.u32 %0, 1, 0, p;\n"
        "}\n"
        : "=r"(done)
        : "r"(smem_u32(bar)), "r"(parity)
        : "memory");
    return done != 0;
}

// Producer warp: issues bulk copies global -> ring stage as the consumers free slots, in program order and
// without ever stopping at an op boundary. (An L2-prefetch cursor running ahead of the ring while the consumers
// sit at a barrier was measured not to help - 2.88 -> 3.3 ms always on, no gain when stalled only: an SM ingests
// 2 KB bulk copies at 47.8 GB/s from L2 exactly as from HBM, DESIGN section 4 - and is gone from the code.)
__device__ __noinline__ void step_producer(const StepParams& p, Ring rg, int lane) {
    unsigned char* const r_base = rg.base;
    uint64_t* const r_full = rg.full;
    uint64_t* const r_empty = rg.empty;
    const int r_stages = rg.stages;
    const int n_ops = p.n_ops, head_size = p.head_size;
    const StepOp* ops = p.ops;

    int slot = 0, pass = 0;   // pass = number of times the ring has wrapped
    Cursor rc{};
    rc.oi = 0;
    cursor_seek(rc, ops, n_ops, lane, head_size);
    while (rc.valid) {
        if (pass > 0) mbar_wait(&r_empty[slot], (uint32_t)(pass - 1) & 1u);
        unsigned char* const dst = r_base + (size_t)slot * kStageB;
        if (!rc.phase2) {
            const uint32_t seg_bytes = (rc.seg == rc.nseg - 1) ? rc.last_seg_bytes : (uint32_t)kSegB;
            if (lane == 0) mbar_expect_tx(&r_full[slot], seg_bytes * rc.rows_here);
            if (rc.src) bulk_g2s(dst + (size_t)lane * kSegB, rc.src + (size_t)rc.seg * kSegB, seg_bytes, &r_full[slot]);
        } else {
            const long long off = (long long)rc.s2 * kStageB;
            const long long left = rc.bytes2 - off;
            const uint32_t stage_bytes = left < (long long)kStageB ? (uint32_t)left : (uint32_t)kStageB;
            if (lane == 0) mbar_expect_tx(&r_full[slot], stage_bytes);
            const uint32_t lo = (uint32_t)lane * kSegB;
            if (lo < stage_bytes)
                bulk_g2s(dst + lo, rc.sl.W2 + (size_t)rc.sl.u0 * rc.sl.K * 2 + off + lo,
                         stage_bytes - lo < (uint32_t)kSegB ? stage_bytes - lo : (uint32_t)kSegB, &r_full[slot]);
        }
        cursor_next(rc, ops, n_ops, lane, head_size);
        if (++slot == r_stages) {
            slot = 0;
            ++pass;
        }
    }
}

// ------------------------------------------------------------------------------------------
// consumers: x staging (+ RMSNorm / embedding / split-KV combine prologues)
// ------------------------------------------------------------------------------------------
template <int NB>
__device__ __noinline__ void stage_x(const StepOp& op, const StepParams& p, __half* xs, float* red, float* scratch) {
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const int K = op.K, pk = K / 8;
    float ss[NB];
#pragma unroll
    for (int b = 0; b < NB; ++b) ss[b] = 0.f;

    if (op.xsrc == kXCombine && p.splits > 1) {
        // x = softmax-merged split-KV partials, rounded to fp16 (the unfused path stores the
        // attention output in T before o_proj): fused_decoder_self_attention.cu:203-221,272-287.
        // Pass 1: one thread per (sequence, head) turns the split maxima / sums into merge weights;
        // pass 2: one thread per output element, loads of the S partial rows are independent.
        const int d = p.head_size, S = p.splits;
        const int nbh = p.nb * p.head_num;
        float* wtab = scratch;            // [nbh][S] merge weights
        float* inv_tab = scratch + nbh * S;   // [nbh]
        for (int bh = tid; bh < nbh; bh += kStepConsumers) {
            const float2* ml = reinterpret_cast<const float2*>(p.part_ml) + (size_t)bh * S;   // (m, l) pairs
            float M = -INFINITY;
            for (int s0 = 0; s0 < S; s0 += 4) {   // 4 independent L2 loads in flight
                float2 v[4];
#pragma unroll
                for (int u = 0; u < 4; ++u) v[u] = (s0 + u < S) ? __ldcg(ml + s0 + u) : make_float2(-INFINITY, 0.f);
#pragma unroll
                for (int u = 0; u < 4; ++u) M = fmaxf(M, v[u].x);
            }
            float L = 0.f;
            for (int s0 = 0; s0 < S; s0 += 4) {
                float2 v[4];
#pragma unroll
                for (int u = 0; u < 4; ++u) v[u] = (s0 + u < S) ? __ldcg(ml + s0 + u) : make_float2(-INFINITY, 0.f);
#pragma unroll
                for (int u = 0; u < 4; ++u) {
                    if (s0 + u < S) {
                        const float a = (v[u].x == -INFINITY) ? 0.f : expf(v[u].x - M);
                        wtab[bh * S + s0 + u] = a;
                        L += v[u].y * a;
                    }
                }
            }
            inv_tab[bh] = 1.f / (L + 1e-6f);
        }
        consumers_sync();
        // one thread per 4 consecutive output elements; the S partial rows are independent loads
        const int d4 = d / 4, total4 = nbh * d4;
        for (int idx = tid; idx < total4; idx += kStepConsumers) {
            const int bh = idx / d4, e4 = idx - bh * d4;
            const float4* po = reinterpret_cast<const float4*>(p.part_o + (size_t)bh * S * d) + e4;
            float4 O = make_float4(0.f, 0.f, 0.f, 0.f);
            for (int s0 = 0; s0 < S; s0 += 4) {
                float4 v[4];
#pragma unroll
                for (int u = 0; u < 4; ++u)
                    v[u] = (s0 + u < S) ? __ldcg(po + (size_t)(s0 + u) * d4) : make_float4(0.f, 0.f, 0.f, 0.f);
#pragma unroll
                for (int u = 0; u < 4; ++u) {
                    const float w = (s0 + u < S) ? wtab[bh * S + s0 + u] : 0.f;
                    O.x += v[u].x * w; O.y += v[u].y * w; O.z += v[u].z * w; O.w += v[u].w * w;
                }
            }
            const float inv = inv_tab[bh];
            __half2* dst = reinterpret_cast<__half2*>(xs + (size_t)idx * 4);
            dst[0] = __floats2half2_rn(O.x * inv, O.y * inv);
            dst[1] = __floats2half2_rn(O.z * inv, O.w * inv);
        }
        consumers_sync();
        return;
    }

    // gamma packets of this thread, fetched while x is still in flight
    constexpr int GP = 2;
    Pack16 gpk[GP];
    const __half* gm = reinterpret_cast<const __half*>(op.gamma);
    if (op.norm) {
#pragma unroll
        for (int j = 0; j < GP; ++j) {
            const int i = tid + j * kStepConsumers;
            if (i < pk) gpk[j] = ld16(gm + (size_t)i * 8);
        }
    }

#pragma unroll
    for (int b = 0; b < NB; ++b) {
        if (b >= p.nb) break;
        const __half* xg;
        if (op.xsrc == kXEmbed) {
            int id = p.ids_cur[b];
            id = id < 0 ? 0 : (id >= p.vocab ? p.vocab - 1 : id);
            xg = reinterpret_cast<const __half*>(p.embed) + (size_t)id * K;
        } else {
            // with a single split the attention op has already written the merged fp16 output
            xg = reinterpret_cast<const __half*>(op.xsrc == kXCombine ? p.attn : op.x) + (size_t)b * K;
        }
        if (op.xsrc == kXResidAcc) {
            // the residual add of the K-split projection: T(x + T(sum)), the rounding points of "linear output stored
            // in T, then launchAddResidual" (add_residual.cu:58); CTA 0 publishes the new residual stream to xout
            const float4* ag = reinterpret_cast<const float4*>(op.acc_in) + (size_t)b * (K / 4);
            for (int i0 = tid; i0 < pk; i0 += 2 * kStepConsumers) {
                const int i1 = i0 + kStepConsumers;
                const Pack16 va = ld16_cg_nv(xg + (size_t)i0 * 8);
                const float4 a0 = __ldcg(ag + 2 * i0), a1 = __ldcg(ag + 2 * i0 + 1);
                Pack16 vb = va;
                float4 b0 = a0, b1 = a1;
                if (i1 < pk) {
                    vb = ld16_cg_nv(xg + (size_t)i1 * 8);
                    b0 = __ldcg(ag + 2 * i1);
                    b1 = __ldcg(ag + 2 * i1 + 1);
                }
#pragma unroll
                for (int u = 0; u < 2; ++u) {
                    const int i = u ? i1 : i0;
                    if (i >= pk) break;
                    float f[8];
                    Unpack<__half>::get(u ? vb : va, f);
                    const float4 s0 = u ? b0 : a0, s1 = u ? b1 : a1;
                    f[0] += __half2float(__float2half_rn(s0.x)); f[1] += __half2float(__float2half_rn(s0.y));
                    f[2] += __half2float(__float2half_rn(s0.z)); f[3] += __half2float(__float2half_rn(s0.w));
                    f[4] += __half2float(__float2half_rn(s1.x)); f[5] += __half2float(__float2half_rn(s1.y));
                    f[6] += __half2float(__float2half_rn(s1.z)); f[7] += __half2float(__float2half_rn(s1.w));
                    const Pack16 v = Unpack<__half>::put(f);
                    st16(xs + (size_t)b * K + (size_t)i * 8, v);
                    if (blockIdx.x == 0) st16(reinterpret_cast<__half*>(op.xout) + (size_t)b * K + (size_t)i * 8, v);
                    if (op.norm) {
                        Unpack<__half>::get(v, f);   // the norm sees the rounded sum
#pragma unroll
                        for (int e = 0; e < 8; ++e) ss[b] = fmaf(f[e], f[e], ss[b]);
                    }
                }
            }
            continue;
        }
        for (int i0 = tid; i0 < pk; i0 += 4 * kStepConsumers) {
            Pack16 v4[4];
#pragma unroll
            for (int u = 0; u < 4; ++u) {   // four independent L2 loads in flight per thread
                const int i = i0 + u * kStepConsumers;
                if (i < pk) v4[u] = ld16_cg_nv(xg + (size_t)i * 8);
            }
#pragma unroll
            for (int u = 0; u < 4; ++u) {
                const int i = i0 + u * kStepConsumers;
                if (i >= pk) break;
                const Pack16 v = v4[u];
                st16(xs + (size_t)b * K + (size_t)i * 8, v);
                if (op.xsrc == kXEmbed && blockIdx.x == 0)   // the residual stream starts as the embedding
                    st16(reinterpret_cast<__half*>(op.xout) + (size_t)b * K + (size_t)i * 8, v);
                if (op.norm) {
                    float f[8];
                    Unpack<__half>::get(v, f);
#pragma unroll
                    for (int e = 0; e < 8; ++e) ss[b] = fmaf(f[e], f[e], ss[b]);
                }
            }
        }
    }
    if (op.norm) {
#pragma unroll
        for (int b = 0; b < NB; ++b) {
            const float s = warp_sum(ss[b]);
            if (lane == 0) red[b * kStepWarps + warp] = s;
        }
        consumers_sync();
        if (tid < NB) {
            float s = 0.f;
            for (int w = 0; w < kStepWarps; ++w) s += red[tid * kStepWarps + w];
            red[NB * kStepWarps + tid] = rsqrtf(s / (float)K + p.eps);
        }
        consumers_sync();
        const bool gfirst = op.norm == 2;
        for (int i = tid, j = 0; i < pk; i += kStepConsumers, ++j) {
            float g[8];
            Unpack<__half>::get(j == 0 ? gpk[0] : (j == 1 ? gpk[1] : ld16(gm + (size_t)i * 8)), g);
#pragma unroll
            for (int b = 0; b < NB; ++b) {
                if (b >= p.nb) break;
                const float inv = red[NB * kStepWarps + b];
                float f[8];
                Unpack<__half>::get(ld16(xs + (size_t)b * K + (size_t)i * 8), f);
#pragma unroll
                for (int e = 0; e < 8; ++e) f[e] = gfirst ? (g[e] * f[e]) * inv : (f[e] * inv) * g[e];
                st16(xs + (size_t)b * K + (size_t)i * 8, Unpack<__half>::put(f));
            }
        }
    }
    consumers_sync();
}

// ------------------------------------------------------------------------------------------
// consumers: second phase of kEpiSwigluDown. The CTA holds act[u0, u1) (the SwiGLU outputs it has just computed) in
// shared memory and streams rows [u0, u1) of W2 = down^T [I, H]: y[0, H) += act[j] * W2[j][0, H) - the down projection
// (ffn.cpp:88) split along K over the CTAs. A ring stage is kStageB / (2K) whole rows; thread t owns 16-byte packets
// t, t + 256, .. of every row (rows narrower than 256 packets: several row groups in parallel), so there is no
// cross-lane reduction at all; the CTA's partial [H] goes to the fp32 accumulator with red.global.add.v4.f32 and
// the consumer of the sum (next op's x staging) rounds it to T before the residual add.
// ------------------------------------------------------------------------------------------
__device__ __forceinline__ void red_add_v4(float* gptr, float a, float b, float c, float d) {
    asm volatile("red.global.add.v4.f32 [%0], {%1,%2,%3,%4};" ::"l"(gptr), "f"(a), "f"(b), "f"(c), "f"(d) : "memory");
}

// y[b][0, K) += sum_r xv[b][r] * row_r[0, K) over the n_rows rows (K halves each) this CTA streams through the ring;
// xv = [nb][x_stride] halves in shared memory, out = fp32 [nb][out_stride] in global memory (red.add)
template <int NB, int PPT>
__device__ __forceinline__ uint32_t down_consume(const StepParams& p, uint32_t state, const int K, const int n_rows,
                                                 const __half* const act_s, const int act_stride, float* const out,
                                                 const int out_stride, unsigned long long* const stamp = nullptr) {
    extern __shared__ __align__(128) unsigned char smem[];
    const StepSmem<NB> sm(smem, p.stages, p.xs_bytes);
    unsigned char* const r_base = sm.ring;
    uint64_t* const r_full = sm.full;
    uint64_t* const r_empty = sm.empty;
    const int r_stages = p.stages;
#ifdef B200_STEP_STARVE_TRACE
    long long* const starved = sm.starved;
    const bool traced = p.trace != nullptr && threadIdx.x < 32 && (blockIdx.x == 0 || blockIdx.x == gridDim.x - 1);
#endif
    int slot = (int)(state & 0xffu);
    uint32_t phase = state >> 8;
    const int nb = p.nb;
    const int PR = K / 8;                                              // packets per row
    const int tid = threadIdx.x, lane = tid & 31;
    const int PRc = PR < kStepConsumers ? PR : kStepConsumers;
    const int rg = tid / PRc, nrg = kStepConsumers / PRc;              // row groups working in parallel on a stage
    const int p0 = tid - rg * PRc;
    const int rows_per_stage = kStageB / (K * 2);
    const int n2 = (n_rows + rows_per_stage - 1) / rows_per_stage;
    float acc[NB][PPT][8];
#pragma unroll
    for (int b = 0; b < NB; ++b)
#pragma unroll
        for (int i = 0; i < PPT; ++i)
#pragma unroll
            for (int e = 0; e < 8; ++e) acc[b][i][e] = 0.f;
#ifdef B200_DBG_STAGES
    long long dbg_t[16], dbg_c[16];
    const long long dbg_entry = clock64();
#endif

    for (int s2 = 0; s2 < n2; ++s2) {
#ifdef B200_STEP_STARVE_TRACE
        if (traced) {
            const long long c0 = clock64();
            mbar_wait(&r_full[slot], phase);
            if (lane == 0) *starved += clock64() - c0;
        } else
#endif
            mbar_wait(&r_full[slot], phase);
#ifdef B200_DBG_STAGES
        if (stamp && s2 < 16) dbg_t[s2] = clock64();
#endif
        const int r0 = s2 * rows_per_stage;
        const int r1 = min(n_rows, r0 + rows_per_stage);
        const unsigned char* st = r_base + (size_t)slot * kStageB;
        // four rows per trip: all their shared-memory loads are issued before the first FMA (8 warps per SM cannot hide
        // the load latency of a one-row-at-a-time loop: 150 cycles per row measured)
        constexpr int UR = PPT == 1 ? 4 : (PPT == 2 ? 2 : 1);
        for (int r = r0 + rg; r < r1; r += UR * nrg) {
            uint16_t a[UR][NB];
            Pack16 w[UR][PPT];
#pragma unroll
            for (int u = 0; u < UR; ++u) {
                const int rr = min(r + u * nrg, r1 - 1);   // rows past the end re-read the last row; their FMAs are skipped
#pragma unroll
                for (int b = 0; b < NB; ++b) a[u][b] = __half_as_ushort(act_s[(b < nb ? b : 0) * act_stride + rr]);
                const unsigned char* row = st + (size_t)(rr - r0) * K * 2;
#pragma unroll
                for (int i = 0; i < PPT; ++i) {
                    const int pk = min(p0 + i * kStepConsumers, PR - 1);
                    w[u][i] = ld16(row + (size_t)pk * 16);
                }
            }
#pragma unroll
            for (int u = 0; u < UR; ++u) {
                if (r + u * nrg < r1) {
#pragma unroll
                    for (int i = 0; i < PPT; ++i) {
                        if (p0 + i * kStepConsumers < PR) {
#pragma unroll
                            for (int b = 0; b < NB; ++b) {
#pragma unroll
                                for (int e = 0; e < 4; ++e) {
                                    acc[b][i][2 * e] = fma_h((uint16_t)(w[u][i].w[e] & 0xffffu), a[u][b], acc[b][i][2 * e]);
                                    acc[b][i][2 * e + 1] = fma_h((uint16_t)(w[u][i].w[e] >> 16), a[u][b], acc[b][i][2 * e + 1]);
                                }
                            }
                        }
                    }
                }
            }
        }
#ifdef B200_DBG_STAGES
        if (stamp && s2 < 16) dbg_c[s2] = clock64();
#endif
        __syncwarp();
        if (lane == 0) mbar_arrive(&r_empty[slot]);
        if (++slot == r_stages) {
            slot = 0;
            phase ^= 1;
        }
    }
    if (stamp) *stamp = global_timer_ns();   // trace: all stages consumed
#ifdef B200_DBG_STAGES
    if (stamp) {
        printf("entry->first data %d, total %d; stages (cycles: compute, then gap to the next stage's data):", (int)(dbg_t[0] - dbg_entry), (int)(clock64() - dbg_entry));
        for (int i = 0; i < n2 && i < 16; ++i) printf(" %d/%d", (int)(dbg_c[i] - dbg_t[i]), i + 1 < n2 ? (int)(dbg_t[i + 1] - dbg_c[i]) : 0);
        printf("\n");
    }
#endif
    // the CTA's partial sums -> the fp32 accumulator (L2 adds; order over the CTAs is not fixed)
    if (rg < n_rows) {
#pragma unroll
        for (int b = 0; b < NB; ++b) {
            if (b < nb) {
#pragma unroll
                for (int i = 0; i < PPT; ++i) {
                    const int pk = p0 + i * kStepConsumers;
                    if (pk < PR) {
                        float* dst = out + (size_t)b * out_stride + (size_t)pk * 8;
                        red_add_v4(dst, acc[b][i][0], acc[b][i][1], acc[b][i][2], acc[b][i][3]);
                        red_add_v4(dst + 4, acc[b][i][4], acc[b][i][5], acc[b][i][6], acc[b][i][7]);
                    }
                }
            }
        }
    }
    return (uint32_t)slot | (phase << 8);
}

// ------------------------------------------------------------------------------------------
// consumers: one GEMV op. Warp w owns row pair w of every ring stage; two rows share each x packet.
// Epilogues are those of gemv.cu (same rounding points as the unfused launch sequence).
// ------------------------------------------------------------------------------------------
template <int NB>
__device__ __noinline__ uint32_t gemv_consume(const StepOp& op, const StepParams& p, uint32_t state, const int step,
                                              const int op_index) {
    extern __shared__ __align__(128) unsigned char smem_dyn[];
    const StepSmem<NB> lay(smem_dyn, p.stages, p.xs_bytes);
    unsigned char* const r_base = lay.ring;
    uint64_t* const r_full = lay.full;
    uint64_t* const r_empty = lay.empty;
    const int r_stages = p.stages;
    const __half* const xs = lay.xs;
    float* const scratch = lay.scratch;
    constexpr int KSEG = kSegB / 2;
#ifdef B200_STEP_STARVE_TRACE   // build with -DB200_STEP_STARVE_TRACE: cycles warp 0 waited for weights, per op, into the trace
    long long* const starved = reinterpret_cast<long long*>(
        (reinterpret_cast<uintptr_t>(scratch) + kStepScratchBytes - 16) & ~(uintptr_t)7);
#endif
    // everything the loops need sits in registers: with ~220 KB of shared memory in use there is
    // no L1 left, so any local-memory access (spill, indexed array) would be an L2 round trip
    int slot = (int)(state & 0xffu);
    uint32_t phase = state >> 8;
    const int epi = op.epi;
    const int nb = p.nb, N = op.N, K = op.K, units = op.units, head_size = p.head_size;
    // opaque copies: nothing derived from the thread / block indices may be hoisted out of the kernel's op loop, where it
    // would stay live across every out-of-line call (see decode_step_kernel)
    const int tid_o = opaque_i32((int)threadIdx.x), bid_o = opaque_i32((int)blockIdx.x), grid_o = opaque_i32((int)gridDim.x);
    const int warp = tid_o >> 5, lane = tid_o & 31;
    const int nseg = (K + KSEG - 1) / KSEG;
    const int last_seg_pk = (K - (nseg - 1) * KSEG) / 8;
    const int u0 = (int)(((long long)units * bid_o) / grid_o);
    const int u1 = (int)(((long long)units * (bid_o + 1)) / grid_o);
    const int ngroups = (u1 - u0 + kStepWarps - 1) / kStepWarps;
    const int gsz = group_size(u1 - u0, ngroups);
    __half* const y = reinterpret_cast<__half*>(op.y);
#ifdef B200_STEP_STARVE_TRACE
    const bool traced = p.trace != nullptr && warp == 0 && (blockIdx.x == 0 || blockIdx.x == gridDim.x - 1);
    if (traced && lane == 0) *starved = 0;
#endif

    if (epi == kEpiRopeKV) {
        // cos / sin of this step's rotary angles, once per CTA into shared memory (qkv_bias_and_RoPE.cu:29-46): the
        // epilogue used to fetch rope_den[i] from global memory on lane 0 of every row pair - an L2 round trip with
        // the warp's next stage waiting behind it
        for (int i = tid_o; i < (p.rotary_dim >> 1); i += kStepConsumers) {
            const float ang = (float)(step - 1) / p.rope_den[i];
            scratch[2 * i] = cosf(ang);
            scratch[2 * i + 1] = sinf(ang);
        }
        consumers_sync();
    }
    for (int g = 0; g < ngroups; ++g) {
        const int unit = u0 + g * gsz + warp;
        const bool active = warp < gsz && unit < u1;
        int ra = 0, rb = -1;
        if (active) {
            if (epi == kEpiSwiglu || epi == kEpiSwigluDown) { ra = unit; rb = N + unit; }
            else if (epi == kEpiRopeKV) {
                const int hd = head_size >> 1;
                const int head = unit / hd, i = unit - head * hd;
                ra = head * head_size + i;
                rb = ra + hd;
            } else { ra = 2 * unit; rb = (2 * unit + 1 < N) ? 2 * unit + 1 : -1; }
        }
        // residual epilogue: fetch the old y values now, their L2 latency hides behind the row
        float ya[NB], yb[NB];
#pragma unroll
        for (int b = 0; b < NB; ++b) {
            ya[b] = yb[b] = 0.f;
            if (epi == kEpiResidual && active && lane == 0 && b < nb) {
                ya[b] = __half2float(__ldcg(y + (size_t)b * N + ra));
                if (rb >= 0) yb[b] = __half2float(__ldcg(y + (size_t)b * N + rb));
            }
        }
        float acc_a[NB], acc_b[NB];
#pragma unroll
        for (int b = 0; b < NB; ++b) acc_a[b] = acc_b[b] = 0.f;

        for (int seg = 0; seg < nseg; ++seg) {
#ifdef B200_STEP_STARVE_TRACE
            if (traced) {   // cycles this warp spends starved for weights
                const long long c0 = clock64();
                mbar_wait(&r_full[slot], phase);
                if (lane == 0) *starved += clock64() - c0;
            } else
#endif
                mbar_wait(&r_full[slot], phase);
            if (active) {
                const int seg_pk = (seg == nseg - 1) ? last_seg_pk : KSEG / 8;
                const unsigned char* sa = r_base + (size_t)slot * kStageB + (2 * warp) * kSegB;
                const unsigned char* sb = sa + kSegB;   // row b's slot (stale data when rb < 0: result unused)
                const __half* xseg = xs + (size_t)seg * KSEG;
                if (seg_pk == KSEG / 8) {
                    // whole segment (all but a row's last one): branch-free, every weight packet is loaded before the
                    // first FMA; same summation order as the guarded loop below
                    constexpr int CH = 4;   // packets per lane loaded before their FMAs (register budget)
#pragma unroll
                    for (int j0 = 0; j0 < kJS; j0 += CH) {
                        Pack16 wa[CH], wb[CH];
#pragma unroll
                        for (int j = 0; j < CH; ++j) {
                            wa[j] = ld16(sa + (lane + 32 * (j0 + j)) * 16);
                            wb[j] = ld16(sb + (lane + 32 * (j0 + j)) * 16);
                        }
#pragma unroll
                        for (int b = 0; b < NB; ++b) {
#pragma unroll
                            for (int j = 0; j < CH; ++j) {
                                const Pack16 xv = ld16(xseg + (size_t)b * K + (size_t)(lane + 32 * (j0 + j)) * 8);
                                acc_a[b] = dot16<__half>(wa[j], xv, acc_a[b]);
                                acc_b[b] = dot16<__half>(wb[j], xv, acc_b[b]);
                            }
                        }
                    }
                } else {
#pragma unroll
                    for (int j = 0; j < kJS; ++j) {
                        const int pkt = lane + 32 * j;
                        if (pkt < seg_pk) {
                            const Pack16 wa = ld16(sa + pkt * 16);
                            const Pack16 wb = ld16(sb + pkt * 16);
#pragma unroll
                            for (int b = 0; b < NB; ++b) {
                                const Pack16 xv = ld16(xseg + (size_t)b * K + (size_t)pkt * 8);
                                acc_a[b] = dot16<__half>(wa, xv, acc_a[b]);
                                acc_b[b] = dot16<__half>(wb, xv, acc_b[b]);
                            }
                        }
                    }
                }
            }
            __syncwarp();
            if (lane == 0) mbar_arrive(&r_empty[slot]);
            if (++slot == r_stages) {
                slot = 0;
                phase ^= 1;
            }
        }
        if (active) {
#pragma unroll
            for (int b = 0; b < NB; ++b) {
                acc_a[b] = warp_sum(acc_a[b]);
                acc_b[b] = warp_sum(acc_b[b]);
            }
            if (lane == 0) {
                auto rnd = [](float v) { return __half2float(__float2half_rn(v)); };
                if (epi == kEpiSwiglu || epi == kEpiSwigluDown) {
#pragma unroll
                    for (int b = 0; b < NB; ++b) {
                        if (b < nb) {
                            const float gv = rnd(acc_a[b]), uv = rnd(acc_b[b]);   // gate/up are stored in T unfused
                            const float sg = rnd(gv / (1.f + expf(-gv)));
                            const __half av = __float2half_rn(sg * uv);
                            if (epi == kEpiSwiglu) y[(size_t)b * N + ra] = av;
                            else   // stays in this CTA ([nb][act_stride] in `scratch`): it multiplies its own W2 rows
                                reinterpret_cast<__half*>(scratch)[b * op.act_stride + (unit - u0)] = av;
                        }
                    }
                } else if (epi == kEpiRopeKV) {
                    const int d = head_size, hd = d >> 1;
                    const int head = ra / d, i = ra - head * d;
                    const int qkv_heads = p.head_num + 2 * p.kv_head_num;
                    float c = 1.f, sn = 0.f;
                    if (head < p.head_num + p.kv_head_num && i < (p.rotary_dim >> 1)) {
                        c = scratch[2 * i];   // staged once per op, below (scratch is only 4-byte aligned)
                        sn = scratch[2 * i + 1];
                    }
#pragma unroll
                    for (int b = 0; b < NB; ++b) {
                        if (b < nb) {
                            const float av = rnd(acc_a[b]), bv = rnd(acc_b[b]);   // qkv_buf is stored in T before RoPE
                            const __half oa = __float2half_rn(c * av - sn * bv);
                            const __half ob = __float2half_rn(c * bv + sn * av);
                            __half* yt = y + (size_t)b * qkv_heads * d;
                            yt[ra] = oa;
                            yt[rb] = ob;
                            if (head >= p.head_num) {
                                const bool is_k = head < p.head_num + p.kv_head_num;
                                const int kvh = is_k ? head - p.head_num : head - p.head_num - p.kv_head_num;
                                __half* cache = reinterpret_cast<__half*>(is_k ? op.k_cache : op.v_cache);
                                __half* row = cache + (((size_t)b * p.kv_head_num + kvh) * p.max_seq_len +
                                                       (size_t)(step - 1)) * d;
                                row[i] = oa;
                                row[i + hd] = ob;
                            }
                        }
                    }
                } else {
#pragma unroll
                    for (int b = 0; b < NB; ++b) {
                        if (b < nb) {
                            float av = acc_a[b], bv = acc_b[b];
                            if (epi == kEpiResidual) {   // linear output rounded to T, then T(out + residual)
                                av = rnd(av) + ya[b];
                                bv = rnd(bv) + yb[b];
                            }
                            y[(size_t)b * N + ra] = __float2half_rn(av);
                            if (rb >= 0) y[(size_t)b * N + rb] = __float2half_rn(bv);
                        }
                    }
                }
            }
        }
    }
#ifdef B200_STEP_STARVE_TRACE
    if (traced && lane == 0) p.trace[4 * p.n_ops + 1 + (blockIdx.x == 0 ? 0 : p.n_ops) + op_index] = (unsigned long long)*starved;
#endif
    if (epi == kEpiRopeKV && op.head_cnt != nullptr) {
        // this CTA's rows of q / k / v (and the cache rows appended with them) are complete: credit every qkv head
        // it covers with the number of row pairs it contributed; attention items wait for head_size / 2 per head
        consumers_sync();
        if (threadIdx.x == 0 && u1 > u0) {
            const int hd = head_size >> 1;
            for (int head = u0 / hd; head <= (u1 - 1) / hd; ++head) {
                const int n = min(u1, (head + 1) * hd) - max(u0, head * hd);
                asm volatile("red.release.gpu.global.add.u32 [%0], %1;" ::"l"(op.head_cnt + head), "r"((unsigned)n) : "memory");
            }
        }
    }
    return (uint32_t)slot | (phase << 8);
}

// second phase of a kEpiSwigluDown op: ONE out-of-line entry (own register allocation; called from the kernel loop
// where little is live), the packets-per-thread variants are chosen inside
template <int NB>
__device__ __noinline__ uint32_t down_phase(const StepOp& op, const StepParams& p, uint32_t state, const int op_index) {
    consumers_sync();   // the whole act slice is in shared memory
    const int units = op.units;
    const int u0 = (int)(((long long)units * blockIdx.x) / gridDim.x);
    const int u1 = (int)(((long long)units * (blockIdx.x + 1)) / gridDim.x);
    const int ppt = (op.K / 8 + kStepConsumers - 1) / kStepConsumers;   // 16-byte packets of a W2 row per thread
    extern __shared__ __align__(128) unsigned char smem_dyn[];
    const StepSmem<NB> lay(smem_dyn, p.stages, p.xs_bytes);
    const __half* act_s = reinterpret_cast<const __half*>(lay.scratch);
    if (ppt <= 1) state = down_consume<NB, 1>(p, state, op.K, u1 - u0, act_s, op.act_stride, op.acc_out, op.K);
    else if (ppt == 2) state = down_consume<NB, 2>(p, state, op.K, u1 - u0, act_s, op.act_stride, op.acc_out, op.K);
    else state = down_consume<NB, (NB <= 2 ? 4 : 2)>(p, state, op.K, u1 - u0, act_s, op.act_stride, op.acc_out, op.K);   // K = 8192: batch <= 2 (builder)
#ifdef B200_STEP_STARVE_TRACE
    if (p.trace != nullptr && threadIdx.x == 0 && (blockIdx.x == 0 || blockIdx.x == gridDim.x - 1)) {
        extern __shared__ __align__(128) unsigned char smem[];
        const StepSmem<NB> sm(smem, p.stages, p.xs_bytes);
        p.trace[4 * p.n_ops + 1 + (blockIdx.x == 0 ? 0 : p.n_ops) + op_index] = (unsigned long long)*sm.starved;
    }
#endif
    return state;
}

// ------------------------------------------------------------------------------------------
// consumers: split-KV decode attention (decode_attention.cu's kernel restated for 8 warps inside
// the persistent kernel). Work item = (sequence, kv head, split); head_size 128 (16 packets / row).
// The new token's k,v were written to cache row step-1 by the qkv op's epilogue.
// ------------------------------------------------------------------------------------------
#ifndef B200_STEP_ATTN_GC
#define B200_STEP_ATTN_GC 2   // query heads of one kv head that a work item handles at once (see attn_consume)
#endif
__device__ __forceinline__ float exp_guard_s(float m, float M) { return (m == -INFINITY) ? 0.f : expf(m - M); }

// L2 prefetch of the cache rows [k0, min(k1, step-1)) of this CTA's first attention item (one
// prefetch per 128-byte line); the newest row is written by the qkv op of this step and is not touched
template <int G>
__device__ __forceinline__ void attn_prefetch(const StepOp& op, const StepParams& p, int step) {
    constexpr int D = 128;
    constexpr int NP = G > B200_STEP_ATTN_GC ? G / B200_STEP_ATTN_GC : 1;
    const int splits = p.splits;
    const int items = p.nb * p.kv_head_num * splits * NP;
    if ((int)blockIdx.x >= items) return;
    const int item = blockIdx.x / NP;   // (sequence, kv head, split) of this CTA's first work item
    const int per = (step + splits - 1) / splits;
    const int split = item % splits, bh = item / splits;
    const int b = bh / p.kv_head_num, kvh = bh - b * p.kv_head_num;
    const __half* kc = reinterpret_cast<const __half*>(op.k_cache) + ((size_t)b * p.kv_head_num + kvh) * p.max_seq_len * D;
    const __half* vc = reinterpret_cast<const __half*>(op.v_cache) + ((size_t)b * p.kv_head_num + kvh) * p.max_seq_len * D;
    const int k0 = split * per, kend = min(min(step, k0 + per), step - 1);
    const int lines = (kend - k0) * (D * 2 / 128);
    for (int i = threadIdx.x; i < 2 * lines; i += kStepConsumers) {
        const __half* base = (i < lines ? kc : vc) + (size_t)k0 * D;
        const int li = i < lines ? i : i - lines;
        asm volatile("prefetch.global.L2 [%0];" ::"l"(base + (size_t)li * 64));
    }
}

template <int G>
__device__ __noinline__ void attn_consume(const StepOp& op, const StepParams& p, float* sm, int step) {
    constexpr int PR = 16, E = 8, D = 128, KPI = 2, UN = 4;
    constexpr int GC = G > B200_STEP_ATTN_GC ? B200_STEP_ATTN_GC : G;   // heads per work item
    constexpr int NP = G / GC;                                            // head groups per (sequence, kv head, split)
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const int pi = lane % PR, ks = lane / PR;
    const int splits = p.splits;
    const int items = p.nb * p.kv_head_num * splits * NP;
    const int qkv_heads = p.head_num + 2 * p.kv_head_num;
    const float scale = rsqrtf((float)D);
    const int per = (step + splits - 1) / splits;

    for (int item = blockIdx.x; item < items; item += gridDim.x) {
        const int g0 = (item % NP) * GC, rest = item / NP;
        const int split = rest % splits, bh = rest / splits;
        const int b = bh / p.kv_head_num, kvh = bh - b * p.kv_head_num;
        const __half* tok = reinterpret_cast<const __half*>(p.qkv) + (size_t)b * qkv_heads * D;
        const __half* kc = reinterpret_cast<const __half*>(op.k_cache) + ((size_t)b * p.kv_head_num + kvh) * p.max_seq_len * D;
        const __half* vc = reinterpret_cast<const __half*>(op.v_cache) + ((size_t)b * p.kv_head_num + kvh) * p.max_seq_len * D;
        const int k0 = split * per, k1 = min(step, k0 + per);
        if (op.head_cnt != nullptr) {
            // no grid barrier after the qkv op: wait until the CTAs that own the rows of this item's q heads, k head and
            // v head have released them (D / 2 row pairs per head)
            if (threadIdx.x < GC + 2) {
                const int head = (int)threadIdx.x < GC ? kvh * G + g0 + (int)threadIdx.x
                                                       : ((int)threadIdx.x == GC ? p.head_num + kvh : p.head_num + p.kv_head_num + kvh);
                const unsigned long long t0 = global_timer_ns();
                while (ld_acquire_u32(op.head_cnt + head) < (unsigned)(D / 2)) {
                    if (global_timer_ns() - t0 > 2000000000ull) {
                        atomicExch(&p.sync[2], 3u);
                        __trap();
                    }
                }
            }
            consumers_sync();
        }

        // G = 4 / 8 (8 = the 70B-shaped tensor-parallel ranks): eight heads' q, (m, l) and 64 fp32 outputs per lane do not fit
        // the 168-register cap (G = 8: 472 bytes of stack, 319 local loads / stores in the SASS; G = 4: 136 / 106). A work item
        // therefore covers GC = 2 query heads of its (sequence, kv head, split); the NP = G / GC head groups are separate items
        // that other CTAs take (each re-reads the split's K / V rows, from L2 after the first: a few hundred KB per layer
        // against ~100 MB of weights). Per-head arithmetic and its order are unchanged. Measured on an 8-layer 70B-shaped
        // slice, one GPU, batch 1, context 512 / 4096 (tools/g8_step_bench.py): 2.350 / 2.484 ms per step with all eight heads
        // in one item, 2.310 / 2.443 with passes of four, 2.278 / 2.430 with passes of two inside the item.
        {
            Pack16 q[GC];
#pragma unroll
            for (int g = 0; g < GC; ++g) q[g] = ld16_cg(tok + (size_t)(kvh * G + g0 + g) * D + pi * E);
            float m[GC], l[GC], o[GC][E];
#pragma unroll
            for (int g = 0; g < GC; ++g) {
                m[g] = -INFINITY;
                l[g] = 0.f;
#pragma unroll
                for (int e = 0; e < E; ++e) o[g][e] = 0.f;
            }
            for (int base = k0 + warp * KPI * UN; base < k1; base += kStepWarps * KPI * UN) {
                Pack16 kp[UN], vp[UN];
                bool ok[UN];
#pragma unroll
                for (int u = 0; u < UN; ++u) {
                    const int key = base + u * KPI + ks;
                    ok[u] = key < k1;
                    if (ok[u]) kp[u] = ld16_cg(kc + (size_t)key * D + pi * E);
                }
#pragma unroll
                for (int u = 0; u < UN; ++u) {
                    const int key = base + u * KPI + ks;
                    if (ok[u]) vp[u] = ld16_cg(vc + (size_t)key * D + pi * E);
                }
                float s[GC][UN];
#pragma unroll
                for (int u = 0; u < UN; ++u) {
#pragma unroll
                    for (int g = 0; g < GC; ++g) {
                        float dsum = ok[u] ? dot16<__half>(kp[u], q[g], 0.f) : 0.f;
#pragma unroll
                        for (int off = PR >> 1; off > 0; off >>= 1) dsum += __shfl_xor_sync(0xffffffffu, dsum, off);
                        s[g][u] = ok[u] ? dsum * scale : -INFINITY;
                    }
                }
#pragma unroll
                for (int g = 0; g < GC; ++g) {
                    float mn = m[g];
#pragma unroll
                    for (int u = 0; u < UN; ++u) mn = fmaxf(mn, s[g][u]);
                    if (mn != -INFINITY) {
                        const float corr = exp_guard_s(m[g], mn);
                        l[g] *= corr;
#pragma unroll
                        for (int e = 0; e < E; ++e) o[g][e] *= corr;
                        m[g] = mn;
#pragma unroll
                        for (int u = 0; u < UN; ++u) {
                            if (ok[u]) {
                                const float pr = __expf(s[g][u] - mn);
                                l[g] += pr;
                                float vf[E];
                                Unpack<__half>::get(vp[u], vf);
#pragma unroll
                                for (int e = 0; e < E; ++e) o[g][e] = fmaf(pr, vf[e], o[g][e]);
                            }
                        }
                    }
                }
            }
            // merge the two key lanes-groups of the warp
#pragma unroll
            for (int g = 0; g < GC; ++g) {
                const float mo = __shfl_xor_sync(0xffffffffu, m[g], PR);
                const float lo = __shfl_xor_sync(0xffffffffu, l[g], PR);
                const float M = fmaxf(m[g], mo);
                const float a = exp_guard_s(m[g], M), bb = exp_guard_s(mo, M);
                l[g] = l[g] * a + lo * bb;
#pragma unroll
                for (int e = 0; e < E; ++e) {
                    const float oo = __shfl_xor_sync(0xffffffffu, o[g][e], PR);
                    o[g][e] = o[g][e] * a + oo * bb;
                }
                m[g] = M;
            }
            if (ks == 0) {
#pragma unroll
                for (int g = 0; g < GC; ++g) {
                    float* dst = sm + ((size_t)warp * GC + g) * (2 + D);
                    if (pi == 0) { dst[0] = m[g]; dst[1] = l[g]; }
#pragma unroll
                    for (int e = 0; e < E; ++e) dst[2 + pi * E + e] = o[g][e];
                }
            }
            consumers_sync();
            for (int idx = threadIdx.x; idx < GC * D; idx += kStepConsumers) {
                const int g = idx / D, e = idx - g * D;
                float M = -INFINITY;
#pragma unroll
                for (int w = 0; w < kStepWarps; ++w) M = fmaxf(M, sm[((size_t)w * GC + g) * (2 + D)]);
                float L = 0.f, O = 0.f;
#pragma unroll
                for (int w = 0; w < kStepWarps; ++w) {
                    const float* src = sm + ((size_t)w * GC + g) * (2 + D);
                    const float a = exp_guard_s(src[0], M);
                    L += src[1] * a;
                    O += src[2 + e] * a;
                }
                const int head = kvh * G + g0 + g;
                if (splits == 1) {
                    reinterpret_cast<__half*>(p.attn)[((size_t)b * p.head_num + head) * D + e] =
                        __float2half_rn(O / (L + 1e-6f));
                } else {
                    const size_t slot = ((size_t)b * p.head_num + head) * splits + split;
                    p.part_o[slot * D + e] = O;
                    if (e == 0) { p.part_ml[slot * 2] = M; p.part_ml[slot * 2 + 1] = L; }
                }
            }
        }
        consumers_sync();   // sm is reused by the next item; every thread's partial stores precede thread 0's release
        if (op.type == kOpAttnO && threadIdx.x == 0)
            asm volatile("red.release.gpu.global.add.u32 [%0], 1;" ::"l"(reinterpret_cast<unsigned*>(op.y) + bh) : "memory");
    }
}

// second half of a kOpAttnO op: o-projection item (head, part) of this CTA. Waits for the attention items of its
// (sequence, kv head) pairs, merges the head's split partials (fused_decoder_self_attention.cu:203-221,272-287; the
// attention output is rounded to T as the unfused path stores it before o_proj), then streams its block of the
// packed o weights: rows = the head's d input columns, N / parts output values each.
template <int NB>
__device__ __noinline__ uint32_t attn_o_phase(const StepOp& op, const StepParams& p, uint32_t state, const int op_index) {
    extern __shared__ __align__(128) unsigned char smem[];
    const int item = blockIdx.x;
    if (item >= op.units) return state;   // no o item (and no ring stages) for this CTA
    const StepSmem<NB> sm(smem, p.stages, p.xs_bytes);
    const int parts = op.act_stride, d = p.head_size, nb = p.nb, S = p.splits;
    const int head = item / parts, part = item - head * parts;
    const int kvh = head / (p.head_num / p.kv_head_num);
    const int tid = threadIdx.x;
    const bool tr = p.trace != nullptr && blockIdx.x == 0 && tid == 0;
    if (tr) p.trace[4 * p.n_ops + 1 + op_index] = global_timer_ns();   // attention items of this CTA done
    const int gq = p.head_num / p.kv_head_num;
    const int np = gq > B200_STEP_ATTN_GC ? gq / B200_STEP_ATTN_GC : 1;   // attention work items per (sequence, kv head, split)
    if (tid == 0) {
        const unsigned* cnt = reinterpret_cast<const unsigned*>(op.y);
        const unsigned long long t0 = global_timer_ns();
        for (int b = 0; b < nb; ++b) {
            while (ld_acquire_u32(cnt + b * p.kv_head_num + kvh) < (unsigned)(S * np)) {
                if (global_timer_ns() - t0 > 2000000000ull) {
                    atomicExch(&p.sync[2], 2u);
                    __trap();
                }
            }
        }
    }
    consumers_sync();
    __half* xh = sm.xs;   // [nb][d]: the head's attention output
    for (int idx = tid; idx < nb * d; idx += kStepConsumers) {
        const int b = idx / d, j = idx - b * d;
        const size_t bh = (size_t)b * p.head_num + head;
        __half v;
        if (S == 1) {
            v = ldh_cg(reinterpret_cast<const __half*>(p.attn) + bh * d + j);
        } else {
            // online merge over the splits, four at a time: the eight loads of a group are independent (one L2
            // round trip per group instead of one per load)
            const float2* ml = reinterpret_cast<const float2*>(p.part_ml) + bh * S;
            const float* po = p.part_o + bh * S * d + j;
            float M = -INFINITY, L = 0.f, O = 0.f;
            for (int s0 = 0; s0 < S; s0 += 4) {
                float2 t[4];
                float o4[4];
#pragma unroll
                for (int u = 0; u < 4; ++u) {
                    const bool in = s0 + u < S;
                    t[u] = in ? __ldcg(ml + s0 + u) : make_float2(-INFINITY, 0.f);
                    o4[u] = in ? __ldcg(po + (size_t)(s0 + u) * d) : 0.f;
                }
                float Mn = M;
#pragma unroll
                for (int u = 0; u < 4; ++u) Mn = fmaxf(Mn, t[u].x);
                if (Mn != -INFINITY) {
                    const float c = (M == -INFINITY) ? 0.f : expf(M - Mn);
                    L *= c;
                    O *= c;
                    M = Mn;
#pragma unroll
                    for (int u = 0; u < 4; ++u) {
                        const float a = (t[u].x == -INFINITY) ? 0.f : expf(t[u].x - M);
                        L += t[u].y * a;
                        O += o4[u] * a;
                    }
                }
            }
            v = __float2half_rn(O * (1.f / (L + 1e-6f)));
        }
        xh[b * d + j] = v;
    }
    consumers_sync();
    if (tr) {
        p.trace[5 * p.n_ops + 1 + op_index] = global_timer_ns();   // dependencies met, head output merged
        int sl = (int)(state & 0xffu), landed = 0;                 // ring stages of the o block already landed
        uint32_t ph = state >> 8;
        for (int i = 0; i < p.stages; ++i) {
            landed += mbar_test(&sm.full[sl], ph) ? 1 : 0;
            if (++sl == p.stages) { sl = 0; ph ^= 1; }
        }
        p.trace[2 * p.n_ops + op_index] = (unsigned long long)landed;
    }
    const int row_len = op.N / parts;
    const int ppt = (row_len / 8 + kStepConsumers - 1) / kStepConsumers;
    float* out = op.acc_out + (size_t)part * row_len;
    unsigned long long* stamp = tr ? p.trace + 3 * p.n_ops + op_index : nullptr;
    if (ppt <= 1) state = down_consume<NB, 1>(p, state, row_len, d, xh, d, out, op.N, stamp);
    else state = down_consume<NB, 2>(p, state, row_len, d, xh, d, out, op.N, stamp);
    return state;
}

// kOpExchange: see decode_step.h. One thread per 16-byte line of this CTA's slice; the push of a line and the wait for
// the same line of every rank are done by the same thread, so no intra-CTA synchronisation is needed. fp16 only.
template <int NB>
__device__ __noinline__ void exchange_op(const StepOp& op, const StepParams& p, const unsigned call_base) {
    const int world = p.wire.world, me = p.wire.rank;
    const unsigned call = call_base + (unsigned)op.units + 1u;
    const size_t buf_off = kP2PHeader + (size_t)(call & 1u) * world * p.wire.slot_bytes;
    const int lines = p.nb * op.N / 4;
    const int per = (lines + (int)gridDim.x - 1) / (int)gridDim.x;
    const int l0 = per * blockIdx.x, l1 = min(lines, l0 + per);
    float4* acc4 = reinterpret_cast<float4*>(op.acc_out);
    uint2* hid2 = reinterpret_cast<uint2*>(op.y);
    for (int i = l0 + (int)threadIdx.x; i < l1; i += kStepConsumers) {
        // ---- push: T(partial) of line i into slot `me` of every rank (its own included), then clear the accumulator
        const float4 a = __ldcg(acc4 + i);
        const __half2 h0 = __floats2half2_rn(a.x, a.y), h1 = __floats2half2_rn(a.z, a.w);
        const unsigned d0 = *reinterpret_cast<const unsigned*>(&h0), d1 = *reinterpret_cast<const unsigned*>(&h1);
#pragma unroll 1
        for (int r = 0; r < world; ++r)
            asm volatile("st.volatile.global.v4.u32 [%0], {%1,%2,%3,%4};" ::"l"(p.wire.peer[r] + buf_off + (size_t)me * p.wire.slot_bytes + (size_t)i * 16),
                         "r"(d0), "r"(call), "r"(d1), "r"(call)
                         : "memory");
        acc4[i] = make_float4(0.f, 0.f, 0.f, 0.f);
        // ---- receive: the same line from every rank, summed in rank order (identical on all ranks), + residual
        float s0 = 0.f, s1 = 0.f, s2 = 0.f, s3 = 0.f;
        const unsigned long long t0 = global_timer_ns();
        for (int r = 0; r < world; ++r) {
            const unsigned char* src = p.wire.peer[me] + buf_off + (size_t)r * p.wire.slot_bytes + (size_t)i * 16;
            unsigned w0, w1, w2, w3;
            for (;;) {
                asm volatile("ld.volatile.global.v4.u32 {%0,%1,%2,%3}, [%4];" : "=r"(w0), "=r"(w1), "=r"(w2), "=r"(w3) : "l"(src) : "memory");
                if (w1 == call && w3 == call) break;
                // a peer never arrived: fail instead of hanging. 10 s, not the 2 s of the intra-GPU waits: the other rank is
                // another PROCESS, whose host thread may be late by a scheduling hiccup or a lazy weight re-pack
                if (global_timer_ns() - t0 > 10000000000ull) {
                    atomicExch(&p.sync[2], 4u);
                    __trap();
                }
            }
            const float2 x = __half22float2(*reinterpret_cast<const __half2*>(&w0));
            const float2 y = __half22float2(*reinterpret_cast<const __half2*>(&w2));
            s0 += x.x; s1 += x.y; s2 += y.x; s3 += y.y;
        }
        // hidden <- T(hidden + T(sum)): "allreduce, then launchAddResidual" (oracle/model.py _ffn)
        const uint2 hv = __ldcg(hid2 + i);   // written by another CTA earlier in this launch (embedding row): L2, not L1
        const float2 ha = __half22float2(*reinterpret_cast<const __half2*>(&hv.x));
        const float2 hb = __half22float2(*reinterpret_cast<const __half2*>(&hv.y));
        auto rnd = [](float v) { return __half2float(__float2half_rn(v)); };
        const __half2 o0 = __floats2half2_rn(ha.x + rnd(s0), ha.y + rnd(s1));
        const __half2 o1 = __floats2half2_rn(hb.x + rnd(s2), hb.y + rnd(s3));
        uint2 ov;
        ov.x = *reinterpret_cast<const unsigned*>(&o0);
        ov.y = *reinterpret_cast<const unsigned*>(&o1);
        hid2[i] = ov;
    }
}

// greedy pick (topK.cu:147 with K = 1; ties -> lower index) + advance (ids_cur <- ids_next, step+1)
__device__ __noinline__ void final_consume(const StepParams& p, float* sv, int* si) {
    const int lane = threadIdx.x & 31, wid = threadIdx.x >> 5;
    for (int b = blockIdx.x; b < p.nb; b += gridDim.x) {
        const __half* row = reinterpret_cast<const __half*>(p.logits) + (size_t)b * p.vocab;
        float bv = -INFINITY;
        int bi = 0x7fffffff;
        auto take = [&](float v, int j) {
            if (v > bv || (v == bv && j < bi)) { bv = v; bi = j; }
        };
        const bool vec = (p.vocab & 7) == 0;
        const int npk = vec ? p.vocab / 8 : 0;
        for (int i0 = threadIdx.x; i0 < npk; i0 += 4 * kStepConsumers) {
            Pack16 v4[4];
#pragma unroll
            for (int u = 0; u < 4; ++u) {
                const int i = i0 + u * kStepConsumers;
                if (i < npk) v4[u] = ld16_cg_nv(row + (size_t)i * 8);
            }
#pragma unroll
            for (int u = 0; u < 4; ++u) {
                const int i = i0 + u * kStepConsumers;
                if (i >= npk) break;
                float f[8];
                Unpack<__half>::get(v4[u], f);
#pragma unroll
                for (int e = 0; e < 8; ++e) take(f[e], i * 8 + e);
            }
        }
        for (int j = npk * 8 + threadIdx.x; j < p.vocab; j += kStepConsumers) take(__half2float(ldh_cg(row + j)), j);
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) {
            const float ov = __shfl_xor_sync(0xffffffffu, bv, o);
            const int oi = __shfl_xor_sync(0xffffffffu, bi, o);
            if (ov > bv || (ov == bv && oi < bi)) { bv = ov; bi = oi; }
        }
        if (lane == 0) { sv[wid] = bv; si[wid] = bi; }
        consumers_sync();
        if (threadIdx.x == 0) {
            for (int w = 1; w < kStepWarps; ++w)
                if (sv[w] > bv || (sv[w] == bv && si[w] < bi)) { bv = sv[w]; bi = si[w]; }
            if (bi == 0x7fffffff) bi = 0;
            p.ids_next[b] = bi;
            p.ids_cur[b] = bi;
        }
        consumers_sync();
    }
}

// trace mode: x staged stamp + ring stages already landed when the GEMV of op `oi` starts (CTA 0, thread 0)
__device__ __noinline__ void trace_gemv_start(const StepParams& p, uint32_t ring_state, int oi) {
    extern __shared__ __align__(128) unsigned char smem[];
    uint64_t* full_bar = reinterpret_cast<uint64_t*>(smem + (size_t)p.stages * kStageB + p.xs_bytes);
    p.trace[2 * p.n_ops + oi] = global_timer_ns();
    int sl = (int)(ring_state & 0xffu), landed = 0;
    uint32_t ph = ring_state >> 8;
    for (int i = 0; i < p.stages; ++i) {
        landed += mbar_test(&full_bar[sl], ph) ? 1 : 0;
        if (++sl == p.stages) { sl = 0; ph ^= 1; }
    }
    p.trace[3 * p.n_ops + oi] = (unsigned long long)landed;
}

// ------------------------------------------------------------------------------------------
// the kernel
// ------------------------------------------------------------------------------------------
template <int NB, int G>
__global__ void __launch_bounds__(kStepThreads, 1) decode_step_kernel(const __grid_constant__ StepParams gp) {
    extern __shared__ __align__(128) unsigned char smem[];
    // parameters and the consumers' current op live in shared memory: with ~214 KB of smem in use
    // the L1 is too small to hold per-thread local copies, and every mbarrier wait (a compiler
    // memory barrier) would otherwise re-fetch them from L2
    __shared__ StepParams p;
    __shared__ __align__(16) StepOp opbuf[2];   // current op / next op (fetched while the current one runs)
    unsigned char* ring = smem;
    uint64_t* full_bar = reinterpret_cast<uint64_t*>(smem + (size_t)gp.stages * kStageB + gp.xs_bytes);
    uint64_t* empty_bar = full_bar + kStepMaxStages;

    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    if (threadIdx.x == 0) {
        p = gp;
        for (int s = 0; s < gp.stages; ++s) {
            mbar_init(&full_bar[s], 1);
            mbar_init(&empty_bar[s], kStepWarps);
        }
        mbar_fence_init();
    }
    __syncthreads();

    Ring rg{ring, full_bar, empty_bar, gp.stages, 0, 0};
    if (warp == kStepWarps) {
        step_producer(p, rg, lane);
        return;
    }

    // ---- consumers
#ifdef B200_DBG_STAGES
    const long long dbg_c0 = clock64();
    const unsigned long long dbg_g0 = global_timer_ns();
#endif
    const int step = *gp.step_dev;
    const unsigned base = gp.sync[1];
    // tensor parallel: the communicator's call counter as every CTA of this launch finds it (advanced by the last CTA)
    const unsigned call_base = gp.wire.world > 0 ? *reinterpret_cast<volatile unsigned*>(gp.wire.peer[gp.wire.rank]) : 0u;
    const int n_ops = gp.n_ops;
    unsigned epoch = 0;
    uint32_t ring_state = 0;   // slot | phase << 8
    if (threadIdx.x == 0) opbuf[0] = gp.ops[0];
    consumers_sync();
    for (int oi = 0; oi < n_ops; ++oi) {
        const StepOp& op = opbuf[oi & 1];
        // the next op's descriptor travels while this op runs (its L2 round trip used to sit between
        // the grid barrier and the prologue); the barrier's first sync publishes it to the CTA
        static_assert(sizeof(StepOp) % 16 == 0, "StepOp is copied in 16-byte cp.async pieces");
        if (threadIdx.x < sizeof(StepOp) / 16 && oi + 1 < n_ops)
            asm volatile("cp.async.ca.shared.global [%0], [%1], 16;" ::"r"(smem_u32(&opbuf[(oi + 1) & 1]) + 16u * threadIdx.x),
                         "l"(reinterpret_cast<const char*>(gp.ops + oi + 1) + 16 * threadIdx.x)
                         : "memory");
        asm volatile("cp.async.commit_group;" ::: "memory");
        const int type = op.type;
        if (op.zero) {   // an accumulator consumed one op earlier is cleared for its next use (ops in between: >= 2 barriers)
            const int per = (((op.n_zero + (int)gridDim.x - 1) / (int)gridDim.x) + 3) & ~3;
            const int z0 = per * blockIdx.x, z1 = min(op.n_zero, z0 + per);
            for (int i = z0 + 4 * (int)threadIdx.x; i < z1; i += 4 * kStepConsumers)
                *reinterpret_cast<float4*>(op.zero + i) = make_float4(0.f, 0.f, 0.f, 0.f);
        }
        if (op.zero2) {
            const int per = (((op.n_zero2 + (int)gridDim.x - 1) / (int)gridDim.x) + 3) & ~3;
            const int z0 = per * blockIdx.x, z1 = min(op.n_zero2, z0 + per);
            for (int i = z0 + 4 * (int)threadIdx.x; i < z1; i += 4 * kStepConsumers)
                *reinterpret_cast<float4*>(op.zero2 + i) = make_float4(0.f, 0.f, 0.f, 0.f);
        }
        if (type == kOpGemv) {
            // the shared-memory pointers are re-derived at every call site from an opaque copy of the two layout
            // parameters: kept live across the out-of-line calls they (and whatever else is live) compete with the
            // callees for the 168 registers a 9-warp CTA can have, and anything pushed to local memory is an L2 trip
            {
                const StepSmem<NB> L(smem, opaque_i32(gp.stages), (size_t)opaque_i32((int)gp.xs_bytes));
                stage_x<NB>(op, p, L.xs, L.red, L.scratch);
            }
            if (gp.trace && blockIdx.x == 0 && threadIdx.x == 0) trace_gemv_start(p, ring_state, oi);
            ring_state = gemv_consume<NB>(op, p, ring_state, step, oi);
            if (op.epi == kEpiSwigluDown) ring_state = down_phase<NB>(op, p, ring_state, oi);
        } else if (type == kOpAttn) {
            attn_consume<G>(op, p, reinterpret_cast<float*>(smem + (size_t)opaque_i32(gp.stages) * kStageB), step);
        } else if (type == kOpExchange) {
            exchange_op<NB>(op, p, call_base);
        } else if (type == kOpAttnO) {
            attn_consume<G>(op, p, reinterpret_cast<float*>(smem + (size_t)opaque_i32(gp.stages) * kStageB), step);
            ring_state = attn_o_phase<NB>(op, p, ring_state, oi);
        } else {
            const StepSmem<NB> L(smem, opaque_i32(gp.stages), (size_t)opaque_i32((int)gp.xs_bytes));
            final_consume(p, L.red, reinterpret_cast<int*>(L.red + 16));
        }
        if (gp.trace && blockIdx.x == 0 && threadIdx.x == 0) gp.trace[2 * oi] = global_timer_ns();
        if (gp.trace && threadIdx.x == 0) gp.trace[6 * n_ops + 1 + (size_t)oi * gridDim.x + blockIdx.x] = global_timer_ns();
        if (oi + 1 < n_ops && !op.no_barrier) {
            ++epoch;
            asm volatile("cp.async.wait_group 0;" ::: "memory");   // the next op's descriptor has landed
            grid_barrier_arrive(gp.sync);
            // while the other CTAs arrive: warm L2 with what the next op reads first and that does not
            // depend on this step (the KV rows of this CTA's attention item)
            const StepOp& nop = opbuf[(oi + 1) & 1];
            if (nop.type == kOpAttn || nop.type == kOpAttnO) attn_prefetch<G>(nop, p, step);
            grid_barrier_wait(gp.sync, base + epoch * gridDim.x);
        } else if (oi + 1 < n_ops) {
            // no grid barrier: the next op's consumers wait on the counters this op released. The next descriptor
            // still has to be in shared memory for every thread of the CTA.
            asm volatile("cp.async.wait_group 0;" ::: "memory");
            consumers_sync();
            const StepOp& nop = opbuf[(oi + 1) & 1];
            if (nop.type == kOpAttnO) attn_prefetch<G>(nop, p, step);
        }
        if (gp.trace && blockIdx.x == 0 && threadIdx.x == 0) gp.trace[2 * oi + 1] = global_timer_ns();
    }
#ifdef B200_DBG_STAGES
    if (blockIdx.x == 0 && threadIdx.x == 0)
        printf("kernel: %lld cycles in %llu ns -> %.3f GHz\n", clock64() - dbg_c0, global_timer_ns() - dbg_g0,
               (double)(clock64() - dbg_c0) / (double)(global_timer_ns() - dbg_g0));
#endif
    // last op done by this CTA: the final arrival advances the launch base and the step
    consumers_sync();
    if (threadIdx.x == 0) {
        __threadfence();
        const unsigned total = (epoch + 1) * gridDim.x;
        const unsigned prev = atomicAdd(&gp.sync[0], 1u);
        if (prev + 1 == base + total) {   // every CTA has finished its final op
            gp.sync[1] = base + total;
            *gp.step_dev = step + 1;
            if (gp.wire.world > 0)
                *reinterpret_cast<volatile unsigned*>(gp.wire.peer[gp.wire.rank]) = call_base + (unsigned)gp.n_exchanges;
        }
    }
}

// ------------------------------------------------------------------------------------------
// host side: the program is built once per engine; a launch = one cooperative kernel
// ------------------------------------------------------------------------------------------
static size_t step_smem(int stages, size_t xs_bytes, int nb_t) {
    return (size_t)stages * kStageB + xs_bytes + 2 * kStepMaxStages * sizeof(uint64_t) +
           (size_t)(nb_t * kStepWarps + nb_t + 32) * sizeof(float) + kStepScratchBytes + 64;
}

template <int NB, int G>
static cudaError_t step_launch_t(const StepProgram& sp, cudaStream_t s) {
    auto kern = decode_step_kernel<NB, G>;
    static unsigned long long attr_done = 0;   // per instantiation, one bit per device
    {
        cudaError_t e = ensure_max_dynamic_smem(reinterpret_cast<const void*>(kern), 227 * 1024 - 1024, attr_done);   // static smem holds params + op
        if (e != cudaSuccess) return e;
    }
    cudaLaunchConfig_t cfg{};
    cfg.gridDim = dim3(sp.grid);
    cfg.blockDim = dim3(kStepThreads);
    cfg.dynamicSmemBytes = sp.smem;
    cfg.stream = s;
    cudaLaunchAttribute attr[1];
    attr[0].id = cudaLaunchAttributeCooperative;   // all CTAs co-resident: the grid barrier cannot deadlock
    attr[0].val.cooperative = 1;
    cfg.attrs = attr;
    cfg.numAttrs = 1;
    return cudaLaunchKernelEx(&cfg, kern, sp.params);
}

// one round of CTAs over (sequence, kv head, split) items, >= 64 keys per split
int step_pick_splits(int batch, int kv_head_num, int step) {
    int splits = sm_count() / (batch * kv_head_num);
    const int by_len = (step + 63) / 64;
    if (splits > by_len) splits = by_len;
    if (splits > 64) splits = 64;
    if (splits < 1) splits = 1;
    return splits;
}

int step_program_launch(const StepProgram* sp, int step, cudaStream_t s) {
    StepProgram local = *sp;
    const int splits = step_pick_splits(local.params.nb, local.params.kv_head_num, step);
    local.params.splits = splits;
    const size_t ml = (size_t)local.params.nb * local.params.head_num * splits * 2 * sizeof(float);
    local.params.part_ml = reinterpret_cast<float*>(local.attn_ws);
    local.params.part_o = reinterpret_cast<float*>((char*)local.attn_ws + ((ml + 255) & ~(size_t)255));
    cudaError_t e = cudaErrorInvalidValue;
#define STEP_CASE(NB_, G_) if (local.nb_t == NB_ && local.g == G_) e = step_launch_t<NB_, G_>(local, s);
    STEP_CASE(1, 1) STEP_CASE(2, 1) STEP_CASE(4, 1) STEP_CASE(1, 2) STEP_CASE(2, 2) STEP_CASE(4, 2)
    STEP_CASE(1, 4) STEP_CASE(2, 4) STEP_CASE(4, 4) STEP_CASE(1, 8) STEP_CASE(2, 8) STEP_CASE(4, 8)
#undef STEP_CASE
    count_launch();
    return (int)e;
}

int step_program_trace(const StepProgram* sp, unsigned long long* host_dst, int max_entries) {
    if (!sp || !sp->params.trace) return 0;
    int n = 6 * sp->params.n_ops + 1 + sp->params.n_ops * sp->grid;
    if (n > max_entries) n = max_entries;
    cudaMemcpy(host_dst, sp->params.trace, (size_t)n * sizeof(unsigned long long), cudaMemcpyDeviceToHost);
    return n;
}

void step_program_destroy(StepProgram* sp) {
    if (!sp) return;
    cudaFree(sp->params.trace);
    cudaFree(sp->ops_dev);
    cudaFree(sp->sync_dev);
    delete sp;
}

// returns nullptr when the configuration is outside what the persistent kernel covers
StepProgram* step_program_create(const std::vector<StepOp>& ops, StepParams params, int batch, int head_num,
                                 int kv_head_num, int head_size, int k_max, void* attn_ws) {
    const int G = head_num / kv_head_num;
    if (batch > 4 || head_size != 128 || (G != 1 && G != 2 && G != 4 && G != 8) || head_num % kv_head_num) return nullptr;
    static int disabled = -1;
    if (disabled < 0) { const char* e = getenv("B200_NO_PERSISTENT"); disabled = (e && e[0] == '1') ? 1 : 0; }
    if (disabled) return nullptr;
    const int nb_t = batch <= 1 ? 1 : batch <= 2 ? 2 : 4;
    size_t xs_bytes = ((size_t)nb_t * k_max * 2 + 127) & ~(size_t)127;
    const size_t attn_bytes = ((size_t)kStepWarps * G * (2 + 128) * sizeof(float) + 127) & ~(size_t)127;
    if (xs_bytes < attn_bytes) xs_bytes = attn_bytes;
    int stages = kStepMaxStages;
    while (stages >= 2 && step_smem(stages, xs_bytes, nb_t) > 227 * 1024 - 1024) --stages;   // 1 KB: static smem
    if (stages < 2) return nullptr;
    int dev = 0, coop = 0;
    cudaGetDevice(&dev);
    cudaDeviceGetAttribute(&coop, cudaDevAttrCooperativeLaunch, dev);
    if (!coop) return nullptr;
    StepProgram* sp = new StepProgram();
    sp->nb_t = nb_t;
    sp->g = G;
    sp->grid = sm_count();
    sp->smem = step_smem(stages, xs_bytes, nb_t);
    if (cudaMalloc(&sp->ops_dev, ops.size() * sizeof(StepOp)) != cudaSuccess ||
        cudaMalloc(&sp->sync_dev, 64) != cudaSuccess) {
        step_program_destroy(sp);
        return nullptr;
    }
    cudaMemcpy(sp->ops_dev, ops.data(), ops.size() * sizeof(StepOp), cudaMemcpyHostToDevice);
    cudaMemset(sp->sync_dev, 0, 64);
    sp->params = params;
    sp->attn_ws = attn_ws;
    sp->params.ops = sp->ops_dev;
    sp->params.n_ops = (int)ops.size();
    sp->params.nb = batch;
    sp->params.sync = sp->sync_dev;
    sp->params.stages = stages;
    sp->params.xs_bytes = xs_bytes;
    {
        const char* t = getenv("B200_STEP_TRACE");
        if (t && t[0] == '1') {
            const size_t n = (6 * ops.size() + 1 + ops.size() * (size_t)sp->grid) * sizeof(unsigned long long);
            if (cudaMalloc(&sp->params.trace, n) == cudaSuccess) cudaMemset(sp->params.trace, 0, n);
        }
    }
    return sp;
}

}  // namespace b200
