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
#include <vector>

#include "common.cuh"
#include "decode_step.h"
#include "launch_count.h"

namespace b200 {

constexpr int kStepWarps = 8;                         // consumer warps
constexpr int kStepThreads = (kStepWarps + 1) * 32;   // + producer warp
constexpr int kStepConsumers = kStepWarps * 32;
constexpr int kSegB = 2048;                           // K-segment bytes per row per stage
constexpr int kStageB = kStepWarps * 2 * kSegB;       // 32 KB: 8 row pairs
constexpr int kStepMaxStages = 6;
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
__device__ __forceinline__ void grid_barrier(unsigned* sync, unsigned target) {
    consumers_sync();
    if (threadIdx.x == 0) {
        __threadfence();
        atomicAdd(&sync[0], 1u);
        const unsigned long long t0 = global_timer_ns();
        while ((int)(ld_acquire_u32(&sync[0]) - target) < 0) {
            if (global_timer_ns() - t0 > 2000000000ull) {
                atomicExch(&sync[2], 1u);
                __trap();
            }
        }
        __threadfence();
    }
    consumers_sync();
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
    if (EPI == kEpiSwiglu) {
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
    if (op.epi == kEpiSwiglu) op_unit_rows<kEpiSwiglu>(op, head_size, u, ra, rb);
    else if (op.epi == kEpiRopeKV) op_unit_rows<kEpiRopeKV>(op, head_size, u, ra, rb);
    else op_unit_rows<kEpiPlain>(op, head_size, u, ra, rb);
}

__device__ __noinline__ void step_producer(const StepParams& p, Ring rg, int lane) {
    bool wrapped = false;
    for (int oi = 0; oi < p.n_ops; ++oi) {
        const StepOp& op = p.ops[oi];
        if (op.type != kOpGemv) continue;
        const int KSEG = kSegB / 2;
        const int nseg = (op.K + KSEG - 1) / KSEG;
        const unsigned char* Wb = reinterpret_cast<const unsigned char*>(op.W);
        const size_t row_bytes = (size_t)op.K * 2;
        const uint32_t last_seg_bytes = (uint32_t)(op.K - (nseg - 1) * KSEG) * 2u;
        const int u0 = (int)(((long long)op.units * blockIdx.x) / gridDim.x);
        const int u1 = (int)(((long long)op.units * (blockIdx.x + 1)) / gridDim.x);
        const int ngroups = (u1 - u0 + kStepWarps - 1) / kStepWarps;
        for (int g = 0; g < ngroups; ++g) {
            const int ubase = u0 + g * kStepWarps;
            const int nu = min(kStepWarps, u1 - ubase);
            const unsigned char* src = nullptr;
            if (lane < 2 * nu) {
                int ra, rb;
                op_rows_rt(op, p.head_size, ubase + (lane >> 1), ra, rb);
                const int r = (lane & 1) ? rb : ra;
                if (r >= 0) src = Wb + (size_t)r * row_bytes;
            }
            const uint32_t rows_here = __popc(__ballot_sync(0xffffffffu, src != nullptr));
            unsigned char* dst_lane = rg.base + (size_t)lane * kSegB;
            for (int seg = 0; seg < nseg; ++seg) {
                if (wrapped) mbar_wait(&rg.empty[rg.slot], rg.phase);
                const uint32_t seg_bytes = (seg == nseg - 1) ? last_seg_bytes : (uint32_t)kSegB;
                if (lane == 0) mbar_expect_tx(&rg.full[rg.slot], seg_bytes * rows_here);
                if (src)
                    bulk_g2s(dst_lane + (size_t)rg.slot * kStageB, src + (size_t)seg * kSegB, seg_bytes,
                             &rg.full[rg.slot]);
                if (++rg.slot == rg.stages) {
                    rg.slot = 0;
                    if (wrapped) rg.phase ^= 1;
                    wrapped = true;
                }
            }
        }
    }
}

// ------------------------------------------------------------------------------------------
// consumers: x staging (+ RMSNorm / embedding / split-KV combine prologues)
// ------------------------------------------------------------------------------------------
template <int NB>
__device__ __noinline__ void stage_x(const StepOp& op, const StepParams& p, __half* xs, float* red) {
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const int K = op.K, pk = K / 8;
    float ss[NB];
#pragma unroll
    for (int b = 0; b < NB; ++b) ss[b] = 0.f;

    if (op.xsrc == kXCombine && p.splits > 1) {
        // x = softmax-merged split-KV partials, rounded to fp16 (the unfused path stores the
        // attention output in T before o_proj): fused_decoder_self_attention.cu:203-221,272-287
        const int d = p.head_size, S = p.splits;
        const int total = p.nb * p.head_num * d;   // == nb * K
        for (int idx = tid; idx < total; idx += kStepConsumers) {
            const int bh = idx / d, e = idx - bh * d;
            const float* ml = p.part_ml + (size_t)bh * S * 2;
            float M = -INFINITY;
            for (int s = 0; s < S; ++s) M = fmaxf(M, ldf_cg(ml + 2 * s));
            float L = 0.f, O = 0.f;
            for (int s = 0; s < S; ++s) {
                const float m = ldf_cg(ml + 2 * s);
                const float a = (m == -INFINITY) ? 0.f : expf(m - M);
                L += ldf_cg(ml + 2 * s + 1) * a;
                O += ldf_cg(p.part_o + ((size_t)bh * S + s) * d + e) * a;
            }
            xs[idx] = __float2half_rn(O * (1.f / (L + 1e-6f)));
        }
        consumers_sync();
        return;
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
        for (int i = tid; i < pk; i += kStepConsumers) {
            const Pack16 v = ld16_cg(xg + (size_t)i * 8);
            st16(xs + (size_t)b * K + (size_t)i * 8, v);
            if (op.xsrc == kXEmbed && blockIdx.x == 0)   // the residual stream starts as the embedding
                st16(reinterpret_cast<__half*>(p.hidden_buf) + (size_t)b * K + (size_t)i * 8, v);
            if (op.norm) {
                float f[8];
                Unpack<__half>::get(v, f);
#pragma unroll
                for (int e = 0; e < 8; ++e) ss[b] = fmaf(f[e], f[e], ss[b]);
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
        const __half* gm = reinterpret_cast<const __half*>(op.gamma);
        const bool gfirst = op.norm == 2;
        for (int i = tid; i < pk; i += kStepConsumers) {
            float g[8];
            Unpack<__half>::get(ld16(gm + (size_t)i * 8), g);
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
// consumers: one GEMV op. Warp w owns row pair w of every ring stage; two rows share each x packet.
// Epilogues are those of gemv.cu (same rounding points as the unfused launch sequence).
// ------------------------------------------------------------------------------------------
template <int NB, int EPI>
__device__ __noinline__ void gemv_consume(const StepOp& op, const StepParams& p, Ring& rg, const __half* xs, int step) {
    constexpr int KSEG = kSegB / 2;
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const int K = op.K;
    const int nseg = (K + KSEG - 1) / KSEG;
    const int last_seg_pk = (K - (nseg - 1) * KSEG) / 8;
    const int u0 = (int)(((long long)op.units * blockIdx.x) / gridDim.x);
    const int u1 = (int)(((long long)op.units * (blockIdx.x + 1)) / gridDim.x);
    const int ngroups = (u1 - u0 + kStepWarps - 1) / kStepWarps;
    __half* y = reinterpret_cast<__half*>(op.y);

    float acc_a[NB], acc_b[NB];
#pragma unroll
    for (int b = 0; b < NB; ++b) acc_a[b] = acc_b[b] = 0.f;

    for (int g = 0; g < ngroups; ++g) {
        const int unit = u0 + g * kStepWarps + warp;
        const bool active = unit < u1;
        int ra = 0, rb = -1;
        if (active) op_unit_rows<EPI>(op, p.head_size, unit, ra, rb);
        for (int seg = 0; seg < nseg; ++seg) {
            mbar_wait(&rg.full[rg.slot], rg.phase);
            if (active) {
                const int seg_pk = (seg == nseg - 1) ? last_seg_pk : KSEG / 8;
                const unsigned char* sa = rg.base + (size_t)rg.slot * kStageB + (2 * warp) * kSegB;
                const unsigned char* sb = sa + kSegB;
                const __half* xseg = xs + (size_t)seg * KSEG;
                Pack16 wa[kJS], wb[kJS];
#pragma unroll
                for (int j = 0; j < kJS; ++j) {
                    const int pkt = lane + 32 * j;
                    if (pkt < seg_pk) {
                        wa[j] = ld16(sa + pkt * 16);
                        if (rb >= 0) wb[j] = ld16(sb + pkt * 16);
                    }
                }
#pragma unroll
                for (int j = 0; j < kJS; ++j) {
                    const int pkt = lane + 32 * j;
                    if (pkt < seg_pk) {
#pragma unroll
                        for (int b = 0; b < NB; ++b) {
                            const Pack16 xv = ld16(xseg + (size_t)b * K + (size_t)pkt * 8);
                            acc_a[b] = dot16<__half>(wa[j], xv, acc_a[b]);
                            if (rb >= 0) acc_b[b] = dot16<__half>(wb[j], xv, acc_b[b]);
                        }
                    }
                }
            }
            __syncwarp();
            if (lane == 0) mbar_arrive(&rg.empty[rg.slot]);
            if (++rg.slot == rg.stages) {
                rg.slot = 0;
                rg.phase ^= 1;
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
                if (EPI == kEpiSwiglu) {
#pragma unroll
                    for (int b = 0; b < NB; ++b) {
                        if (b >= p.nb) break;
                        const float gv = rnd(acc_a[b]), uv = rnd(acc_b[b]);   // gate/up are stored in T unfused
                        const float s = rnd(gv / (1.f + expf(-gv)));
                        y[(size_t)b * op.N + ra] = __float2half_rn(s * uv);
                    }
                } else if (EPI == kEpiRopeKV) {
                    const int d = p.head_size, hd = d >> 1;
                    const int head = ra / d, i = ra - head * d;
                    const int qkv_heads = p.head_num + 2 * p.kv_head_num;
                    float c = 1.f, s = 0.f;
                    if (head < p.head_num + p.kv_head_num && i < (p.rotary_dim >> 1)) {
                        const float ang = (float)(step - 1) / p.rope_den[i];
                        c = cosf(ang);
                        s = sinf(ang);
                    }
#pragma unroll
                    for (int b = 0; b < NB; ++b) {
                        if (b >= p.nb) break;
                        const float av = rnd(acc_a[b]), bv = rnd(acc_b[b]);   // qkv_buf is stored in T before RoPE
                        const __half oa = __float2half_rn(c * av - s * bv);
                        const __half ob = __float2half_rn(c * bv + s * av);
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
                } else {
#pragma unroll
                    for (int b = 0; b < NB; ++b) {
                        if (b >= p.nb) break;
                        float av = acc_a[b], bv = acc_b[b];
                        if (EPI == kEpiResidual) {   // linear output rounded to T, then T(out + residual)
                            av = rnd(av) + __half2float(ldh_cg(y + (size_t)b * op.N + ra));
                            if (rb >= 0) bv = rnd(bv) + __half2float(ldh_cg(y + (size_t)b * op.N + rb));
                        }
                        y[(size_t)b * op.N + ra] = __float2half_rn(av);
                        if (rb >= 0) y[(size_t)b * op.N + rb] = __float2half_rn(bv);
                    }
                }
            }
#pragma unroll
            for (int b = 0; b < NB; ++b) acc_a[b] = acc_b[b] = 0.f;
        }
    }
}

// ------------------------------------------------------------------------------------------
// consumers: split-KV decode attention (decode_attention.cu's kernel restated for 8 warps inside
// the persistent kernel). Work item = (sequence, kv head, split); head_size 128 (16 packets / row).
// The new token's k,v were written to cache row step-1 by the qkv op's epilogue.
// ------------------------------------------------------------------------------------------
__device__ __forceinline__ float exp_guard_s(float m, float M) { return (m == -INFINITY) ? 0.f : expf(m - M); }

template <int G>
__device__ __noinline__ void attn_consume(const StepOp& op, const StepParams& p, float* sm, int step) {
    constexpr int PR = 16, E = 8, D = 128, KPI = 2, UN = 4;
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const int pi = lane % PR, ks = lane / PR;
    const int splits = p.splits;
    const int items = p.nb * p.kv_head_num * splits;
    const int qkv_heads = p.head_num + 2 * p.kv_head_num;
    const float scale = rsqrtf((float)D);
    const int per = (step + splits - 1) / splits;

    for (int item = blockIdx.x; item < items; item += gridDim.x) {
        const int split = item % splits, bh = item / splits;
        const int b = bh / p.kv_head_num, kvh = bh - b * p.kv_head_num;
        const __half* tok = reinterpret_cast<const __half*>(p.qkv) + (size_t)b * qkv_heads * D;
        const __half* kc = reinterpret_cast<const __half*>(op.k_cache) + ((size_t)b * p.kv_head_num + kvh) * p.max_seq_len * D;
        const __half* vc = reinterpret_cast<const __half*>(op.v_cache) + ((size_t)b * p.kv_head_num + kvh) * p.max_seq_len * D;
        const int k0 = split * per, k1 = min(step, k0 + per);

        Pack16 q[G];
#pragma unroll
        for (int g = 0; g < G; ++g) q[g] = ld16_cg(tok + (size_t)(kvh * G + g) * D + pi * E);
        float m[G], l[G], o[G][E];
#pragma unroll
        for (int g = 0; g < G; ++g) {
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
            float s[G][UN];
#pragma unroll
            for (int u = 0; u < UN; ++u) {
#pragma unroll
                for (int g = 0; g < G; ++g) {
                    float dsum = ok[u] ? dot16<__half>(kp[u], q[g], 0.f) : 0.f;
#pragma unroll
                    for (int off = PR >> 1; off > 0; off >>= 1) dsum += __shfl_xor_sync(0xffffffffu, dsum, off);
                    s[g][u] = ok[u] ? dsum * scale : -INFINITY;
                }
            }
#pragma unroll
            for (int g = 0; g < G; ++g) {
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
        for (int g = 0; g < G; ++g) {
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
            for (int g = 0; g < G; ++g) {
                float* dst = sm + ((size_t)warp * G + g) * (2 + D);
                if (pi == 0) { dst[0] = m[g]; dst[1] = l[g]; }
#pragma unroll
                for (int e = 0; e < E; ++e) dst[2 + pi * E + e] = o[g][e];
            }
        }
        consumers_sync();
        for (int idx = threadIdx.x; idx < G * D; idx += kStepConsumers) {
            const int g = idx / D, e = idx - g * D;
            float M = -INFINITY;
#pragma unroll
            for (int w = 0; w < kStepWarps; ++w) M = fmaxf(M, sm[((size_t)w * G + g) * (2 + D)]);
            float L = 0.f, O = 0.f;
#pragma unroll
            for (int w = 0; w < kStepWarps; ++w) {
                const float* src = sm + ((size_t)w * G + g) * (2 + D);
                const float a = exp_guard_s(src[0], M);
                L += src[1] * a;
                O += src[2 + e] * a;
            }
            const int head = kvh * G + g;
            if (splits == 1) {
                reinterpret_cast<__half*>(p.attn)[((size_t)b * p.head_num + head) * D + e] =
                    __float2half_rn(O / (L + 1e-6f));
            } else {
                const size_t slot = ((size_t)b * p.head_num + head) * splits + split;
                p.part_o[slot * D + e] = O;
                if (e == 0) { p.part_ml[slot * 2] = M; p.part_ml[slot * 2 + 1] = L; }
            }
        }
        consumers_sync();   // sm is reused by the next item
    }
}

// greedy pick (topK.cu:147 with K = 1; ties -> lower index) + advance (ids_cur <- ids_next, step+1)
__device__ __noinline__ void final_consume(const StepParams& p, float* sv, int* si) {
    const int lane = threadIdx.x & 31, wid = threadIdx.x >> 5;
    for (int b = blockIdx.x; b < p.nb; b += gridDim.x) {
        const __half* row = reinterpret_cast<const __half*>(p.logits) + (size_t)b * p.vocab;
        float bv = -INFINITY;
        int bi = 0x7fffffff;
        for (int j = threadIdx.x; j < p.vocab; j += kStepConsumers) {
            const float v = __half2float(ldh_cg(row + j));
            if (v > bv || (v == bv && j < bi)) { bv = v; bi = j; }
        }
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

// ------------------------------------------------------------------------------------------
// the kernel
// ------------------------------------------------------------------------------------------
template <int NB, int G>
__global__ void __launch_bounds__(kStepThreads, 1) decode_step_kernel(const StepParams p) {
    extern __shared__ __align__(128) unsigned char smem[];
    unsigned char* ring = smem;
    __half* xs = reinterpret_cast<__half*>(smem + (size_t)p.stages * kStageB);
    uint64_t* full_bar = reinterpret_cast<uint64_t*>(smem + (size_t)p.stages * kStageB + p.xs_bytes);
    uint64_t* empty_bar = full_bar + kStepMaxStages;
    float* red = reinterpret_cast<float*>(empty_bar + kStepMaxStages);   // [NB*8 + NB] floats, also argmax scratch

    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    if (threadIdx.x == 0) {
        for (int s = 0; s < p.stages; ++s) {
            mbar_init(&full_bar[s], 1);
            mbar_init(&empty_bar[s], kStepWarps);
        }
        mbar_fence_init();
    }
    __syncthreads();

    Ring rg{ring, full_bar, empty_bar, p.stages, 0, 0};
    if (warp == kStepWarps) {
        step_producer(p, rg, lane);
        return;
    }

    // ---- consumers
    const int step = *p.step_dev;
    const unsigned base = p.sync[1];
    unsigned epoch = 0;
    for (int oi = 0; oi < p.n_ops; ++oi) {
        const StepOp& op = p.ops[oi];
        if (op.type == kOpGemv) {
            stage_x<NB>(op, p, xs, red);
            switch (op.epi) {
                case kEpiResidual: gemv_consume<NB, kEpiResidual>(op, p, rg, xs, step); break;
                case kEpiSwiglu: gemv_consume<NB, kEpiSwiglu>(op, p, rg, xs, step); break;
                case kEpiRopeKV: gemv_consume<NB, kEpiRopeKV>(op, p, rg, xs, step); break;
                default: gemv_consume<NB, kEpiPlain>(op, p, rg, xs, step); break;
            }
        } else if (op.type == kOpAttn) {
            attn_consume<G>(op, p, reinterpret_cast<float*>(xs), step);
        } else {
            final_consume(p, red, reinterpret_cast<int*>(red + 16));
        }
        if (oi + 1 < p.n_ops) {
            ++epoch;
            grid_barrier(p.sync, base + epoch * gridDim.x);
        }
    }
    // last op done by this CTA: the final arrival advances the launch base and the step
    consumers_sync();
    if (threadIdx.x == 0) {
        __threadfence();
        const unsigned total = (epoch + 1) * gridDim.x;
        const unsigned prev = atomicAdd(&p.sync[0], 1u);
        if (prev + 1 == base + total) {   // every CTA has finished its final op
            p.sync[1] = base + total;
            *p.step_dev = step + 1;
        }
    }
}

// ------------------------------------------------------------------------------------------
// host side: the program is built once per engine; a launch = one cooperative kernel
// ------------------------------------------------------------------------------------------
static size_t step_smem(int stages, size_t xs_bytes, int nb_t) {
    return (size_t)stages * kStageB + xs_bytes + 2 * kStepMaxStages * sizeof(uint64_t) +
           (size_t)(nb_t * kStepWarps + nb_t + 32) * sizeof(float) + 64;
}

template <int NB, int G>
static cudaError_t step_launch_t(const StepProgram& sp, cudaStream_t s) {
    auto kern = decode_step_kernel<NB, G>;
    static bool attr_set = false;
    if (!attr_set) {
        cudaError_t e = cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, 227 * 1024);
        if (e != cudaSuccess) return e;
        attr_set = true;
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

void step_program_destroy(StepProgram* sp) {
    if (!sp) return;
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
    while (stages >= 2 && step_smem(stages, xs_bytes, nb_t) > 227 * 1024) --stages;
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
    return sp;
}

}  // namespace b200
