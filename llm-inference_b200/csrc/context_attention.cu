// Prefill (context) attention on tcgen05 / TMEM without materialised scores, mask or repeated KV.
//
// Replaces the five-launch sequence of LLaMAContextAttentionLayer::forward
// (src/layers/attention/context_attention.cpp:136-161): launchRepeatKVCache (repeat_kv.cu:76) ->
// launchLinearStridedBatchGemm q*k^T (linear.cu:127) -> launchScaleMaskAndSoftmax
// (attn_softmax_kernel.cu:337; scale, +(1-mask)*-10000, exp(x-max)/(sum+1e-6), :109-173) ->
// launchLinearStridedBatchGemm p*v -> launchTransposeOutRemovePadding
// (fused_transpose_and_remv_pad.cu:51). Semantics kept: query t of sequence b sits at absolute
// position history_len[b] + t and sees cache rows [0, history_len[b] + t]; fp32 softmax with the
// 1e-6 in the denominator; a masked key contributes exp(-10000 + ...) which is 0 in fp32, so masked
// keys (and whole key tiles above the diagonal) are skipped instead of added; GQA reads kv head
// head / (h / hk) directly (RepeatKV never materialised); the output is written token-major with the
// padding removed.
//
// One CTA per (sequence, head, 128-query tile), 320 threads, heaviest (last) query tiles first:
//   warp 0      TMA producer: Q tile once, then K_j / V_j tiles (128 keys x 128 dims, two 128B-swizzled
//               [128 x 64] boxes each) into 2-stage rings;
//   warp 1      MMA issuer: S_j = Q K_j^T (both K-major, 8 x UMMA K=16) into one of two TMEM score
//               buffers; O_j = P_j V_j (A = P from smem, B = V as MN-major operand - V rows are keys,
//               exactly what the PV contraction runs over) into one of two TMEM output buffers;
//               S_{j+1} is issued before P_j V_j so the tensor pipe works under the softmax of tile j;
//   warps 2-9   compute: thread = (query row, 64-column half); warps w and w+4 share a TMEM lane quarter.
//               The half row of scores is read from TMEM once into registers, the two halves exchange the
//               row maximum through smem (one named barrier per tile), P_j written as fp16 into the
//               swizzled K-major A-operand layout; then o = o * alpha_{j-1} + O_{j-1} for the thread's 64
//               output columns (tcgen05.ld), so P_{j-1} V_{j-1} runs under the softmax of tile j;
//               final o / (l0 + l1 + 1e-6) -> fp16 -> out.
//               Measured history (ncu at cfg3, tools/ctx_trace.py): v1 = 4 softmax + 4 accumulate warps with two
//               TMEM passes, 324 us per layer; this version 261 us, a key tile = 2850 clocks of the compute warps
//               (425 score read + max exchange, 1580 exp2 - the 128x128 ex2 alone are 1024 clocks of the 16/clk
//               MUFU pipe -, 670 fold) against 1024 tensor clocks. Tried and dropped: two warp groups on
//               alternating key tiles with private (m, l, O) streams merged at the end (so that one group's MUFU
//               phase would run under the other's TMEM phases): 128 fp32 outputs per thread cap the kernel at
//               168 registers for 320 threads (register allocation in 4-warp units) -> spills, 317 us;
//               ex2.approx.f16x2 (one MUFU op per pair, MUFU.EX2.F16 in SASS) off the diagonal: 271 us, no gain
//               (the extra cvt / unpack instructions cost what the MUFU pipe saves).
// TMEM: 2 x 128 score columns + 2 x 128 output columns = 512. Shared memory: Q 32 KB + K 2 x 32 KB +
// V 2 x 32 KB + P 2 x 32 KB + 3 KB of alpha / l / mbarriers = 226.7 KB (one CTA per SM).
//
// Roofline: tensor; 4 * d flops per (query, visible key) pair = 4 * B * h * d * S(S+1)/2 at history 0.
#include "common.cuh"
#include "launch_count.h"
#include "tc_common.cuh"

namespace b200 {

constexpr int kCtxThreads = 320;
constexpr int kCtxTile = 128;                       // queries per CTA = keys per tile = head size
constexpr uint32_t kBoxBytes = 128 * 64 * 2;        // one [128 rows x 64 fp16] swizzled box
constexpr uint32_t kTileBytes = 2 * kBoxBytes;      // 128 x 128 fp16
constexpr uint32_t kOffQ = 0;
constexpr uint32_t kOffK = kOffQ + kTileBytes;      // [2]
constexpr uint32_t kOffV = kOffK + 2 * kTileBytes;  // [2]
constexpr uint32_t kOffP = kOffV + 2 * kTileBytes;  // [2]
constexpr uint32_t kOffAlpha = kOffP + 2 * kTileBytes;          // float [2][128][2]: row-max / row-sum exchange between column halves
constexpr uint32_t kOffL = kOffAlpha + 4 * 128 * 4;             // (spare)
constexpr uint32_t kOffBar = kOffL + 128 * 4;                   // 20 mbarriers
constexpr uint32_t kOffSlot = kOffBar + 20 * 8;
constexpr uint32_t kCtxSmem = kOffSlot + 16;
static_assert(kCtxSmem <= 232448, "context attention exceeds the 227 KB shared-memory limit");

struct CtxParams {
    __half* out;              // [token_num, head_num * 128]
    const int* input_len;     // [batch]
    const int* history_len;   // [batch]
    int batch, seq_len, head_num, kv_head_num, max_seq_len, q_tiles;
    float scale_log2;         // 1/sqrt(d) * log2(e)
    unsigned long long* trace;   // debug (b200_context_attention_set_trace): %globaltimer stamps of CTA 0
    // paged cache (b200_context_attention_paged): K / V are page pools [n_pages, hk, 128, d]; key tile j of sequence b is
    // page page_table[b * max_pages + j] - a page IS one 128-key tile, so the only change is the TMA row coordinate
    const int* page_table;
    int max_pages;
};

// MN-major operand with the 128B swizzle (canonical layout ((8,n),(8,k)):((1,LBO),(8,SBO)) in 16-byte
// units): rows of the smem tile are K indices (128 bytes = 64 MN elements each), 8 rows form a 1024-byte
// swizzle atom; SBO = distance between 8-row groups along K, LBO = distance between 64-element
// chunks along MN (checked on the B200: with the two swapped the P V product is wrong).
__device__ __forceinline__ uint64_t make_mnmajor_sw128_desc(uint32_t smem_addr, uint32_t lbo_bytes,
                                                            uint32_t sbo_bytes) {
    uint64_t d = 0;
    d |= (uint64_t)((smem_addr & 0x3FFFFu) >> 4);
    d |= (uint64_t)((lbo_bytes >> 4) & 0x3FFFu) << 16;
    d |= (uint64_t)((sbo_bytes >> 4) & 0x3FFFu) << 32;
    d |= (uint64_t)1 << 46;
    d |= (uint64_t)2 << 61;
    return d;
}

// mbarrier wait that traps after ~2 s instead of hanging the device if a pipeline invariant is broken
__device__ __forceinline__ void ctx_wait(uint64_t* bar, uint32_t parity) {
    uint32_t done, spins = 0;
    uint64_t t0 = 0;
    for (;;) {
        asm volatile(
            "{\n"
            ".reg .pred p;\n"
            "mbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\n"
            "selp.u32 %0, 1, 0, p;\n"
            "}\n"
            : "=r"(done)
            : "r"(smem_u32(bar)), "r"(parity)
            : "memory");
        if (done) return;
        if ((++spins & 0xFFFu) == 0) {
            uint64_t now;
            asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(now));
            if (t0 == 0) t0 = now;
            else if (now - t0 > 2000000000ull) __trap();
        }
    }
}

__device__ __forceinline__ void ctx_stamp(const CtxParams& p, int slot) {
    if (p.trace && blockIdx.x == 0) {
        unsigned long long t;
        asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t));
        p.trace[slot] = t;
    }
}

__device__ __forceinline__ float ex2_approx(float x) {
    float y;
    asm("ex2.approx.ftz.f32 %0, %1;" : "=f"(y) : "f"(x));
    return y;
}
// packed fp32x2 arithmetic (sm_100: one FFMA2 / FADD2 issue slot for two lanes): the softmax is issue-bound
__device__ __forceinline__ uint64_t pack2(uint32_t lo, uint32_t hi) {
    uint64_t r;
    asm("mov.b64 %0, {%1, %2};" : "=l"(r) : "r"(lo), "r"(hi));
    return r;
}
__device__ __forceinline__ uint64_t pack2f(float lo, float hi) { return pack2(__float_as_uint(lo), __float_as_uint(hi)); }
__device__ __forceinline__ void unpack2(uint64_t v, float& lo, float& hi) {
    uint32_t a, b;
    asm("mov.b64 {%0, %1}, %2;" : "=r"(a), "=r"(b) : "l"(v));
    lo = __uint_as_float(a);
    hi = __uint_as_float(b);
}
__device__ __forceinline__ uint64_t fma2(uint64_t a, uint64_t b, uint64_t c) {
    uint64_t d;
    asm("fma.rn.f32x2 %0, %1, %2, %3;" : "=l"(d) : "l"(a), "l"(b), "l"(c));
    return d;
}
__device__ __forceinline__ uint64_t add2(uint64_t a, uint64_t b) {
    uint64_t d;
    asm("add.rn.f32x2 %0, %1, %2;" : "=l"(d) : "l"(a), "l"(b));
    return d;
}
__device__ __forceinline__ void fence_proxy_async_smem() {
    asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
}
__device__ __forceinline__ void st_shared_v4(uint32_t addr, uint32_t a, uint32_t b, uint32_t c, uint32_t d) {
    asm volatile("st.shared.v4.b32 [%0], {%1, %2, %3, %4};" ::"r"(addr), "r"(a), "r"(b), "r"(c), "r"(d) : "memory");
}
__device__ __forceinline__ uint32_t pack_h2(float a, float b) {
    const __half2 h = __floats2half2_rn(a, b);
    return *reinterpret_cast<const uint32_t*>(&h);
}

__global__ void __launch_bounds__(kCtxThreads, 1)
context_attn_tcgen05_kernel(const __grid_constant__ CUtensorMap map_q, const __grid_constant__ CUtensorMap map_k,
                            const __grid_constant__ CUtensorMap map_v, const CtxParams p) {
    extern __shared__ __align__(1024) unsigned char smem[];
    pdl_wait();   // lengths, q and the caches come from the preceding kernels
    // ---- which (query tile, sequence, head) ----
    const int bh = p.batch * p.head_num;
    const int qt = p.q_tiles - 1 - (int)(blockIdx.x / bh);
    const int b = (int)(blockIdx.x % bh) / p.head_num, head = (int)(blockIdx.x % bh) % p.head_num;
    const int in_len = p.input_len[b], hist = p.history_len[b];
    const int q0 = qt * kCtxTile;
    if (q0 >= in_len) return;   // the whole CTA leaves before touching barriers / TMEM
    const int ctx_len = hist + in_len;
    const int q_hi = min(q0 + kCtxTile, in_len);
    const int kv_end = hist + q_hi;                     // keys [0, kv_end) are visible to this tile
    const int n_kt = (kv_end + kCtxTile - 1) / kCtxTile;
    const int kvh = head / (p.head_num / p.kv_head_num);

    uint64_t* bars = reinterpret_cast<uint64_t*>(smem + kOffBar);
    uint64_t* q_full = bars;
    uint64_t* k_full = bars + 1;    // [2]
    uint64_t* k_empty = bars + 3;   // [2]
    uint64_t* v_full = bars + 5;    // [2]
    uint64_t* v_empty = bars + 7;   // [2]
    uint64_t* s_full = bars + 9;    // [2]
    uint64_t* s_empty = bars + 11;  // [2]
    uint64_t* p_full = bars + 13;   // [2]
    uint64_t* o_full = bars + 15;   // [2]
    uint64_t* o_empty = bars + 17;  // [2]
    float* alpha_s = reinterpret_cast<float*>(smem + kOffAlpha);
    uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(smem + kOffSlot);

    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    if (threadIdx.x == 0) {
        if (smem_u32(smem) & 1023u) __trap();   // swizzled tiles need a 1024-byte aligned base
        mbar_init(q_full, 1);
        for (int s = 0; s < 2; ++s) {
            mbar_init(&k_full[s], 1);
            mbar_init(&k_empty[s], 1);
            mbar_init(&v_full[s], 1);
            mbar_init(&v_empty[s], 1);
            mbar_init(&s_full[s], 1);
            mbar_init(&s_empty[s], 8);
            mbar_init(&p_full[s], 8);
            mbar_init(&o_full[s], 1);
            mbar_init(&o_empty[s], 8);
        }
        mbar_fence_init();
        tma_prefetch_desc(&map_q);
        tma_prefetch_desc(&map_k);
        tma_prefetch_desc(&map_v);
    }
    if (warp == 1) tmem_alloc(tmem_slot, 512);
    tc_fence_before();
    __syncthreads();
    tc_fence_after();
    const uint32_t tmem_base = *tmem_slot;
    pdl_launch_dependents();

    if (warp == 0) {
        // ================= TMA producer =================
        const int qrow = (b * p.head_num + head) * p.seq_len + q0;
        const int krow = (b * p.kv_head_num + kvh) * p.max_seq_len;
        const int* pages = p.page_table ? p.page_table + (size_t)b * p.max_pages : nullptr;
        auto tile_row = [&](int j) {   // first row of key tile j in the K / V tensor map
            return pages ? (__ldg(pages + j) * p.kv_head_num + kvh) * kCtxTile : krow + j * kCtxTile;
        };
        if (lane == 0) {
            mbar_expect_tx(q_full, kTileBytes);
            tma_load_2d(smem + kOffQ, &map_q, 0, qrow, q_full);
            tma_load_2d(smem + kOffQ + kBoxBytes, &map_q, 64, qrow, q_full);
        }
        for (int j = 0; j < n_kt; ++j) {
            const int s = j & 1;
            const uint32_t ph = (uint32_t)(j >> 1) & 1u;
            ctx_wait(&k_empty[s], ph ^ 1u);
            if (lane == 0) {
                unsigned char* dst = smem + kOffK + (uint32_t)s * kTileBytes;
                mbar_expect_tx(&k_full[s], kTileBytes);
                tma_load_2d(dst, &map_k, 0, tile_row(j), &k_full[s]);
                tma_load_2d(dst + kBoxBytes, &map_k, 64, tile_row(j), &k_full[s]);
            }
            ctx_wait(&v_empty[s], ph ^ 1u);
            if (lane == 0) {
                unsigned char* dst = smem + kOffV + (uint32_t)s * kTileBytes;
                mbar_expect_tx(&v_full[s], kTileBytes);
                tma_load_2d(dst, &map_v, 0, tile_row(j), &v_full[s]);
                tma_load_2d(dst + kBoxBytes, &map_v, 64, tile_row(j), &v_full[s]);
            }
            __syncwarp();
        }
    } else if (warp == 1) {
        // ================= MMA issuer =================
        constexpr uint32_t idesc_s = make_idesc_f16(kCtxTile, kCtxTile);       // A, B K-major
        constexpr uint32_t idesc_pv = idesc_s | (1u << 16);                    // B (= V) MN-major
        const uint32_t q_addr = smem_u32(smem + kOffQ);
        auto issue_s = [&](int j) {
            const int s = j & 1;
            const uint32_t ph = (uint32_t)(j >> 1) & 1u;
            ctx_wait(&k_full[s], ph);
            ctx_wait(&s_empty[s], ph ^ 1u);
            tc_fence_after();
            if (lane == 0) {
                const uint32_t k_addr = smem_u32(smem + kOffK + (uint32_t)s * kTileBytes);
#pragma unroll
                for (int ks = 0; ks < kCtxTile / kUmmaK; ++ks) {
                    const uint32_t off = (uint32_t)(ks >> 2) * kBoxBytes;
                    umma_f16(tmem_base + (uint32_t)(s * kCtxTile),
                             make_kmajor_sw128_desc(q_addr + off) + (uint64_t)(2 * (ks & 3)),
                             make_kmajor_sw128_desc(k_addr + off) + (uint64_t)(2 * (ks & 3)), idesc_s, ks > 0 ? 1u : 0u);
                }
                umma_commit(&k_empty[s]);
                umma_commit(&s_full[s]);
                ctx_stamp(p, 8 * j + 0);   // S_j issued
            }
            __syncwarp();
        };
        auto issue_pv = [&](int j) {
            const int s = j & 1;
            const uint32_t ph = (uint32_t)(j >> 1) & 1u;
            ctx_wait(&v_full[s], ph);
            ctx_wait(&p_full[s], ph);
            ctx_wait(&o_empty[s], ph ^ 1u);
            tc_fence_after();
            if (lane == 0) {
                const uint32_t p_addr = smem_u32(smem + kOffP + (uint32_t)s * kTileBytes);
                const uint32_t v_addr = smem_u32(smem + kOffV + (uint32_t)s * kTileBytes);
#pragma unroll
                for (int ks = 0; ks < kCtxTile / kUmmaK; ++ks) {
                    // 16 keys = two 8-row groups of 1024 bytes; the two 64-wide halves of d are 16 KB apart
                    const uint64_t dv = make_mnmajor_sw128_desc(v_addr + (uint32_t)ks * 2048u, kBoxBytes, 1024u);
                    umma_f16(tmem_base + (uint32_t)(256 + s * kCtxTile),
                             make_kmajor_sw128_desc(p_addr + (uint32_t)(ks >> 2) * kBoxBytes) + (uint64_t)(2 * (ks & 3)), dv,
                             idesc_pv, ks > 0 ? 1u : 0u);
                }
                umma_commit(&v_empty[s]);
                umma_commit(&o_full[s]);
                ctx_stamp(p, 8 * j + 1);   // P_j V_j issued
            }
            __syncwarp();
        };
        ctx_wait(q_full, 0);
        issue_s(0);
        for (int j = 0; j < n_kt; ++j) {
            if (j + 1 < n_kt) issue_s(j + 1);
            issue_pv(j);
        }
    } else {
        // ================= compute warps 2..9: thread = (query row, column half) =================
        // Warps w and w + 4 share TMEM lane quarter w & 3: the first takes score / output columns [0, 64),
        // the second [64, 128). Two warps per scheduler hide each other's TMEM / MUFU latency; the row
        // maximum is the only value the two halves exchange per tile (smem, one named barrier); the row
        // sums stay per half and are added once at the end.
        const int quarter = warp & 3, half = (warp - 2) >> 2;
        const int row = quarter * 32 + lane;
        const int qpos = min(hist + q0 + row, ctx_len - 1);   // pad rows behave like the last real row
        const float sl = p.scale_log2;
        const uint32_t lane_addr = tmem_base + ((uint32_t)(quarter * 32) << 16) + (uint32_t)(half * 64);
        float m = -1e30f, l = 0.f, alpha_prev = 0.f;
        uint64_t o2[32];   // this thread's 64 output columns as fp32 pairs
#pragma unroll
        for (int e = 0; e < 32; ++e) o2[e] = 0ull;
        auto accumulate = [&](int jj, float alpha) {   // o = o * alpha + O_jj (this thread's 64 columns)
            const int s = jj & 1;
            ctx_wait(&o_full[s], (uint32_t)(jj >> 1) & 1u);
            tc_fence_after();
            const uint64_t a2 = pack2f(alpha, alpha);
#pragma unroll
            for (int c = 0; c < 2; ++c) {
                uint32_t v[32];
                tmem_ld_32x32(lane_addr + (uint32_t)(256 + s * kCtxTile + c * 32), v);
#pragma unroll
                for (int e = 0; e < 16; ++e) o2[c * 16 + e] = fma2(o2[c * 16 + e], a2, pack2(v[2 * e], v[2 * e + 1]));
            }
            tc_fence_before();
            __syncwarp();
            if (lane == 0) mbar_arrive(&o_empty[s]);
        };
        for (int j = 0; j < n_kt; ++j) {
            const int s = j & 1;
            const uint32_t ph = (uint32_t)(j >> 1) & 1u;
            const int kbase = j * kCtxTile + half * 64;       // first key of this thread's columns
            const bool need_mask = j * kCtxTile + kCtxTile - 1 > hist + q0;
            ctx_wait(&s_full[s], ph);
            tc_fence_after();
            if (threadIdx.x == 64) ctx_stamp(p, 8 * j + 2);   // S_j visible to the softmax
            const uint32_t taddr = lane_addr + (uint32_t)(s * kCtxTile);
            // this thread's 64 scores, read from TMEM ONCE (TMEM reads run at 64 B/clk per SM: the two-pass
            // version read 192 KB per tile = 3072 clocks against 1024 clocks of MMA work)
            uint32_t sv[64];
            tmem_ld_32x32(taddr, *reinterpret_cast<uint32_t(*)[32]>(&sv[0]));
            tmem_ld_32x32(taddr + 32u, *reinterpret_cast<uint32_t(*)[32]>(&sv[32]));
            tc_fence_before();
            __syncwarp();
            if (lane == 0) mbar_arrive(&s_empty[s]);   // the score buffer is free for S_{j+2}
            float mx = -1e30f;
            if (need_mask) {
#pragma unroll
                for (int e = 0; e < 64; ++e) mx = (kbase + e <= qpos) ? fmaxf(mx, __uint_as_float(sv[e])) : mx;
            } else {
#pragma unroll
                for (int e = 0; e < 64; ++e) mx = fmaxf(mx, __uint_as_float(sv[e]));
            }
            float* xch = alpha_s + (j & 1) * 256;             // [128 rows][2 halves], double buffered per tile
            xch[row * 2 + half] = mx;
            asm volatile("bar.sync 1, 256;" ::: "memory");
            mx = fmaxf(fmaxf(mx, xch[row * 2 + (half ^ 1)]), m);
            if (threadIdx.x == 64) ctx_stamp(p, 8 * j + 3);   // row maxima exchanged
            const float alpha = ex2_approx((m - mx) * sl);
            m = mx;
            const float msl = m * sl;
            // p = exp2(x * sl - m * sl), row sum, fp16 P into the swizzled K-major A-operand layout.
            // P[s] was the A operand of P_{j-2} V_{j-2}; this thread has already consumed O_{j-2} (below).
            const uint32_t p_row = smem_u32(smem + kOffP + (uint32_t)s * kTileBytes) + (uint32_t)half * kBoxBytes + (uint32_t)row * 128u;
            float rs = 0.f;
            if (need_mask) {   // diagonal tiles: per-element visibility
#pragma unroll
                for (int c8 = 0; c8 < 8; ++c8) {   // 16-byte chunk (8 keys) inside this half's 64-key box
                    float pe[8];
#pragma unroll
                    for (int e = 0; e < 8; ++e) {
                        const float pv = ex2_approx(fmaf(__uint_as_float(sv[c8 * 8 + e]), sl, -msl));
                        pe[e] = (kbase + c8 * 8 + e <= qpos) ? pv : 0.f;
                        rs += pe[e];
                    }
                    st_shared_v4(p_row + (uint32_t)((c8 ^ (row & 7)) << 4), pack_h2(pe[0], pe[1]), pack_h2(pe[2], pe[3]),
                                 pack_h2(pe[4], pe[5]), pack_h2(pe[6], pe[7]));
                }
            } else {           // every key of the tile is visible: packed fp32x2 arithmetic, no selects
                const uint64_t sl2 = pack2f(sl, sl), nm2 = pack2f(-msl, -msl);
                uint64_t rs2 = 0ull;
#pragma unroll
                for (int c8 = 0; c8 < 8; ++c8) {
                    uint32_t h[4];
#pragma unroll
                    for (int e = 0; e < 4; ++e) {
                        float a0, a1;
                        unpack2(fma2(pack2(sv[c8 * 8 + 2 * e], sv[c8 * 8 + 2 * e + 1]), sl2, nm2), a0, a1);
                        const float p0 = ex2_approx(a0), p1 = ex2_approx(a1);
                        rs2 = add2(rs2, pack2f(p0, p1));
                        h[e] = pack_h2(p0, p1);
                    }
                    st_shared_v4(p_row + (uint32_t)((c8 ^ (row & 7)) << 4), h[0], h[1], h[2], h[3]);
                }
                float r0, r1;
                unpack2(rs2, r0, r1);
                rs = r0 + r1;
            }
            l = l * alpha + rs;
            if (j == n_kt - 1) {
                // cache rows at or beyond the context length are not ours: 0 * garbage must stay 0
                const int valid = ctx_len - j * kCtxTile;
                if (valid < kCtxTile) {
                    ctx_wait(&v_full[s], ph);
                    if (row >= valid) {
                        const uint32_t v_row = smem_u32(smem + kOffV + (uint32_t)s * kTileBytes) + (uint32_t)half * kBoxBytes + (uint32_t)row * 128u;
#pragma unroll
                        for (int q = 0; q < 8; ++q) st_shared_v4(v_row + (uint32_t)(q << 4), 0u, 0u, 0u, 0u);
                    }
                }
            }
            fence_proxy_async_smem();   // P (and the zeroed V rows) become visible to the tensor core
            __syncwarp();
            if (lane == 0) mbar_arrive(&p_full[s]);
            if (threadIdx.x == 64) ctx_stamp(p, 8 * j + 4);   // P_j published
            if (j > 0) accumulate(j - 1, alpha_prev);   // P_{j-1} V_{j-1} ran under this tile's softmax
            if (threadIdx.x == 64) ctx_stamp(p, 8 * j + 5);   // O_{j-1} folded in
            alpha_prev = alpha;
        }
        accumulate(n_kt - 1, alpha_prev);
        // ---- row sum over both halves, normalise, packed token-major store
        float* xch = alpha_s + (n_kt & 1) * 256;   // the buffer tile n_kt would use: its last readers passed a barrier since
        xch[row * 2 + half] = l;
        asm volatile("bar.sync 1, 256;" ::: "memory");
        const float inv = 1.f / (l + xch[row * 2 + (half ^ 1)] + 1e-6f);
        const int t = q0 + row;
        if (t < in_len) {
            int tok = t;
            for (int i = 0; i < b; ++i) tok += p.input_len[i];
            __half* dst = p.out + ((size_t)tok * p.head_num + head) * kCtxTile + half * 64;
#pragma unroll
            for (int q = 0; q < 8; ++q) {
                float f[8];
#pragma unroll
                for (int e = 0; e < 4; ++e) {
                    unpack2(o2[q * 4 + e], f[2 * e], f[2 * e + 1]);
                    f[2 * e] *= inv;
                    f[2 * e + 1] *= inv;
                }
                st16(dst + q * 8, Unpack<__half>::put(f));
            }
        }
    }
    tc_fence_before();
    __syncthreads();
    if (warp == 1) {
        tc_fence_after();
        tmem_dealloc(tmem_base, 512);
    }
}

}  // namespace b200

using namespace b200;

static unsigned long long* g_ctx_trace = nullptr;
/* debug: device buffer of >= 8 * tiles u64; CTA 0 (the last query tile of sequence 0, head 0) stamps %globaltimer per
 * key tile j at [8j + 0..5] = S_j issued, P_j V_j issued, S_j visible, maxima exchanged, P_j published, O_{j-1} folded */
extern "C" void b200_context_attention_set_trace(void* dev_buf) { g_ctx_trace = (unsigned long long*)dev_buf; }

static int context_attention_impl(const void* q, const void* k_cache_layer, const void* v_cache_layer,
                                  void* out, const int* padding_offset, const int* input_len,
                                  const int* history_len, int batch, int seq_len, int token_num,
                                  int head_num, int kv_head_num, int head_size, int max_seq_len,
                                  int dtype, const int* page_table, int max_pages, long long n_pages, void* stream) {
    (void)padding_offset;   // the packed token index is the prefix sum of input_len (cal_paddingoffset.cu:51-72)
    if (!q || !k_cache_layer || !v_cache_layer || !out || !input_len || !history_len) return B200_EINVAL;
    if (batch < 1 || seq_len < 1 || token_num < 0 || head_num < 1 || kv_head_num < 1 || head_num % kv_head_num ||
        max_seq_len < 1)
        return B200_EINVAL;
    // tensor-core tiles exist for fp16 with 128-wide heads; anything else takes the reference-order path
    if (dtype != B200_F16 || head_size != kCtxTile) return B200_EUNSUPPORTED;
    auto al = [](const void* x) { return (reinterpret_cast<uintptr_t>(x) & 15u) == 0; };
    if (!al(q) || !al(k_cache_layer) || !al(v_cache_layer) || !al(out)) return B200_EUNSUPPORTED;
    static int disabled = -1;
    if (disabled < 0) {
        const char* e = getenv("B200_NO_TCGEN05");
        disabled = (e && e[0] == '1') ? 1 : 0;
    }
    if (disabled) return B200_EUNSUPPORTED;
    const long long q_rows = (long long)batch * head_num * seq_len;
    const long long kv_rows = page_table ? n_pages * kv_head_num * kCtxTile : (long long)batch * kv_head_num * max_seq_len;
    if (q_rows + kCtxTile >= (1ll << 31) || kv_rows + kCtxTile >= (1ll << 31)) return B200_EUNSUPPORTED;
    if (token_num == 0) return 0;
    CUtensorMap mq, mk, mv;
    if (!make_tma_map_2d(&mq, q, q_rows, kCtxTile, kCtxTile) || !make_tma_map_2d(&mk, k_cache_layer, kv_rows, kCtxTile, kCtxTile) ||
        !make_tma_map_2d(&mv, v_cache_layer, kv_rows, kCtxTile, kCtxTile))
        return B200_EUNSUPPORTED;
    CtxParams p{};
    p.out = reinterpret_cast<__half*>(out);
    p.input_len = input_len;
    p.history_len = history_len;
    p.batch = batch; p.seq_len = seq_len; p.head_num = head_num; p.kv_head_num = kv_head_num;
    p.max_seq_len = max_seq_len;
    p.q_tiles = (seq_len + kCtxTile - 1) / kCtxTile;
    p.scale_log2 = (1.0f / sqrtf((float)head_size)) * 1.4426950408889634f;
    p.trace = g_ctx_trace;
    p.page_table = page_table;
    p.max_pages = max_pages;
    static unsigned long long attr_done = 0;   // one bit per device
    {
        cudaError_t e = ensure_max_dynamic_smem(reinterpret_cast<const void*>(context_attn_tcgen05_kernel), (int)kCtxSmem, attr_done);
        if (e != cudaSuccess) return (int)e;
    }
    const long long grid = (long long)p.q_tiles * batch * head_num;
    if (grid >= (1ll << 31)) return B200_EUNSUPPORTED;
    cudaError_t e = launch_k(context_attn_tcgen05_kernel, dim3((unsigned)grid), dim3(kCtxThreads), (size_t)kCtxSmem,
                             (cudaStream_t)stream, true, mq, mk, mv, p);
    count_launch();
    return (int)e;
}

extern "C" int b200_context_attention(const void* q, const void* k_cache_layer, const void* v_cache_layer,
                                      void* out, const int* padding_offset, const int* input_len,
                                      const int* history_len, int batch, int seq_len, int token_num,
                                      int head_num, int kv_head_num, int head_size, int max_seq_len,
                                      int dtype, void* stream) {
    return context_attention_impl(q, k_cache_layer, v_cache_layer, out, padding_offset, input_len, history_len, batch, seq_len,
                                  token_num, head_num, kv_head_num, head_size, max_seq_len, dtype, nullptr, 0, 0, stream);
}

/* the same over a paged cache: K / V page pools [n_pages, kv_heads, 128, d] and the sequences' page lists (DEVICE int
 * [batch, max_pages]); a page is exactly one 128-key tile of the kernel, so a tile costs one list lookup in the TMA producer.
 * Every page a sequence's context touches (history + this pass) must be in its list and hold this pass's rows already. */
extern "C" int b200_context_attention_paged(const void* q, const void* k_pool_layer, const void* v_pool_layer, const int* page_table,
                                            int max_pages, int n_pages, void* out, const int* padding_offset, const int* input_len,
                                            const int* history_len, int batch, int seq_len, int token_num, int head_num,
                                            int kv_head_num, int head_size, int dtype, void* stream) {
    if (!page_table || max_pages < 1 || n_pages < 1) return B200_EINVAL;
    static_assert(kCtxTile == B200_KV_PAGE_ROWS, "a page must be one key tile");
    return context_attention_impl(q, k_pool_layer, v_pool_layer, out, padding_offset, input_len, history_len, batch, seq_len,
                                  token_num, head_num, kv_head_num, head_size, max_pages * kCtxTile, dtype, page_table, max_pages,
                                  n_pages, stream);
}
