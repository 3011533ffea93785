// Masked decode self-attention over the KV cache, split-KV (flash-decoding) form.
// Replaces launchDecoderMaskedMHA (src/kernels/fused_decoder_self_attention.cu:80-291, 317-365):
//   append k,v of the new token at row step-1; s_j = (q . K_j) * rsqrt(d), j < step;
//   p = exp(s - max s) / (sum + 1e-6) (:203-221); out = sum_j p_j V_j (:272-287).
// SURVEY defects D1-D6 resolved: kv head = q head / (h/hk), element-correct addressing, layer
// slice includes kv_head_num, softmax over the full length, token-major qkv, fp16 implemented.
//
// HBM-bound: algorithmic bytes = 2*B*hk*step*d*sizeof(T). Layout of the work:
//  * grid (B*hk, n_splits): one CTA streams one contiguous chunk of one kv head's keys/values
//    and serves all G = h/hk query heads of that group from a single read (GQA reuse, so
//    RepeatKV is never materialised);
//  * a key row is PR 16-byte packets; PR lanes own one row, 32/PR rows per warp-wide load,
//    kUnroll loads in flight per lane; dot products reduce with warp shuffles across the PR lanes;
//  * every lane group runs its own online softmax (m, l, o[]) in fp32, merged by shuffles, then
//    across warps through shared memory, then across splits by the combine kernel.
#include <cuda_fp8.h>

#include "common.cuh"
#include "launch_count.h"

namespace b200 {

constexpr int kAttnWarps = 4;
constexpr int kUnroll = 4;

struct AttnParams {
    const void* qkv;  // [B, h+2hk, d]
    void* k_cache;    // [B, hk, max_seq, d]
    void* v_cache;
    void* out;        // [B, h*d]
    float* part_ml;   // [B, h, splits, 2]
    float* part_o;    // [B, h, splits, d]
    int batch, head_num, kv_head_num, head_size, max_seq_len, step;
    const int* step_dev;
    int step_stride;  // 0: *step_dev is the one step all sequences share; 1: step_dev[b] (ragged batch)
    int append_kv;
    const float* k_scale;   // fp8 cache only: one scale per cached row, [B, hk, max_seq]
    const float* v_scale;
    const unsigned char* finished;   // optional [B]: a retired sequence neither appends nor reads its cache (output 0)
    // paged cache (b200_decode_attention_paged): k_cache / v_cache are page pools [n_pages, hk, kPageRows, d] and row
    // `key` of sequence b lives in page page_table[b * max_pages + key / kPageRows]
    const int* page_table;
    int max_pages;
};
constexpr int kPageRows = B200_KV_PAGE_ROWS;
constexpr int kPageShift = 7;
static_assert(kPageRows == 1 << kPageShift, "page size");

// `finished` flags for the launches that follow on this host thread (NULL = off): the reference passes a
// TensorWrapper<bool>* finished to launchDecoderMaskedMHA (fused_decoder_self_attention.h:10-19) and never reads it;
// here a set flag retires the sequence - no K/V append, no cache reads, zero output row.
static thread_local const unsigned char* g_attn_finished = nullptr;

// fp32 data keeps the accurate expf (1e-5 tolerance); half data uses the MUFU fast path
template <typename T> __device__ __forceinline__ float fexp(float x);
template <> __device__ __forceinline__ float fexp<float>(float x) { return expf(x); }
template <> __device__ __forceinline__ float fexp<__half>(float x) { return __expf(x); }
__device__ __forceinline__ float exp_guard(float m, float M) {
    return (m == -INFINITY) ? 0.f : expf(m - M);
}

template <typename T, int G, int PR, bool PAGED = false>
__global__ void __launch_bounds__(kAttnWarps * 32)
decode_attn_split_kernel(const AttnParams p) {
    constexpr int E = DT<T>::kPer16B;
    constexpr int KPI = 32 / PR;  // key rows per warp-wide load
    constexpr int D = PR * E;     // head size
    extern __shared__ float sm[];  // [kAttnWarps][G][2 + D]

    pdl_wait();
    const int bh = blockIdx.x;
    const int b = bh / p.kv_head_num, kvh = bh - b * p.kv_head_num;
    const int step = p.step_dev ? p.step_dev[b * p.step_stride] : p.step;
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const int pi = lane % PR, ks = lane / PR;
    const int qkv_heads = p.head_num + 2 * p.kv_head_num;
    const T* tok = reinterpret_cast<const T*>(p.qkv) + (size_t)b * qkv_heads * D;
    // dense: the (b, kvh) slice of [B, hk, max_seq, d]; paged: the pool, rows found through the sequence's page list
    T* kc = reinterpret_cast<T*>(p.k_cache) + (PAGED ? (size_t)0 : ((size_t)b * p.kv_head_num + kvh) * p.max_seq_len * D);
    T* vc = reinterpret_cast<T*>(p.v_cache) + (PAGED ? (size_t)0 : ((size_t)b * p.kv_head_num + kvh) * p.max_seq_len * D);
    const int* pages = PAGED ? p.page_table + (size_t)b * p.max_pages : nullptr;
    auto row_off = [&](int key) -> size_t {   // element offset of cache row `key` from kc / vc
        if (PAGED) {
            const int pg = __ldg(pages + (key >> kPageShift));
            return (((size_t)pg * p.kv_head_num + kvh) * kPageRows + (size_t)(key & (kPageRows - 1))) * D;
        }
        return (size_t)key * D;
    };

    // keys of this split
    const int splits = gridDim.y;
    const int per = (step + splits - 1) / splits;
    const int k0 = blockIdx.y * per;
    const bool retired = p.finished != nullptr && p.finished[b] != 0;
    const int k1 = retired ? k0 : min(step, k0 + per);   // a retired sequence: empty key range -> (m, l, o) = (-inf, 0, 0)

    // append the new token's k,v (reference :170-172, :274-276) - only the CTA that will read it
    if (p.append_kv && k0 <= step - 1 && step - 1 < k1) {
        if (threadIdx.x < PR) {
            st16(kc + row_off(step - 1) + threadIdx.x * E,
                 ld16(tok + (size_t)(p.head_num + kvh) * D + threadIdx.x * E));
        } else if (threadIdx.x >= 32 && threadIdx.x < 32 + PR) {
            const int t = threadIdx.x - 32;
            st16(vc + row_off(step - 1) + t * E,
                 ld16(tok + (size_t)(p.head_num + p.kv_head_num + kvh) * D + t * E));
        }
    }

    const T* knew = tok + (size_t)(p.head_num + kvh) * D;
    const T* vnew = tok + (size_t)(p.head_num + p.kv_head_num + kvh) * D;
    const int new_row = p.append_kv ? step - 1 : -1;

    // q packets for the G heads of this kv group stay packed in T
    Pack16 q[G];
#pragma unroll
    for (int g = 0; g < G; ++g) q[g] = ld16(tok + (size_t)(kvh * G + g) * D + pi * E);
    const float scale = rsqrtf((float)D);

    float m[G], l[G], o[G][E];
#pragma unroll
    for (int g = 0; g < G; ++g) {
        m[g] = -INFINITY;
        l[g] = 0.f;
#pragma unroll
        for (int e = 0; e < E; ++e) o[g][e] = 0.f;
    }

    // warp w takes key rows k0 + (it*kAttnWarps + w)*KPI*kUnroll + u*KPI + ks
    for (int base = k0 + warp * KPI * kUnroll; base < k1; base += kAttnWarps * KPI * kUnroll) {
        Pack16 kp[kUnroll], vp[kUnroll];
        bool ok[kUnroll];
        size_t ro[kUnroll];
#pragma unroll
        for (int u = 0; u < kUnroll; ++u) {
            const int key = base + u * KPI + ks;
            ok[u] = key < k1;
            ro[u] = (PAGED && ok[u]) ? row_off(key) : 0;   // dense: recomputed from `key` below, nothing kept
        }
#pragma unroll
        for (int u = 0; u < kUnroll; ++u) {
            const int key = base + u * KPI + ks;
            // the row this kernel appended is taken from qkv, never re-read from the cache through the
            // read-only (.nc) path: .nc is only defined for data no thread of the kernel writes
            if (ok[u]) kp[u] = ld16_stream((key == new_row ? knew : kc + (PAGED ? ro[u] : (size_t)key * D)) + pi * E);
        }
#pragma unroll
        for (int u = 0; u < kUnroll; ++u) {
            const int key = base + u * KPI + ks;
            if (ok[u]) vp[u] = ld16_stream((key == new_row ? vnew : vc + (PAGED ? ro[u] : (size_t)key * D)) + pi * E);
        }
        float s[G][kUnroll];
#pragma unroll
        for (int u = 0; u < kUnroll; ++u) {
#pragma unroll
            for (int g = 0; g < G; ++g) {
                float d = ok[u] ? dot16<T>(kp[u], q[g], 0.f) : 0.f;
#pragma unroll
                for (int off = PR >> 1; off > 0; off >>= 1) d += __shfl_xor_sync(0xffffffffu, d, off);
                s[g][u] = ok[u] ? d * scale : -INFINITY;
            }
        }
#pragma unroll
        for (int g = 0; g < G; ++g) {
            float mn = m[g];
#pragma unroll
            for (int u = 0; u < kUnroll; ++u) mn = fmaxf(mn, s[g][u]);
            if (mn != -INFINITY) {
                const float corr = exp_guard(m[g], mn);
                l[g] *= corr;
#pragma unroll
                for (int e = 0; e < E; ++e) o[g][e] *= corr;
                m[g] = mn;
#pragma unroll
                for (int u = 0; u < kUnroll; ++u) {
                    if (ok[u]) {
                        const float pr = fexp<T>(s[g][u] - mn);
                        l[g] += pr;
                        float vf[E];
                        Unpack<T>::get(vp[u], vf);
#pragma unroll
                        for (int e = 0; e < E; ++e) o[g][e] = fmaf(pr, vf[e], o[g][e]);
                    }
                }
            }
        }
    }

    // merge the KPI lane groups of the warp
#pragma unroll
    for (int off = PR; off < 32; off <<= 1) {
#pragma unroll
        for (int g = 0; g < G; ++g) {
            const float mo = __shfl_xor_sync(0xffffffffu, m[g], off);
            const float lo = __shfl_xor_sync(0xffffffffu, l[g], off);
            const float M = fmaxf(m[g], mo);
            const float a = exp_guard(m[g], M), bb = exp_guard(mo, M);
            l[g] = l[g] * a + lo * bb;
#pragma unroll
            for (int e = 0; e < E; ++e) {
                const float oo = __shfl_xor_sync(0xffffffffu, o[g][e], off);
                o[g][e] = o[g][e] * a + oo * bb;
            }
            m[g] = M;
        }
    }
    // across warps through shared memory
    if (ks == 0) {
#pragma unroll
        for (int g = 0; g < G; ++g) {
            float* dst = sm + ((size_t)warp * G + g) * (2 + D);
            if (pi == 0) { dst[0] = m[g]; dst[1] = l[g]; }
#pragma unroll
            for (int e = 0; e < E; ++e) dst[2 + pi * E + e] = o[g][e];
        }
    }
    __syncthreads();
    pdl_launch_dependents();
    for (int idx = threadIdx.x; idx < G * D; idx += blockDim.x) {
        const int g = idx / D, e = idx - g * D;
        float M = -INFINITY;
#pragma unroll
        for (int w = 0; w < kAttnWarps; ++w) M = fmaxf(M, sm[((size_t)w * G + g) * (2 + D)]);
        float L = 0.f, O = 0.f;
#pragma unroll
        for (int w = 0; w < kAttnWarps; ++w) {
            const float* src = sm + ((size_t)w * G + g) * (2 + D);
            const float a = exp_guard(src[0], M);
            L += src[1] * a;
            O += src[2 + e] * a;
        }
        const int head = kvh * G + g;
        if (splits == 1) {
            reinterpret_cast<T*>(p.out)[((size_t)b * p.head_num + head) * D + e] =
                DT<T>::from_f(O / (L + 1e-6f));
        } else {
            const size_t slot = ((size_t)b * p.head_num + head) * splits + blockIdx.y;
            p.part_o[slot * D + e] = O;
            if (e == 0) { p.part_ml[slot * 2] = M; p.part_ml[slot * 2 + 1] = L; }
        }
    }
}

// ------------------------------------------------------------------------------------------
// The same split-KV kernel over an fp8 (e4m3) cache with one fp32 scale per row (kv_fp8.cu): head size 128, a key row
// is 128 B = eight 16-byte packets of 16 elements, eight lanes own a row, four rows per warp-wide load. q stays fp16;
// a cache element is widened to half exactly (e4m3 is a subset of fp16) and enters the same fp32 FMA; the row scale
// multiplies the finished dot product (K) or the softmax weight (V). No append here: kv_append_fp8 runs before it.
// ------------------------------------------------------------------------------------------
__device__ __forceinline__ uint32_t fp8x2_to_half2(unsigned short v) {
    const __half2_raw h = __nv_cvt_fp8x2_to_halfraw2((__nv_fp8x2_storage_t)v, __NV_E4M3);
    return *reinterpret_cast<const uint32_t*>(&h);
}

// PAGED: kc / vc / ksc / vsc are the POOLS and row `key` of this (sequence, kv head) is row
// (pages[key / 128] * kv_heads + kvh) * 128 + key % 128 of them (data x d, scales x 1)
template <int G, bool FULL, bool PAGED>
__device__ __forceinline__ void fp8_tile(const unsigned char* kc, const unsigned char* vc, const float* ksc, const float* vsc,
                                         const Pack16 (&q)[G][2], float (&m)[G], float (&l)[G], float (&o)[G][16], const int base,
                                         const int k1, const int pi, const int ks, const float scale, const int* pages,
                                         const int kv_heads, const int kvh) {
    constexpr int D = 128, PR = 8, E = 16, KPI = 32 / PR;
        Pack16 kp[kUnroll], vp[kUnroll];
        float ks_[kUnroll], vs_[kUnroll];
        size_t rid[kUnroll];
        bool ok[kUnroll];
#pragma unroll
        for (int u = 0; u < kUnroll; ++u) {
            const int key = FULL ? base + u * KPI + ks : min(base + u * KPI + ks, k1 - 1);   // rows past the end re-read the last row
            ok[u] = FULL || base + u * KPI + ks < k1;
            rid[u] = PAGED ? ((size_t)__ldg(pages + (key >> kPageShift)) * kv_heads + kvh) * kPageRows + (size_t)(key & (kPageRows - 1))
                           : (size_t)key;
            kp[u] = ld16_stream(kc + rid[u] * D + pi * E);
            ks_[u] = __ldg(ksc + rid[u]);
        }
#pragma unroll
        for (int u = 0; u < kUnroll; ++u) {
            vp[u] = ld16_stream(vc + rid[u] * D + pi * E);
            vs_[u] = __ldg(vsc + rid[u]);
        }
        float s[G][kUnroll];
#pragma unroll
        for (int u = 0; u < kUnroll; ++u) {
            uint32_t kh[8];   // the packet as eight half2
#pragma unroll
            for (int i = 0; i < 4; ++i) {
                kh[2 * i] = fp8x2_to_half2((unsigned short)(kp[u].w[i] & 0xffffu));
                kh[2 * i + 1] = fp8x2_to_half2((unsigned short)(kp[u].w[i] >> 16));
            }
#pragma unroll
            for (int g = 0; g < G; ++g) {
                float d = 0.f;
#pragma unroll
                for (int i = 0; i < 8; ++i) {
                    const uint32_t qw = q[g][i >> 2].w[i & 3];
                    d = fma_h((uint16_t)(kh[i] & 0xffffu), (uint16_t)(qw & 0xffffu), d);
                    d = fma_h((uint16_t)(kh[i] >> 16), (uint16_t)(qw >> 16), d);
                }
#pragma unroll
                for (int off = PR >> 1; off > 0; off >>= 1) d += __shfl_xor_sync(0xffffffffu, d, off);
                s[g][u] = (FULL || ok[u]) ? d * ks_[u] * scale : -INFINITY;
            }
        }
#pragma unroll
        for (int g = 0; g < G; ++g) {
            float mn = m[g];
#pragma unroll
            for (int u = 0; u < kUnroll; ++u) mn = fmaxf(mn, s[g][u]);
            if (FULL || mn != -INFINITY) {
                const float corr = exp_guard(m[g], mn);
                l[g] *= corr;
#pragma unroll
                for (int e = 0; e < E; ++e) o[g][e] *= corr;
                m[g] = mn;
#pragma unroll
                for (int u = 0; u < kUnroll; ++u) {
                    if (FULL || ok[u]) {
                        const float pr = __expf(s[g][u] - mn);
                        l[g] += pr;
                        // the softmax weight times the row scale enters the mixed-precision FMA as a half (11 bits: far
                        // below the cache's own 4-bit rounding); the e4m3 element widens to half exactly
                        const uint16_t pv = __half_as_ushort(__float2half_rn(pr * vs_[u]));
#pragma unroll
                        for (int i = 0; i < 4; ++i) {
                            const uint32_t lo = fp8x2_to_half2((unsigned short)(vp[u].w[i] & 0xffffu));
                            const uint32_t hi = fp8x2_to_half2((unsigned short)(vp[u].w[i] >> 16));
                            o[g][4 * i] = fma_h((uint16_t)(lo & 0xffffu), pv, o[g][4 * i]);
                            o[g][4 * i + 1] = fma_h((uint16_t)(lo >> 16), pv, o[g][4 * i + 1]);
                            o[g][4 * i + 2] = fma_h((uint16_t)(hi & 0xffffu), pv, o[g][4 * i + 2]);
                            o[g][4 * i + 3] = fma_h((uint16_t)(hi >> 16), pv, o[g][4 * i + 3]);
                        }
                    }
                }
            }
        }
}

template <int G, bool PAGED = false>
__global__ void __launch_bounds__(kAttnWarps * 32)
decode_attn_split_fp8_kernel(const AttnParams p) {
    constexpr int D = 128, PR = 8, E = 16, KPI = 32 / PR;
    extern __shared__ float sm[];  // [kAttnWarps][G][2 + D]

    pdl_wait();
    const int bh = blockIdx.x;
    const int b = bh / p.kv_head_num, kvh = bh - b * p.kv_head_num;
    const int step = p.step_dev ? p.step_dev[b * p.step_stride] : p.step;
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const int pi = lane % PR, ks = lane / PR;
    const int qkv_heads = p.head_num + 2 * p.kv_head_num;
    const __half* tok = reinterpret_cast<const __half*>(p.qkv) + (size_t)b * qkv_heads * D;
    const size_t head_rows = PAGED ? (size_t)0 : ((size_t)b * p.kv_head_num + kvh) * p.max_seq_len;
    const int* pages = PAGED ? p.page_table + (size_t)b * p.max_pages : nullptr;
    const unsigned char* kc = reinterpret_cast<const unsigned char*>(p.k_cache) + head_rows * D;
    const unsigned char* vc = reinterpret_cast<const unsigned char*>(p.v_cache) + head_rows * D;
    const float* ksc = p.k_scale + head_rows;
    const float* vsc = p.v_scale + head_rows;

    const int splits = gridDim.y;
    const int per = (step + splits - 1) / splits;
    const int k0 = blockIdx.y * per;
    const bool retired = p.finished != nullptr && p.finished[b] != 0;
    const int k1 = retired ? k0 : min(step, k0 + per);

    Pack16 q[G][2];   // the lane's 16 elements of every query head of the group
#pragma unroll
    for (int g = 0; g < G; ++g) {
        q[g][0] = ld16(tok + (size_t)(kvh * G + g) * D + pi * E);
        q[g][1] = ld16(tok + (size_t)(kvh * G + g) * D + pi * E + 8);
    }
    const float scale = rsqrtf((float)D);

    float m[G], l[G], o[G][E];
#pragma unroll
    for (int g = 0; g < G; ++g) {
        m[g] = -INFINITY;
        l[g] = 0.f;
#pragma unroll
        for (int e = 0; e < E; ++e) o[g][e] = 0.f;
    }

    for (int base = k0 + warp * KPI * kUnroll; base < k1; base += kAttnWarps * KPI * kUnroll) {
        // whole tiles (every row of the 16 this warp loads exists) take a branch-free path: no per-row predicates, and
        // the shuffles stay plain SHFLs instead of divergence-safe collectives
        if (base + KPI * kUnroll <= k1) fp8_tile<G, true, PAGED>(kc, vc, ksc, vsc, q, m, l, o, base, k1, pi, ks, scale, pages, p.kv_head_num, kvh);
        else fp8_tile<G, false, PAGED>(kc, vc, ksc, vsc, q, m, l, o, base, k1, pi, ks, scale, pages, p.kv_head_num, kvh);
    }

    // merge the KPI lane groups of the warp
#pragma unroll
    for (int off = PR; off < 32; off <<= 1) {
#pragma unroll
        for (int g = 0; g < G; ++g) {
            const float mo = __shfl_xor_sync(0xffffffffu, m[g], off);
            const float lo = __shfl_xor_sync(0xffffffffu, l[g], off);
            const float M = fmaxf(m[g], mo);
            const float a = exp_guard(m[g], M), bb = exp_guard(mo, M);
            l[g] = l[g] * a + lo * bb;
#pragma unroll
            for (int e = 0; e < E; ++e) {
                const float oo = __shfl_xor_sync(0xffffffffu, o[g][e], off);
                o[g][e] = o[g][e] * a + oo * bb;
            }
            m[g] = M;
        }
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
    __syncthreads();
    pdl_launch_dependents();
    for (int idx = threadIdx.x; idx < G * D; idx += blockDim.x) {
        const int g = idx / D, e = idx - g * D;
        float M = -INFINITY;
#pragma unroll
        for (int w = 0; w < kAttnWarps; ++w) M = fmaxf(M, sm[((size_t)w * G + g) * (2 + D)]);
        float L = 0.f, O = 0.f;
#pragma unroll
        for (int w = 0; w < kAttnWarps; ++w) {
            const float* src = sm + ((size_t)w * G + g) * (2 + D);
            const float a = exp_guard(src[0], M);
            L += src[1] * a;
            O += src[2 + e] * a;
        }
        const int head = kvh * G + g;
        if (splits == 1) {
            reinterpret_cast<__half*>(p.out)[((size_t)b * p.head_num + head) * D + e] = __float2half_rn(O / (L + 1e-6f));
        } else {
            const size_t slot = ((size_t)b * p.head_num + head) * splits + blockIdx.y;
            p.part_o[slot * D + e] = O;
            if (e == 0) { p.part_ml[slot * 2] = M; p.part_ml[slot * 2 + 1] = L; }
        }
    }
}

// merge the per-split partials: grid B*h, one thread per output element
template <typename T>
__global__ void __launch_bounds__(128)
decode_attn_combine_kernel(const float* __restrict__ part_ml, const float* __restrict__ part_o,
                           T* __restrict__ out, int splits, int d) {
    pdl_wait();
    const size_t bh = blockIdx.x;
    const float* ml = part_ml + bh * splits * 2;
    float M = -INFINITY;
    for (int s = 0; s < splits; ++s) M = fmaxf(M, ml[2 * s]);
    float L = 0.f;
    for (int s = 0; s < splits; ++s) L += ml[2 * s + 1] * exp_guard(ml[2 * s], M);
    const float inv = 1.f / (L + 1e-6f);
    pdl_launch_dependents();
    for (int e = threadIdx.x; e < d; e += blockDim.x) {
        float O = 0.f;
        for (int s = 0; s < splits; ++s)
            O += part_o[(bh * splits + s) * d + e] * exp_guard(ml[2 * s], M);
        out[bh * d + e] = DT<T>::from_f(O * inv);
    }
}

// ------------------------------------------------------------------------------------------
// generic path (any head size, e.g. the reference unit test's d = 4): one CTA per (b, head),
// scores staged in a global fp32 scratch row. Not a performance path.
// ------------------------------------------------------------------------------------------
template <typename T>
__global__ void __launch_bounds__(128)
decode_attn_generic_kernel(const AttnParams p, float* __restrict__ scores) {
    __shared__ float scratch[33];
    pdl_wait();
    const int d = p.head_size;
    const int b = blockIdx.x / p.head_num, head = blockIdx.x - b * p.head_num;
    const int step = p.step_dev ? p.step_dev[b * p.step_stride] : p.step;
    const int kvh = head / (p.head_num / p.kv_head_num);
    const int qkv_heads = p.head_num + 2 * p.kv_head_num;
    const T* tok = reinterpret_cast<const T*>(p.qkv) + (size_t)b * qkv_heads * d;
    const T* q = tok + (size_t)head * d;
    const T* knew = tok + (size_t)(p.head_num + kvh) * d;
    const T* vnew = tok + (size_t)(p.head_num + p.kv_head_num + kvh) * d;
    T* kc = reinterpret_cast<T*>(p.k_cache) + ((size_t)b * p.kv_head_num + kvh) * p.max_seq_len * d;
    T* vc = reinterpret_cast<T*>(p.v_cache) + ((size_t)b * p.kv_head_num + kvh) * p.max_seq_len * d;
    float* sc = scores + (size_t)blockIdx.x * p.max_seq_len;
    const float scale = rsqrtf((float)d);
    const bool use_new = p.append_kv != 0;
    if (p.finished != nullptr && p.finished[b] != 0) {   // retired sequence: no append, no reads, zero output
        for (int e = threadIdx.x; e < d; e += blockDim.x)
            reinterpret_cast<T*>(p.out)[((size_t)b * p.head_num + head) * d + e] = DT<T>::from_f(0.f);
        pdl_launch_dependents();
        return;
    }

    float mx = -INFINITY;
    for (int j = threadIdx.x; j < step; j += blockDim.x) {
        const T* kr = (use_new && j == step - 1) ? knew : kc + (size_t)j * d;
        float acc = 0.f;
        for (int e = 0; e < d; ++e) acc = fmaf(DT<T>::to_f(q[e]), DT<T>::to_f(kr[e]), acc);
        acc *= scale;
        sc[j] = acc;
        mx = fmaxf(mx, acc);
    }
    mx = block_max(mx, scratch);
    float sum = 0.f;
    for (int j = threadIdx.x; j < step; j += blockDim.x) {
        const float e = expf(sc[j] - mx);
        sc[j] = e;
        sum += e;
    }
    sum = block_sum(sum, scratch);  // barriers inside also publish sc[] to the whole CTA
    const float inv = 1.f / (sum + 1e-6f);
    for (int e = threadIdx.x; e < d; e += blockDim.x) {
        float acc = 0.f;
        for (int j = 0; j < step; ++j) {
            const T* vr = (use_new && j == step - 1) ? vnew : vc + (size_t)j * d;
            acc = fmaf(sc[j], DT<T>::to_f(vr[e]), acc);
        }
        reinterpret_cast<T*>(p.out)[((size_t)b * p.head_num + head) * d + e] = DT<T>::from_f(acc * inv);
    }
    // append after all reads of this CTA; only the first q head of the kv group writes
    if (use_new && head % (p.head_num / p.kv_head_num) == 0) {
        for (int e = threadIdx.x; e < d; e += blockDim.x) {
            kc[(size_t)(step - 1) * d + e] = knew[e];
            vc[(size_t)(step - 1) * d + e] = vnew[e];
        }
    }
    pdl_launch_dependents();
}

// ------------------------------------------------------------------------------------------
// host side
// ------------------------------------------------------------------------------------------
static int attn_pick_splits(int batch, int kv_head_num, int step, int group) {
    const int ctas = batch * kv_head_num;
    const int target = 2 * sm_count();
    int splits = (target + ctas - 1) / ctas;
    // keep >= 64 (key, query head) pairs per split, at least 16 keys: with GQA every key row serves `group`
    // query heads, and tensor-parallel shards leave few kv heads per rank (70B tp8: ONE) to spread over the SMs
    int min_keys = 64 / (group < 1 ? 1 : group);
    if (min_keys < 16) min_keys = 16;
    const int max_by_len = (step + min_keys - 1) / min_keys;
    if (splits > max_by_len) splits = max_by_len;
    if (splits < 1) splits = 1;
    if (splits > 64) splits = 64;
    return splits;
}

const unsigned char* attn_finished_ptr() { return g_attn_finished; }   // kv_fp8.cu's append honours the same flags

int attn_pick_splits_public(int batch, int kv_head_num, int step, int group) {
    return attn_pick_splits(batch, kv_head_num, step, group);
}

static bool attn_fast_shape(int head_num, int kv_head_num, int head_size, int esz) {
    const int G = head_num / kv_head_num;
    const int row = head_size * esz;
    const bool g_ok = G == 1 || G == 2 || G == 4 || G == 8;
    const bool pr_ok = row == 128 || row == 256 || row == 512;
    return g_ok && pr_ok && head_num % kv_head_num == 0;
}

template <typename T, int G, int PR>
static cudaError_t attn_launch_split(const AttnParams& p, int splits, cudaStream_t s) {
    constexpr int D = PR * DT<T>::kPer16B;
    const size_t smem = (size_t)kAttnWarps * G * (2 + D) * sizeof(float);
    return launch_k(decode_attn_split_kernel<T, G, PR>, dim3(p.batch * p.kv_head_num, splits),
                    dim3(kAttnWarps * 32), smem, s, true, p);
}

template <typename T>
static int attn_dispatch(const AttnParams& p, int splits, cudaStream_t s) {
    const int G = p.head_num / p.kv_head_num;
    const int PR = p.head_size * (int)sizeof(T) / 16;
    cudaError_t e = cudaErrorInvalidValue;
#define CASE(g, pr) \
    if (G == g && PR == pr) e = attn_launch_split<T, g, pr>(p, splits, s);
    CASE(1, 8) CASE(1, 16) CASE(1, 32) CASE(2, 8) CASE(2, 16) CASE(2, 32)
    CASE(4, 8) CASE(4, 16) CASE(4, 32) CASE(8, 8) CASE(8, 16) CASE(8, 32)
#undef CASE
    count_launch();
    if (e != cudaSuccess) return (int)e;
    if (splits > 1) {
        e = launch_k(decode_attn_combine_kernel<T>, dim3(p.batch * p.head_num), dim3(128), 0, s,
                     true, (const float*)p.part_ml, (const float*)p.part_o, (T*)p.out, splits,
                     p.head_size);
        count_launch();
    }
    return (int)e;
}

}  // namespace b200

using namespace b200;

extern "C" size_t b200_decode_attention_workspace_bytes(int batch, int head_num, int head_size,
                                                        int step) {
    if (batch <= 0 || head_num <= 0 || head_size <= 0 || step <= 0) return 0;
    // split partials (worst case 64 splits) or the generic path's score rows (max_seq <= step
    // is not known here, so the generic path sizes its rows by `step` rounded up by the caller:
    // callers pass step = max_seq_len when they want one buffer for the whole generation).
    const size_t part = (size_t)batch * head_num * 64 * (2 + (size_t)head_size) * sizeof(float);
    const size_t generic = (size_t)batch * head_num * (size_t)step * sizeof(float);
    return (part > generic ? part : generic) + 256;
}

static int decode_attention_impl(const void* qkv, void* k_cache_layer, void* v_cache_layer,
                                 void* out, void* workspace, size_t workspace_bytes,
                                 int batch, int head_num, int kv_head_num, int head_size,
                                 int max_seq_len, int step, const int* step_dev, int step_stride,
                                 int append_kv, int dtype, void* stream) {
    B200_RETURN_IF(!qkv || !k_cache_layer || !v_cache_layer || !out, B200_EINVAL);
    B200_RETURN_IF(batch <= 0 || head_num <= 0 || kv_head_num <= 0 || head_size <= 0 ||
                       max_seq_len <= 0 || head_num % kv_head_num != 0,
                   B200_EINVAL);
    B200_RETURN_IF(step < 1 || step > max_seq_len, B200_EINVAL);
    B200_RETURN_IF(dtype != B200_F32 && dtype != B200_F16, B200_EINVAL);
    cudaStream_t s = (cudaStream_t)stream;
    const int esz = dtype == B200_F16 ? 2 : 4;
    AttnParams p{};
    p.qkv = qkv; p.k_cache = k_cache_layer; p.v_cache = v_cache_layer; p.out = out;
    p.batch = batch; p.head_num = head_num; p.kv_head_num = kv_head_num; p.head_size = head_size;
    p.max_seq_len = max_seq_len; p.step = step; p.step_dev = step_dev; p.step_stride = step_stride; p.append_kv = append_kv;
    p.finished = g_attn_finished;
    auto al = [](const void* q) { return (reinterpret_cast<uintptr_t>(q) & 15u) == 0; };
    const bool fast = attn_fast_shape(head_num, kv_head_num, head_size, esz) && al(qkv) &&
                      al(k_cache_layer) && al(v_cache_layer);
    if (fast) {
        const int splits = attn_pick_splits(batch, kv_head_num, step, head_num / kv_head_num);
        if (splits > 1) {
            const size_t ml = (size_t)batch * head_num * splits * 2 * sizeof(float);
            const size_t need = ((ml + 255) & ~(size_t)255) +
                                (size_t)batch * head_num * splits * head_size * sizeof(float);
            B200_RETURN_IF(!workspace || workspace_bytes < need, B200_EWORKSPACE);
            p.part_ml = reinterpret_cast<float*>(workspace);
            p.part_o = reinterpret_cast<float*>((char*)workspace + ((ml + 255) & ~(size_t)255));
        }
        return dtype == B200_F16 ? attn_dispatch<__half>(p, splits, s)
                                 : attn_dispatch<float>(p, splits, s);
    }
    const size_t need = (size_t)batch * head_num * (size_t)max_seq_len * sizeof(float);
    B200_RETURN_IF(!workspace || workspace_bytes < need, B200_EWORKSPACE);
    cudaError_t e;
    if (dtype == B200_F16)
        e = launch_k(decode_attn_generic_kernel<__half>, dim3(batch * head_num), dim3(128), 0, s,
                     true, p, (float*)workspace);
    else
        e = launch_k(decode_attn_generic_kernel<float>, dim3(batch * head_num), dim3(128), 0, s,
                     true, p, (float*)workspace);
    count_launch();
    return (int)e;
}

extern "C" int b200_decode_attention_ex(const void* qkv, void* k_cache_layer, void* v_cache_layer,
                                        void* out, void* workspace, size_t workspace_bytes,
                                        int batch, int head_num, int kv_head_num, int head_size,
                                        int max_seq_len, int step, const int* step_dev,
                                        int append_kv, int dtype, void* stream) {
    return decode_attention_impl(qkv, k_cache_layer, v_cache_layer, out, workspace, workspace_bytes, batch,
                                 head_num, kv_head_num, head_size, max_seq_len, step, step_dev, 0, append_kv,
                                 dtype, stream);
}

/* ragged batch: sequence b attends over seq_steps_dev[b] keys (its new token at row seq_steps_dev[b]-1);
 * max_step >= every entry sizes the split count and the workspace, nothing else */
extern "C" int b200_decode_attention_ragged(const void* qkv, void* k_cache_layer, void* v_cache_layer,
                                            void* out, void* workspace, size_t workspace_bytes,
                                            int batch, int head_num, int kv_head_num, int head_size,
                                            int max_seq_len, int max_step, const int* seq_steps_dev,
                                            int append_kv, int dtype, void* stream) {
    B200_RETURN_IF(!seq_steps_dev, B200_EINVAL);
    return decode_attention_impl(qkv, k_cache_layer, v_cache_layer, out, workspace, workspace_bytes, batch,
                                 head_num, kv_head_num, head_size, max_seq_len, max_step, seq_steps_dev, 1,
                                 append_kv, dtype, stream);
}

extern "C" int b200_decode_attention(const void* qkv, void* k_cache_layer, void* v_cache_layer,
                                     void* out, void* workspace, size_t workspace_bytes, int batch,
                                     int head_num, int kv_head_num, int head_size, int max_seq_len,
                                     int step, int append_kv, int dtype, void* stream) {
    return b200_decode_attention_ex(qkv, k_cache_layer, v_cache_layer, out, workspace,
                                    workspace_bytes, batch, head_num, kv_head_num, head_size,
                                    max_seq_len, step, nullptr, append_kv, dtype, stream);
}

// ------------------------------------------------------------------------------------------
// Paged cache (SURVEY 8 f4 "paged blocks"): the reference's dense [layers, B, hk, max_seq, d] (llama.cpp:77-78) reserves
// max_seq rows for every sequence slot; here a layer's K (V) is a pool of pages [n_pages, hk, 128 rows, d] and a
// sequence owns the pages its page list names ([batch, max_pages] int, shared by all layers). Same kernel, same
// split / combine, same semantics as b200_decode_attention_ragged; the only difference is the row address. fp16,
// head size 128 (a 256-byte row), head_num / kv_head_num in {1, 2, 4, 8}.
// ------------------------------------------------------------------------------------------
extern "C" int b200_decode_attention_paged(const void* qkv, void* k_pool_layer, void* v_pool_layer, const int* page_table,
                                           int max_pages, void* out, void* workspace, size_t workspace_bytes, int batch,
                                           int head_num, int kv_head_num, int head_size, int max_step,
                                           const int* seq_steps_dev, int step_stride, int append_kv, void* stream) {
    B200_RETURN_IF(!qkv || !k_pool_layer || !v_pool_layer || !page_table || !out, B200_EINVAL);
    B200_RETURN_IF(batch <= 0 || head_num <= 0 || kv_head_num <= 0 || max_pages <= 0 || head_num % kv_head_num != 0, B200_EINVAL);
    B200_RETURN_IF(max_step < 1 || max_step > max_pages * kPageRows, B200_EINVAL);
    B200_RETURN_IF(step_stride != 0 && step_stride != 1, B200_EINVAL);
    const int G = head_num / kv_head_num;
    B200_RETURN_IF(head_size != 128 || (G != 1 && G != 2 && G != 4 && G != 8), B200_EUNSUPPORTED);
    auto al = [](const void* q) { return (reinterpret_cast<uintptr_t>(q) & 15u) == 0; };
    B200_RETURN_IF(!al(qkv) || !al(k_pool_layer) || !al(v_pool_layer), B200_EINVAL);
    cudaStream_t s = (cudaStream_t)stream;
    AttnParams p{};
    p.qkv = qkv; p.k_cache = k_pool_layer; p.v_cache = v_pool_layer; p.out = out;
    p.batch = batch; p.head_num = head_num; p.kv_head_num = kv_head_num; p.head_size = head_size;
    p.max_seq_len = max_pages * kPageRows; p.step = max_step; p.step_dev = seq_steps_dev; p.step_stride = step_stride;
    p.append_kv = append_kv;
    p.finished = g_attn_finished;
    p.page_table = page_table; p.max_pages = max_pages;
    const int splits = attn_pick_splits(batch, kv_head_num, max_step, G);
    if (splits > 1) {
        const size_t ml = (size_t)batch * head_num * splits * 2 * sizeof(float);
        const size_t need = ((ml + 255) & ~(size_t)255) + (size_t)batch * head_num * splits * head_size * sizeof(float);
        B200_RETURN_IF(!workspace || workspace_bytes < need, B200_EWORKSPACE);
        p.part_ml = reinterpret_cast<float*>(workspace);
        p.part_o = reinterpret_cast<float*>((char*)workspace + ((ml + 255) & ~(size_t)255));
    }
    constexpr int PR = 16, D = 128;
    cudaError_t e = cudaErrorInvalidValue;
#define CASE(g)                                                                                                   \
    if (G == g)                                                                                                   \
        e = launch_k(decode_attn_split_kernel<__half, g, PR, true>, dim3(batch * kv_head_num, splits),            \
                     dim3(kAttnWarps * 32), (size_t)kAttnWarps * g * (2 + D) * sizeof(float), s, true, p);
    CASE(1) CASE(2) CASE(4) CASE(8)
#undef CASE
    count_launch();
    if (e != cudaSuccess) return (int)e;
    if (splits > 1) {
        e = launch_k(decode_attn_combine_kernel<__half>, dim3(batch * head_num), dim3(128), 0, s, true,
                     (const float*)p.part_ml, (const float*)p.part_o, (__half*)p.out, splits, head_size);
        count_launch();
    }
    return (int)e;
}

// ------------------------------------------------------------------------------------------
// Allocation-free variant (no counterpart in the reference; added
// for callers that cannot call cudaMalloc: CUDA-graph capture, tight serving loops).
// The split-KV partials / generic score rows live in a file-scope device array instead of a
// caller-supplied workspace. Every GPU that loads the library holds the array until exit.
//  * The array's name is only a host-side stand-in in host code: its device address is looked up
//    once per device, with that device current, by cudaGetSymbolAddress, and that pointer is what
//    the kernels get.
//  * Nothing here (or below it) calls cudaMalloc / cudaFree or an allocating library routine.
//  * All calls share the array. Nothing in it has to start at zero: every (head, split) partial
//    the combine kernel reads is written by the split kernel of the same call first (an empty split
//    writes m = -inf, l = 0), and the generic path writes each score row before it reads it. Two
//    calls that overlap on the GPU (different streams, no dependency) race on it.
// ------------------------------------------------------------------------------------------
#include <cstdio>
static __device__ __align__(256) unsigned char g_attn_static_scratch[B200_DECODE_ATTENTION_STATIC_MAX_BYTES];

namespace {
constexpr int kMaxStaticDevices = 64;
std::mutex g_attn_static_mu;
void* g_attn_static_ptr[kMaxStaticDevices] = {};

// device address of g_attn_static_scratch on the CURRENT device (looked up once per device)
int attn_static_scratch(void** out) {
    int dev = -1;
    cudaError_t e = cudaGetDevice(&dev);
    if (e != cudaSuccess) return (int)e;
    if (dev < 0 || dev >= kMaxStaticDevices) return B200_EUNSUPPORTED;
    std::lock_guard<std::mutex> g(g_attn_static_mu);
    if (!g_attn_static_ptr[dev]) {
        void* ptr = nullptr;
        e = cudaGetSymbolAddress(&ptr, g_attn_static_scratch);
        if (e != cudaSuccess) return (int)e;
        if (!ptr) return (int)cudaErrorInvalidSymbol;
        g_attn_static_ptr[dev] = ptr;
    }
    *out = g_attn_static_ptr[dev];
    return 0;
}
}  // namespace

extern "C" int b200_decode_attention_static(const void* qkv, void* k_cache_layer, void* v_cache_layer,
                                            void* out, int batch, int head_num, int kv_head_num,
                                            int head_size, int max_seq_len, int step, int append_kv,
                                            int dtype, void* stream) {
    B200_RETURN_IF(batch <= 0 || head_num <= 0 || kv_head_num <= 0 || head_size <= 0 || max_seq_len <= 0 ||
                       head_num % kv_head_num != 0 || step < 1 || step > max_seq_len,
                   B200_EINVAL);
    B200_RETURN_IF(dtype != B200_F32 && dtype != B200_F16, B200_EINVAL);
    // scratch this shape needs: the same two formulas decode_attention_impl checks its workspace against
    const int esz = dtype == B200_F16 ? 2 : 4;
    size_t need;
    if (attn_fast_shape(head_num, kv_head_num, head_size, esz)) {
        const int splits = attn_pick_splits(batch, kv_head_num, step, head_num / kv_head_num);
        const size_t ml = (size_t)batch * head_num * splits * 2 * sizeof(float);
        need = splits > 1 ? ((ml + 255) & ~(size_t)255) + (size_t)batch * head_num * splits * head_size * sizeof(float)
                          : 0;
        // a misaligned pointer sends a fast shape down the generic path
        const size_t generic = (size_t)batch * head_num * (size_t)max_seq_len * sizeof(float);
        auto al = [](const void* q) { return (reinterpret_cast<uintptr_t>(q) & 15u) == 0; };
        if (!(al(qkv) && al(k_cache_layer) && al(v_cache_layer))) need = generic;
    } else {
        need = (size_t)batch * head_num * (size_t)max_seq_len * sizeof(float);
    }
    if (need > (size_t)B200_DECODE_ATTENTION_STATIC_MAX_BYTES) {
        fprintf(stderr,
                "[b200][ERROR] b200_decode_attention_static: this shape needs %zu bytes of scratch, the built-in "
                "array holds %zu (B200_DECODE_ATTENTION_STATIC_MAX_BYTES); use b200_decode_attention with a "
                "caller-owned workspace\n",
                need, (size_t)B200_DECODE_ATTENTION_STATIC_MAX_BYTES);
        return B200_EWORKSPACE;
    }
    void* scratch = nullptr;
    const int rc = attn_static_scratch(&scratch);
    if (rc != 0) {
        fprintf(stderr, "[b200][ERROR] b200_decode_attention_static: cudaGetSymbolAddress failed (%d)\n", rc);
        return rc;
    }
    return decode_attention_impl(qkv, k_cache_layer, v_cache_layer, out, scratch,
                                 (size_t)B200_DECODE_ATTENTION_STATIC_MAX_BYTES, batch, head_num, kv_head_num,
                                 head_size, max_seq_len, step, nullptr, 0, append_kv, dtype, stream);
}

// ------------------------------------------------------------------------------------------
// fp8 KV cache (kv_fp8.cu): attention over rows [0, step) of an e4m3 cache with one fp32 scale per row. fp16 q / out,
// head size 128, h/hk in {1, 2, 4}. The new token's row must already be in the cache (b200_kv_append_fp8).
// ------------------------------------------------------------------------------------------
static int decode_attention_fp8_impl(const void* qkv, const void* k_cache8, const void* v_cache8, const float* k_scale,
                                     const float* v_scale, void* out, void* workspace, size_t workspace_bytes,
                                     int batch, int head_num, int kv_head_num, int head_size, int max_seq_len,
                                     int step, const int* step_dev, int step_stride, const int* page_table, int max_pages,
                                     void* stream) {
    B200_RETURN_IF(!qkv || !k_cache8 || !v_cache8 || !k_scale || !v_scale || !out, B200_EINVAL);
    B200_RETURN_IF(batch <= 0 || head_num <= 0 || kv_head_num <= 0 || max_seq_len <= 0 || head_num % kv_head_num != 0,
                   B200_EINVAL);
    B200_RETURN_IF(step < 1 || step > max_seq_len, B200_EINVAL);
    const int G = head_num / kv_head_num;
    B200_RETURN_IF(head_size != 128 || (G != 1 && G != 2 && G != 4), B200_EUNSUPPORTED);
    auto al = [](const void* q) { return (reinterpret_cast<uintptr_t>(q) & 15u) == 0; };
    B200_RETURN_IF(!al(qkv) || !al(k_cache8) || !al(v_cache8), B200_EINVAL);
    AttnParams p{};
    p.qkv = qkv; p.k_cache = const_cast<void*>(k_cache8); p.v_cache = const_cast<void*>(v_cache8); p.out = out;
    p.k_scale = k_scale; p.v_scale = v_scale;
    p.batch = batch; p.head_num = head_num; p.kv_head_num = kv_head_num; p.head_size = head_size;
    p.max_seq_len = max_seq_len; p.step = step; p.step_dev = step_dev; p.step_stride = step_stride; p.append_kv = 0;
    p.finished = g_attn_finished;
    p.page_table = page_table; p.max_pages = max_pages;
    const int splits = attn_pick_splits(batch, kv_head_num, step, G);
    if (splits > 1) {
        const size_t ml = (size_t)batch * head_num * splits * 2 * sizeof(float);
        const size_t need = ((ml + 255) & ~(size_t)255) + (size_t)batch * head_num * splits * head_size * sizeof(float);
        B200_RETURN_IF(!workspace || workspace_bytes < need, B200_EWORKSPACE);
        p.part_ml = reinterpret_cast<float*>(workspace);
        p.part_o = reinterpret_cast<float*>((char*)workspace + ((ml + 255) & ~(size_t)255));
    }
    cudaStream_t s = (cudaStream_t)stream;
    const dim3 grid(batch * kv_head_num, splits), block(kAttnWarps * 32);
    cudaError_t e;
    const size_t sm1 = (size_t)kAttnWarps * (2 + 128) * sizeof(float);
    if (page_table) {
        if (G == 1) e = launch_k(decode_attn_split_fp8_kernel<1, true>, grid, block, sm1, s, true, p);
        else if (G == 2) e = launch_k(decode_attn_split_fp8_kernel<2, true>, grid, block, 2 * sm1, s, true, p);
        else e = launch_k(decode_attn_split_fp8_kernel<4, true>, grid, block, 4 * sm1, s, true, p);
    } else if (G == 1) e = launch_k(decode_attn_split_fp8_kernel<1>, grid, block, sm1, s, true, p);
    else if (G == 2) e = launch_k(decode_attn_split_fp8_kernel<2>, grid, block, 2 * sm1, s, true, p);
    else e = launch_k(decode_attn_split_fp8_kernel<4>, grid, block, 4 * sm1, s, true, p);
    count_launch();
    if (e != cudaSuccess) return (int)e;
    if (splits > 1) {
        e = launch_k(decode_attn_combine_kernel<__half>, dim3(batch * head_num), dim3(128), 0, s, true,
                     (const float*)p.part_ml, (const float*)p.part_o, (__half*)p.out, splits, head_size);
        count_launch();
    }
    return (int)e;
}

extern "C" int b200_decode_attention_fp8(const void* qkv, const void* k_cache8, const void* v_cache8, const float* k_scale,
                                         const float* v_scale, void* out, void* workspace, size_t workspace_bytes,
                                         int batch, int head_num, int kv_head_num, int head_size, int max_seq_len,
                                         int step, const int* step_dev, int step_stride, void* stream) {
    return decode_attention_fp8_impl(qkv, k_cache8, v_cache8, k_scale, v_scale, out, workspace, workspace_bytes, batch, head_num,
                                     kv_head_num, head_size, max_seq_len, step, step_dev, step_stride, nullptr, 0, stream);
}

/* the same over a paged fp8 cache (pools [n_pages, kv_heads, 128, d] + scale pools [n_pages, kv_heads, 128], page lists
 * DEVICE int [batch, max_pages]); the new row must already be in its page (b200_kv_append_fp8_paged) */
extern "C" int b200_decode_attention_fp8_paged(const void* qkv, const void* k_pool8, const void* v_pool8, const float* k_scale_pool,
                                               const float* v_scale_pool, const int* page_table, int max_pages, void* out,
                                               void* workspace, size_t workspace_bytes, int batch, int head_num,
                                               int kv_head_num, int head_size, int step, const int* step_dev, int step_stride,
                                               void* stream) {
    B200_RETURN_IF(!page_table || max_pages <= 0, B200_EINVAL);
    return decode_attention_fp8_impl(qkv, k_pool8, v_pool8, k_scale_pool, v_scale_pool, out, workspace, workspace_bytes, batch,
                                     head_num, kv_head_num, head_size, max_pages * kPageRows, step, step_dev, step_stride,
                                     page_table, max_pages, stream);
}

/* DEVICE flags [batch] (0 / 1) applied to every decode-attention launch issued afterwards from this host thread; NULL
 * switches it off. A set flag retires the sequence: no K/V append, no cache reads, zero output row. */
extern "C" void b200_decode_attention_set_finished(const unsigned char* finished_dev) { g_attn_finished = finished_dev; }
