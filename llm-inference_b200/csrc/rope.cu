// Rotate-half RoPE: decode (in place on the newest token) and prefill
// (bias + RoPE + un-pad/transposition into q/k/v buffers).
// Reference math: src/kernels/qkv_bias_and_RoPE.cu:29-46 (freq + rotation), :323-380 (decode),
// :50-145 (prefill; SURVEY defects D5, D6, D12 resolved: token-major qkv with stride
// (h+2hk)*d, v is copied, timestep = history_len[b] + local token id).
//   theta_i = pos / base^(2i/rot_dim), i < rot_dim/2; pair (i, i + head_size/2)
//   (x_i, x_j) <- (x_i cos - x_j sin, x_j cos + x_i sin)
#include "common.cuh"
#include "launch_count.h"

namespace b200 {

__device__ __forceinline__ float2 rope_cos_sin(int pair, int rot_dim, float base, float pos) {
    const float ang = rope_angle(pair, rot_dim, base, pos);
    return make_float2(cosf(ang), sinf(ang));
}

// ---------------------------------------------------------------- decode
// one CTA per sequence; the cos/sin table of the (shared) position is built once in smem.
template <typename T>
__global__ void __launch_bounds__(256)
rope_decode_kernel(T* __restrict__ qkv, int head_num, int kv_head_num, int head_size, int pos,
                   const int* __restrict__ step_dev, int rot_dim, float base, const float* __restrict__ den,
                   int step_stride) {
    constexpr int E = DT<T>::kPer16B;
    extern __shared__ float2 cs[];  // [head_size/2]; identity beyond rot_dim/2
    const int half_rot = rot_dim >> 1, half_d = head_size >> 1;
    // step_dev (CUDA-graph replay) is only ever written by the LAST kernel of the previous step,
    // which has fully completed before any kernel of this step starts, so it may be read early
    // step_stride 1: step_dev[b] is sequence b's own step (ragged batch)
    if (step_dev) pos = step_dev[blockIdx.x * step_stride] - 1;
    for (int i = threadIdx.x; i < half_d; i += blockDim.x) {
        if (i >= half_rot) { cs[i] = make_float2(1.f, 0.f); continue; }
        // den[i] = base^(2i/rot_dim) from b200_rope_table(): the same correctly rounded float rope_angle() forms
        const float ang = den ? (float)pos / den[i] : rope_angle(i, rot_dim, base, (float)pos);
        cs[i] = make_float2(cosf(ang), sinf(ang));
    }
    pdl_wait();
    __syncthreads();
    const int qkv_heads = head_num + 2 * kv_head_num;
    T* tok = qkv + (size_t)blockIdx.x * qkv_heads * head_size;
    if ((half_d % E) == 0 && (reinterpret_cast<uintptr_t>(qkv) & 15u) == 0) {
        // 16-byte packets: item = packet p of the low half + packet p of the high half of one q / k head
        const int pk_half = half_d / E;
        const int items = (head_num + kv_head_num) * pk_half;
        for (int it = threadIdx.x; it < items; it += blockDim.x) {
            const int head = it / pk_half, pkt = it - head * pk_half;
            T* p = tok + (size_t)head * head_size + pkt * E;
            float lo[E], hi[E];
            Unpack<T>::get(ld16(p), lo);
            Unpack<T>::get(ld16(p + half_d), hi);
#pragma unroll
            for (int i = 0; i < E; ++i) {
                const float2 c = cs[pkt * E + i];
                const float a = lo[i], b = hi[i];
                lo[i] = c.x * a - c.y * b;
                hi[i] = c.x * b + c.y * a;
            }
            st16(p, Unpack<T>::put(lo));
            st16(p + half_d, Unpack<T>::put(hi));
        }
    } else {
        const int items = (head_num + kv_head_num) * half_rot;  // q heads then k heads are contiguous
        for (int it = threadIdx.x; it < items; it += blockDim.x) {
            const int head = it / half_rot, i = it - head * half_rot;
            T* p = tok + (size_t)head * head_size + i;
            const float a = DT<T>::to_f(p[0]), b = DT<T>::to_f(p[half_d]);
            const float2 c = cs[i];
            p[0] = DT<T>::from_f(c.x * a - c.y * b);
            p[half_d] = DT<T>::from_f(c.x * b + c.y * a);
        }
    }
    pdl_launch_dependents();
}

// ---------------------------------------------------------------- prefill
// one CTA per (unpadded) token. 16-byte packets; a pair-item owns packet p of the low half and
// packet p of the high half of one q/k head, a copy-item owns one packet of a v head.
template <typename T, bool kBias>
__global__ void __launch_bounds__(256)
qkv_rope_transpose_kernel(T* __restrict__ q, T* __restrict__ k, T* __restrict__ v,
                          const T* __restrict__ qkv, const T* __restrict__ bias,
                          const int* __restrict__ padding_offset,
                          const int* __restrict__ history_len, int seq_len, int head_num,
                          int kv_head_num, int head_size, int rot_dim, float base, int kv_rows, int to_cache,
                          const int* __restrict__ page_table, int max_pages) {
    constexpr int E = DT<T>::kPer16B;
    extern __shared__ float2 cs[];  // [head_size/2]; identity beyond rot_dim/2
    pdl_wait();
    const int token = blockIdx.x;
    const int dst_token = token + padding_offset[token];
    const int b = dst_token / seq_len;
    const int local = dst_token - b * seq_len;
    const int pos = history_len[b] + local;
    // k / v destination rows: the padded [B, hk, seq_len, d] buffers of the reference, or (to_cache) straight
    // the KV-cache layer slice [B, hk, kv_rows = max_seq, d] at row history + local, which is where
    // launchConcatKVCache (concat_past_kv.cu:17-56) would copy them next
    const int kv_local = to_cache ? pos : local;
    // to_cache: a row beyond the cache slice (history + length > max_seq_len, a caller error the launcher cannot see:
    // the lengths live on the device) is dropped, never written outside the slice
    const bool kv_ok = kv_local >= 0 && kv_local < kv_rows;
    // paged cache (page_table != NULL, to_cache): k / v are page pools [n_pages, hk, 128, d]; the row lives in the page the
    // sequence's list names for it. kv_head_row(h) = index of (kv head h, this token's row) among the rows of k / v.
    const int page = (page_table != nullptr && kv_ok) ? page_table[(size_t)b * max_pages + kv_local / B200_KV_PAGE_ROWS] : 0;
    auto kv_head_row = [&](int h) -> size_t {
        if (page_table != nullptr) return ((size_t)page * kv_head_num + h) * B200_KV_PAGE_ROWS + (size_t)(kv_local % B200_KV_PAGE_ROWS);
        return ((size_t)b * kv_head_num + h) * kv_rows + (size_t)kv_local;
    };
    const int half_d = head_size >> 1, half_rot = rot_dim >> 1;
    for (int i = threadIdx.x; i < half_d; i += blockDim.x)
        cs[i] = (i < half_rot) ? rope_cos_sin(i, rot_dim, base, (float)pos) : make_float2(1.f, 0.f);
    __syncthreads();

    const int qkv_heads = head_num + 2 * kv_head_num;
    const T* src_tok = qkv + (size_t)token * qkv_heads * head_size;
    const int pk_half = half_d / E;          // packets per half head
    const int pair_items = (head_num + kv_head_num) * pk_half;
    const int copy_items = kv_head_num * (head_size / E);

    for (int it = threadIdx.x; it < pair_items + copy_items; it += blockDim.x) {
        if (it < pair_items) {
            const int head = it / pk_half, p = it - head * pk_half;
            const T* s = src_tok + (size_t)head * head_size + p * E;
            float lo[E], hi[E];
            Unpack<T>::get(ld16_stream(s), lo);
            Unpack<T>::get(ld16_stream(s + half_d), hi);
            if (kBias) {
                float bl[E], bh[E];
                Unpack<T>::get(ld16(bias + (size_t)head * head_size + p * E), bl);
                Unpack<T>::get(ld16(bias + (size_t)head * head_size + half_d + p * E), bh);
#pragma unroll
                for (int i = 0; i < E; ++i) { lo[i] += bl[i]; hi[i] += bh[i]; }
            }
#pragma unroll
            for (int i = 0; i < E; ++i) {
                const float2 c = cs[p * E + i];
                const float a = lo[i], bb = hi[i];
                lo[i] = c.x * a - c.y * bb;
                hi[i] = c.x * bb + c.y * a;
            }
            T* d;
            if (head < head_num)
                d = q + (((size_t)b * head_num + head) * seq_len + local) * head_size + p * E;
            else if (kv_ok)
                d = k + kv_head_row(head - head_num) * head_size + p * E;
            else
                continue;
            st16(d, Unpack<T>::put(lo));
            st16(d + half_d, Unpack<T>::put(hi));
        } else {
            const int c = it - pair_items;
            const int pk_head = head_size / E;
            const int head = c / pk_head, p = c - head * pk_head;
            const size_t off = (size_t)(head_num + kv_head_num + head) * head_size + p * E;
            Pack16 val = ld16_stream(src_tok + off);
            if (kBias) {
                float f[E], bf[E];
                Unpack<T>::get(val, f);
                Unpack<T>::get(ld16(bias + off), bf);
#pragma unroll
                for (int i = 0; i < E; ++i) f[i] += bf[i];
                val = Unpack<T>::put(f);
            }
            if (kv_ok) st16(v + kv_head_row(head) * head_size + p * E, val);
        }
    }
    pdl_launch_dependents();
}

template <typename T>
static int rope_decode_impl(void* qkv, int batch, int head_num, int kv_head_num, int head_size,
                            int step, const int* step_dev, int rot_dim, float base, cudaStream_t s,
                            const float* den = nullptr, int step_stride = 0) {
    cudaError_t e = launch_k(rope_decode_kernel<T>, dim3(batch), dim3(256),
                             (size_t)(head_size / 2) * sizeof(float2), s, true, (T*)qkv, head_num,
                             kv_head_num, head_size, step - 1, step_dev, rot_dim, base, den, step_stride);
    count_launch();
    return (int)e;
}

template <typename T>
static int qkv_rope_impl(void* q, void* k, void* v, const void* qkv, const void* bias,
                         const int* padding_offset, const int* history_len, int seq_len,
                         int token_num, int head_num, int kv_head_num, int head_size, int rot_dim,
                         float base, cudaStream_t s, int kv_rows = 0, int to_cache = 0, const int* page_table = nullptr,
                         int max_pages = 0) {
    constexpr int E = DT<T>::kPer16B;
    if (!to_cache) kv_rows = seq_len;
    B200_RETURN_IF((head_size / 2) % E != 0, B200_EUNSUPPORTED);
    const size_t smem = (size_t)(head_size / 2) * sizeof(float2);
    cudaError_t e;
    if (bias)
        e = launch_k(qkv_rope_transpose_kernel<T, true>, dim3(token_num), dim3(256), smem, s, true,
                     (T*)q, (T*)k, (T*)v, (const T*)qkv, (const T*)bias, padding_offset,
                     history_len, seq_len, head_num, kv_head_num, head_size, rot_dim, base, kv_rows, to_cache, page_table, max_pages);
    else
        e = launch_k(qkv_rope_transpose_kernel<T, false>, dim3(token_num), dim3(256), smem, s,
                     true, (T*)q, (T*)k, (T*)v, (const T*)qkv, (const T*)nullptr, padding_offset,
                     history_len, seq_len, head_num, kv_head_num, head_size, rot_dim, base, kv_rows, to_cache, page_table, max_pages);
    count_launch();
    return (int)e;
}

}  // namespace b200

extern "C" int b200_rope_decode_tab(void* qkv, int batch, int head_num, int kv_head_num, int head_size, int step,
                                    const int* step_dev, int rotary_dim, float rotary_base, const float* rope_den,
                                    int dtype, void* stream);
extern "C" int b200_rope_decode_ex(void* qkv, int batch, int head_num, int kv_head_num,
                                   int head_size, int step, const int* step_dev, int rotary_dim,
                                   float rotary_base, int dtype, void* stream) {
    return b200_rope_decode_tab(qkv, batch, head_num, kv_head_num, head_size, step, step_dev, rotary_dim, rotary_base,
                                nullptr, dtype, stream);
}
/* rope_den: optional DEVICE table from b200_rope_table() (saves the double-precision pow per launch) */
extern "C" int b200_rope_decode_tab(void* qkv, int batch, int head_num, int kv_head_num, int head_size, int step,
                                    const int* step_dev, int rotary_dim, float rotary_base, const float* rope_den,
                                    int dtype, void* stream) {
    B200_RETURN_IF(!qkv || batch < 0 || head_num <= 0 || kv_head_num <= 0 || head_size <= 0,
                   B200_EINVAL);
    B200_RETURN_IF(step < 1 || rotary_dim <= 0 || rotary_dim > head_size || (rotary_dim & 1) ||
                       (head_size & 1),
                   B200_EINVAL);
    if (batch == 0) return B200_OK;
    cudaStream_t s = (cudaStream_t)stream;
    if (dtype == B200_F32)
        return b200::rope_decode_impl<float>(qkv, batch, head_num, kv_head_num, head_size, step,
                                             step_dev, rotary_dim, rotary_base, s, rope_den);
    if (dtype == B200_F16)
        return b200::rope_decode_impl<__half>(qkv, batch, head_num, kv_head_num, head_size, step,
                                              step_dev, rotary_dim, rotary_base, s, rope_den);
    return B200_EINVAL;
}

/* ragged batch: token b sits at position seq_steps_dev[b] - 1 */
extern "C" int b200_rope_decode_ragged(void* qkv, int batch, int head_num, int kv_head_num, int head_size,
                                       const int* seq_steps_dev, int rotary_dim, float rotary_base,
                                       const float* rope_den, int dtype, void* stream) {
    B200_RETURN_IF(!qkv || !seq_steps_dev || batch < 0 || head_num <= 0 || kv_head_num <= 0 || head_size <= 0,
                   B200_EINVAL);
    B200_RETURN_IF(rotary_dim <= 0 || rotary_dim > head_size || (rotary_dim & 1) || (head_size & 1), B200_EINVAL);
    if (batch == 0) return B200_OK;
    cudaStream_t s = (cudaStream_t)stream;
    if (dtype == B200_F32)
        return b200::rope_decode_impl<float>(qkv, batch, head_num, kv_head_num, head_size, 1, seq_steps_dev,
                                             rotary_dim, rotary_base, s, rope_den, 1);
    if (dtype == B200_F16)
        return b200::rope_decode_impl<__half>(qkv, batch, head_num, kv_head_num, head_size, 1, seq_steps_dev,
                                              rotary_dim, rotary_base, s, rope_den, 1);
    return B200_EINVAL;
}

extern "C" int b200_rope_decode(void* qkv, int batch, int head_num, int kv_head_num, int head_size,
                                int step, int rotary_dim, float rotary_base, int dtype,
                                void* stream) {
    return b200_rope_decode_ex(qkv, batch, head_num, kv_head_num, head_size, step, nullptr,
                               rotary_dim, rotary_base, dtype, stream);
}

extern "C" int b200_qkv_bias_rope_transpose(void* q, void* k, void* v, const void* qkv,
                                            const void* bias, const int* padding_offset,
                                            const int* history_len, const int* input_len,
                                            int batch, int seq_len, int token_num, int head_num,
                                            int kv_head_num, int head_size, int rotary_dim,
                                            float rotary_base, int dtype, void* stream) {
    (void)input_len;  // carried by the reference signature; positions come from history_len
    B200_RETURN_IF(!q || !k || !v || !qkv || !padding_offset || !history_len, B200_EINVAL);
    B200_RETURN_IF(batch <= 0 || seq_len <= 0 || token_num < 0 || head_num <= 0 ||
                       kv_head_num <= 0 || head_size <= 0,
                   B200_EINVAL);
    B200_RETURN_IF(rotary_dim <= 0 || rotary_dim > head_size || (rotary_dim & 1), B200_EINVAL);
    if (token_num == 0) return B200_OK;
    cudaStream_t s = (cudaStream_t)stream;
    if (dtype == B200_F32)
        return b200::qkv_rope_impl<float>(q, k, v, qkv, bias, padding_offset, history_len, seq_len,
                                          token_num, head_num, kv_head_num, head_size, rotary_dim,
                                          rotary_base, s);
    if (dtype == B200_F16)
        return b200::qkv_rope_impl<__half>(q, k, v, qkv, bias, padding_offset, history_len,
                                           seq_len, token_num, head_num, kv_head_num, head_size,
                                           rotary_dim, rotary_base, s);
    return B200_EINVAL;
}

/* launchAddFusedQKVBiasTransposeAndRoPE (qkv_bias_and_RoPE.cu:248) and launchConcatKVCache
 * (concat_past_kv.cu:103) in one pass: q goes to [batch, head_num, seq_len, head_size] as before, the rotated
 * keys and the values go straight to rows history_len[b] + t of the KV-cache layer slice
 * [batch, kv_head_num, max_seq_len, head_size] (the padded k / v buffers are never written). */
extern "C" int b200_qkv_bias_rope_to_cache(void* q, void* k_cache_layer, void* v_cache_layer, const void* qkv,
                                           const void* bias, const int* padding_offset, const int* history_len,
                                           int batch, int seq_len, int token_num, int head_num, int kv_head_num,
                                           int head_size, int max_seq_len, int rotary_dim, float rotary_base,
                                           int dtype, void* stream) {
    B200_RETURN_IF(!q || !k_cache_layer || !v_cache_layer || !qkv || !padding_offset || !history_len, B200_EINVAL);
    B200_RETURN_IF(batch <= 0 || seq_len <= 0 || token_num < 0 || head_num <= 0 || kv_head_num <= 0 ||
                       head_size <= 0 || max_seq_len <= 0,
                   B200_EINVAL);
    B200_RETURN_IF(rotary_dim <= 0 || rotary_dim > head_size || (rotary_dim & 1), B200_EINVAL);
    if (token_num == 0) return B200_OK;
    cudaStream_t s = (cudaStream_t)stream;
    if (dtype == B200_F32)
        return b200::qkv_rope_impl<float>(q, k_cache_layer, v_cache_layer, qkv, bias, padding_offset, history_len,
                                          seq_len, token_num, head_num, kv_head_num, head_size, rotary_dim,
                                          rotary_base, s, max_seq_len, 1);
    if (dtype == B200_F16)
        return b200::qkv_rope_impl<__half>(q, k_cache_layer, v_cache_layer, qkv, bias, padding_offset, history_len,
                                           seq_len, token_num, head_num, kv_head_num, head_size, rotary_dim,
                                           rotary_base, s, max_seq_len, 1);
    return B200_EINVAL;
}

/* b200_qkv_bias_rope_to_cache into a PAGED cache (kv_paged.cu): the rotated keys and the values of token t of sequence b go
 * to row history_len[b] + t of the sequence's pages (pools [n_pages, kv_head_num, 128, head_size], page lists DEVICE int
 * [batch, max_pages]); fp16 / fp32. The pages must already be in the lists. */
extern "C" int b200_qkv_bias_rope_to_pages(void* q, void* k_pool_layer, void* v_pool_layer, const int* page_table, int max_pages,
                                           const void* qkv, const void* bias, const int* padding_offset,
                                           const int* history_len, int batch, int seq_len, int token_num, int head_num,
                                           int kv_head_num, int head_size, int rotary_dim, float rotary_base, int dtype,
                                           void* stream) {
    B200_RETURN_IF(!q || !k_pool_layer || !v_pool_layer || !page_table || !qkv || !padding_offset || !history_len, B200_EINVAL);
    B200_RETURN_IF(batch <= 0 || seq_len <= 0 || token_num < 0 || head_num <= 0 || kv_head_num <= 0 || head_size <= 0 ||
                       max_pages <= 0,
                   B200_EINVAL);
    B200_RETURN_IF(rotary_dim <= 0 || rotary_dim > head_size || (rotary_dim & 1), B200_EINVAL);
    if (token_num == 0) return B200_OK;
    cudaStream_t s = (cudaStream_t)stream;
    const int rows = max_pages * B200_KV_PAGE_ROWS;
    if (dtype == B200_F32)
        return b200::qkv_rope_impl<float>(q, k_pool_layer, v_pool_layer, qkv, bias, padding_offset, history_len, seq_len,
                                          token_num, head_num, kv_head_num, head_size, rotary_dim, rotary_base, s, rows, 1,
                                          page_table, max_pages);
    if (dtype == B200_F16)
        return b200::qkv_rope_impl<__half>(q, k_pool_layer, v_pool_layer, qkv, bias, padding_offset, history_len, seq_len,
                                           token_num, head_num, kv_head_num, head_size, rotary_dim, rotary_base, s, rows, 1,
                                           page_table, max_pages);
    return B200_EINVAL;
}
