// KV-cache data movement: ConcatKVCache, RepeatKVCache, TransposeOutRemovePadding.
// Pure copies -> bit-exact. One thread moves one 16-byte packet; K and V are handled by the
// same launch (blockIdx.y selects which). 64-bit indexing throughout.
// Reference: src/kernels/concat_past_kv.cu:17-56,103-144; repeat_kv.cu:27-70,76-120;
// fused_transpose_and_remv_pad.cu:18-47.
#include "common.cuh"
#include "launch_count.h"

namespace b200 {

// src [batch, kv_heads, max_q_len, d] -> cache[layer] [batch, kv_heads, max_seq_len, d]
template <int PB>  // packet bytes: 16 (vector) or sizeof(T) (scalar fallback)
__global__ void __launch_bounds__(256)
concat_kv_kernel(const char* __restrict__ k_src, const char* __restrict__ v_src,
                 char* __restrict__ k_dst, char* __restrict__ v_dst,
                 const int* __restrict__ cur_len, const int* __restrict__ history_len,
                 int kv_heads, int max_q_len, int max_seq_len, int row_packets, size_t total) {
    pdl_wait();
    const char* src = blockIdx.y ? v_src : k_src;
    char* dst = blockIdx.y ? v_dst : k_dst;
    for (size_t it = (size_t)blockIdx.x * blockDim.x + threadIdx.x; it < total;
         it += (size_t)gridDim.x * blockDim.x) {
        const int p = (int)(it % row_packets);
        size_t r = it / row_packets;
        const int t = (int)(r % max_q_len);
        r /= max_q_len;
        const int head = (int)(r % kv_heads);
        const int b = (int)(r / kv_heads);
        // rows that would land beyond the cache slice (history + length > max_seq_len: a caller error the launcher
        // cannot see, the lengths live on the device) are dropped, never written outside the slice
        if (t < cur_len[b] && history_len[b] + t < max_seq_len && history_len[b] >= 0) {
            const size_t drow = ((size_t)b * kv_heads + head) * max_seq_len + history_len[b] + t;
            const char* s = src + it * PB;
            char* d = dst + (drow * row_packets + p) * PB;
            if (PB == 16) st16(d, ld16_stream(s));
            else if (PB == 4) *reinterpret_cast<uint32_t*>(d) = *reinterpret_cast<const uint32_t*>(s);
            else *reinterpret_cast<uint16_t*>(d) = *reinterpret_cast<const uint16_t*>(s);
        }
    }
    pdl_launch_dependents();
}

// cache[layer] [batch, kv_heads, max_seq_len, d] -> dst [batch, heads, max_k_len, d]
template <int PB>
__global__ void __launch_bounds__(256)
repeat_kv_kernel(const char* __restrict__ k_src, const char* __restrict__ v_src,
                 char* __restrict__ k_dst, char* __restrict__ v_dst,
                 const int* __restrict__ context_len, int heads, int q_per_kv, int max_seq_len,
                 int max_k_len, int row_packets, size_t total) {
    pdl_wait();
    const char* src = blockIdx.y ? v_src : k_src;
    char* dst = blockIdx.y ? v_dst : k_dst;
    const int kv_heads = heads / q_per_kv;
    for (size_t it = (size_t)blockIdx.x * blockDim.x + threadIdx.x; it < total;
         it += (size_t)gridDim.x * blockDim.x) {
        const int p = (int)(it % row_packets);
        size_t r = it / row_packets;
        const int t = (int)(r % max_k_len);
        r /= max_k_len;
        const int head = (int)(r % heads);
        const int b = (int)(r / heads);
        if (t < context_len[b]) {
            const size_t srow = ((size_t)b * kv_heads + head / q_per_kv) * max_seq_len + t;
            const char* s = src + (srow * row_packets + p) * PB;
            char* d = dst + it * PB;
            if (PB == 16) st16(d, ld16(s));
            else if (PB == 4) *reinterpret_cast<uint32_t*>(d) = *reinterpret_cast<const uint32_t*>(s);
            else *reinterpret_cast<uint16_t*>(d) = *reinterpret_cast<const uint16_t*>(s);
        }
    }
    pdl_launch_dependents();
}

// src [batch, heads, seq_len, d] -> dst [token_num, heads*d]
template <int PB>
__global__ void __launch_bounds__(256)
transpose_remove_pad_kernel(const char* __restrict__ src, char* __restrict__ dst,
                            const int* __restrict__ padding_offset, int seq_len, int heads,
                            int row_packets, size_t total) {
    pdl_wait();
    for (size_t it = (size_t)blockIdx.x * blockDim.x + threadIdx.x; it < total;
         it += (size_t)gridDim.x * blockDim.x) {
        const int p = (int)(it % row_packets);
        size_t r = it / row_packets;
        const int head = (int)(r % heads);
        const int token = (int)(r / heads);
        const int padded = token + padding_offset[token];
        const int b = padded / seq_len, local = padded - b * seq_len;
        const size_t srow = ((size_t)b * heads + head) * seq_len + local;
        const char* s = src + (srow * row_packets + p) * PB;
        char* d = dst + it * PB;
        if (PB == 16) st16(d, ld16_stream(s));
        else if (PB == 4) *reinterpret_cast<uint32_t*>(d) = *reinterpret_cast<const uint32_t*>(s);
        else *reinterpret_cast<uint16_t*>(d) = *reinterpret_cast<const uint16_t*>(s);
    }
    pdl_launch_dependents();
}

static inline unsigned grid_for(size_t total) {
    size_t blocks = (total + 255) / 256;
    const size_t cap = (size_t)sm_count() * 8;
    if (blocks > cap) blocks = cap;
    if (blocks == 0) blocks = 1;
    return (unsigned)blocks;
}
static inline bool a16(const void* p) { return (reinterpret_cast<uintptr_t>(p) & 15u) == 0; }
static inline int esize(int dtype) { return dtype == B200_F16 ? 2 : 4; }

}  // namespace b200

using namespace b200;

extern "C" int b200_concat_kv_cache(const void* k_src, const void* v_src, void* k_cache,
                                    void* v_cache, size_t layer_offset, const int* cur_query_len,
                                    const int* history_len, int batch, int kv_head_num,
                                    int max_q_len, int max_seq_len, int head_size, int dtype,
                                    void* stream) {
    B200_RETURN_IF(!k_src || !v_src || !k_cache || !v_cache || !cur_query_len || !history_len,
                   B200_EINVAL);
    B200_RETURN_IF(batch <= 0 || kv_head_num <= 0 || max_q_len <= 0 || max_seq_len <= 0 ||
                       head_size <= 0 || (dtype != B200_F32 && dtype != B200_F16),
                   B200_EINVAL);
    const int es = esize(dtype);
    char* kd = (char*)k_cache + layer_offset * es;
    char* vd = (char*)v_cache + layer_offset * es;
    const size_t row_bytes = (size_t)head_size * es;
    const bool vec = row_bytes % 16 == 0 && a16(k_src) && a16(v_src) && a16(kd) && a16(vd);
    const int pb = vec ? 16 : es;
    const int row_packets = (int)(row_bytes / pb);
    const size_t total = (size_t)batch * kv_head_num * max_q_len * row_packets;
    cudaStream_t s = (cudaStream_t)stream;
    dim3 grid(grid_for(total), 2);
    cudaError_t e;
#define CALL(PB)                                                                               \
    e = launch_k(concat_kv_kernel<PB>, grid, dim3(256), 0, s, true, (const char*)k_src,        \
                 (const char*)v_src, kd, vd, cur_query_len, history_len, kv_head_num, max_q_len, \
                 max_seq_len, row_packets, total)
    if (pb == 16) CALL(16); else if (pb == 4) CALL(4); else CALL(2);
#undef CALL
    count_launch();
    return (int)e;
}

extern "C" int b200_repeat_kv_cache(const void* k_cache, const void* v_cache, void* k_dst,
                                    void* v_dst, size_t layer_offset, const int* context_len,
                                    int batch, int head_num, int kv_head_num, int max_seq_len,
                                    int max_k_len, int head_size, int dtype, void* stream) {
    B200_RETURN_IF(!k_cache || !v_cache || !k_dst || !v_dst || !context_len, B200_EINVAL);
    B200_RETURN_IF(batch <= 0 || head_num <= 0 || kv_head_num <= 0 || head_num % kv_head_num ||
                       max_seq_len <= 0 || max_k_len <= 0 || head_size <= 0 ||
                       (dtype != B200_F32 && dtype != B200_F16),
                   B200_EINVAL);
    const int es = esize(dtype);
    const char* ks = (const char*)k_cache + layer_offset * es;
    const char* vs = (const char*)v_cache + layer_offset * es;
    const size_t row_bytes = (size_t)head_size * es;
    const bool vec = row_bytes % 16 == 0 && a16(ks) && a16(vs) && a16(k_dst) && a16(v_dst);
    const int pb = vec ? 16 : es;
    const int row_packets = (int)(row_bytes / pb);
    const size_t total = (size_t)batch * head_num * max_k_len * row_packets;
    cudaStream_t s = (cudaStream_t)stream;
    dim3 grid(grid_for(total), 2);
    cudaError_t e;
#define CALL(PB)                                                                                 \
    e = launch_k(repeat_kv_kernel<PB>, grid, dim3(256), 0, s, true, ks, vs, (char*)k_dst,        \
                 (char*)v_dst, context_len, head_num, head_num / kv_head_num, max_seq_len,       \
                 max_k_len, row_packets, total)
    if (pb == 16) CALL(16); else if (pb == 4) CALL(4); else CALL(2);
#undef CALL
    count_launch();
    return (int)e;
}

extern "C" int b200_transpose_remove_padding(const void* src, void* dst,
                                             const int* padding_offset, int token_num, int batch,
                                             int seq_len, int head_num, int head_size, int dtype,
                                             void* stream) {
    B200_RETURN_IF(!src || !dst || !padding_offset, B200_EINVAL);
    B200_RETURN_IF(token_num < 0 || batch <= 0 || seq_len <= 0 || head_num <= 0 || head_size <= 0 ||
                       (dtype != B200_F32 && dtype != B200_F16),
                   B200_EINVAL);
    if (token_num == 0) return B200_OK;
    const int es = esize(dtype);
    const size_t row_bytes = (size_t)head_size * es;
    const bool vec = row_bytes % 16 == 0 && a16(src) && a16(dst);
    const int pb = vec ? 16 : es;
    const int row_packets = (int)(row_bytes / pb);
    const size_t total = (size_t)token_num * head_num * row_packets;
    cudaStream_t s = (cudaStream_t)stream;
    dim3 grid(grid_for(total));
    cudaError_t e;
#define CALL(PB)                                                                              \
    e = launch_k(transpose_remove_pad_kernel<PB>, grid, dim3(256), 0, s, true,                \
                 (const char*)src, (char*)dst, padding_offset, seq_len, head_num, row_packets, \
                 total)
    if (pb == 16) CALL(16); else if (pb == 4) CALL(4); else CALL(2);
#undef CALL
    count_launch();
    return (int)e;
}
