// fp8 (e4m3) KV cache for the batched long-context decode ("KV-cache layout evolution ... fp8 KV").
// The reference keeps K and V in T, [layers, B, hk, max_seq, d] (src/models/llama/llama.cpp:77-78); at cfg4
// (32 sequences x 4096 keys) those rows are 84 % of a decode step's bytes and the split-KV attention already streams
// them at 0.93 of the measured HBM peak - bytes are the only lever left there.
//
// Layout: K8 / V8 [layers, B, hk, max_seq, d] one byte per element + one fp32 scale per cached ROW
// (k_scale / v_scale [layers, B, hk, max_seq]): value = scale * e4m3. The scale of a row is amax(row) / 448, fixed when
// the row is appended (no calibration pass, no re-quantisation of history). A d = 128 row is 128 B + 4 B instead of 256 B.
// Attention semantics are those of fused_decoder_self_attention.cu:80-291 (decode_attention.cu) on the DEQUANTISED rows:
// a row is rounded once, when it enters the cache, and every later step - and the step that appends it - sees the
// same rounded values. Not bit-compatible with the fp16 cache: an engine option, with its own parity test against the
// oracle run on the same rounded rows (tests/test_kv_fp8_gpu.py).
#include <cuda_fp8.h>

#include "common.cuh"
#include "launch_count.h"

namespace b200 {

__device__ __forceinline__ float fp8_to_f(unsigned char v) {
    const __half_raw h = __nv_cvt_fp8_to_halfraw((__nv_fp8_storage_t)v, __NV_E4M3);
    return __half2float(*reinterpret_cast<const __half*>(&h));
}
__device__ __forceinline__ unsigned char f_to_fp8(float v) {
    return (unsigned char)__nv_cvt_float_to_fp8(v, __NV_SATFINITE, __NV_E4M3);
}

// cache row `row` of (sequence b, kv head kvh): its index among the rows of a layer slice (x d = element offset, x 1 = scale
// offset). Dense: [B, hk, max_seq]; paged (kv_paged.cu): [n_pages, hk, 128] through the sequence's page list.
__device__ __forceinline__ size_t kv_row_index(const int* __restrict__ page_table, int max_pages, int b, int kvh, int kv_head_num,
                                               int max_seq, int row) {
    if (page_table == nullptr) return ((size_t)b * kv_head_num + kvh) * max_seq + (size_t)row;
    const int pg = page_table[(size_t)b * max_pages + row / B200_KV_PAGE_ROWS];
    return ((size_t)pg * kv_head_num + kvh) * B200_KV_PAGE_ROWS + (size_t)(row % B200_KV_PAGE_ROWS);
}

// one warp quantises one row of d halves: scale = amax / 448 (1 for an all-zero row)
__device__ __forceinline__ void quantize_row(const __half* __restrict__ src, unsigned char* __restrict__ dst, float* scale_out,
                                             int d, int lane) {
    float amax = 0.f;
    for (int e = lane; e < d; e += 32) amax = fmaxf(amax, fabsf(__half2float(src[e])));
    amax = warp_max(amax);
    const float scale = amax > 0.f ? amax * (1.f / 448.f) : 1.f;
    const float inv = 1.f / scale;
    for (int e = lane; e < d; e += 32) dst[e] = f_to_fp8(__half2float(src[e]) * inv);
    if (lane == 0) *scale_out = scale;
}

// new token of every sequence: k, v of qkv [B, h+2hk, d] (already rotated) -> cache row step-1. grid = B*hk, 64 threads.
__global__ void __launch_bounds__(64)
kv_append_fp8_kernel(const __half* __restrict__ qkv, unsigned char* k8, unsigned char* v8, float* k_scale, float* v_scale,
                     int head_num, int kv_head_num, int d, int max_seq, int step, const int* step_dev, int step_stride,
                     const unsigned char* finished, const int* __restrict__ page_table, int max_pages) {
    pdl_wait();
    const int b = blockIdx.x / kv_head_num, kvh = blockIdx.x - b * kv_head_num;
    if (finished != nullptr && finished[b] != 0) {   // retired sequence: nothing enters its cache any more
        pdl_launch_dependents();
        return;
    }
    const int st = step_dev ? step_dev[b * step_stride] : step;
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const __half* tok = qkv + (size_t)b * (head_num + 2 * kv_head_num) * d;
    if (st >= 1 && st <= max_seq) {
        const size_t row = kv_row_index(page_table, max_pages, b, kvh, kv_head_num, max_seq, st - 1);
        if (warp == 0) quantize_row(tok + (size_t)(head_num + kvh) * d, k8 + row * d, k_scale + row, d, lane);
        else quantize_row(tok + (size_t)(head_num + kv_head_num + kvh) * d, v8 + row * d, v_scale + row, d, lane);
    }
    pdl_launch_dependents();
}

// rows [row0, row0 + n_rows) of an fp16 layer slice [B, hk, src_seq, d] -> the fp8 slice [B, hk, max_seq, d] at
// dst_row0 (the prompt pass quantises the rows it has just attended over in fp16). One warp per row.
__global__ void __launch_bounds__(256)
kv_quantize_rows_kernel(const __half* __restrict__ src, unsigned char* dst, float* scale, int bh_count, int src_seq, int max_seq,
                        int d, const int* __restrict__ n_rows_b, const int* __restrict__ dst_row0_b, int kv_head_num, int max_rows,
                        const int* __restrict__ page_table, int max_pages) {
    const int warp_g = (blockIdx.x * blockDim.x + threadIdx.x) >> 5, lane = threadIdx.x & 31;
    const int total = bh_count * max_rows;
    if (warp_g >= total) return;
    const int bh = warp_g / max_rows, t = warp_g - bh * max_rows;
    const int b = bh / kv_head_num;
    if (t >= n_rows_b[b]) return;
    const int drow = dst_row0_b[b] + t;
    if (drow < 0 || drow >= max_seq) return;
    const size_t r = kv_row_index(page_table, max_pages, b, bh - b * kv_head_num, kv_head_num, max_seq, drow);
    quantize_row(src + ((size_t)bh * src_seq + t) * d, dst + r * d, scale + r, d, lane);
}

// dequantise a whole layer slice to fp16 (tests: the oracle is run on exactly the rows the GPU attends over)
__global__ void __launch_bounds__(256)
kv_dequant_kernel(const unsigned char* __restrict__ src, const float* __restrict__ scale, __half* dst, size_t rows, int d) {
    const size_t n = rows * d;
    for (size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (size_t)gridDim.x * blockDim.x)
        dst[i] = __float2half_rn(fp8_to_f(src[i]) * scale[i / d]);
}

// paged fp8 cache -> dense fp16 [B, hk, dst_seq, d], rows [0, n_rows[b]) of every sequence (tests: the oracle runs on the
// rows the GPU attends over). One thread per element pair would do; one per element keeps it trivially correct.
__global__ void __launch_bounds__(256)
kv_dequant_pages_kernel(const unsigned char* __restrict__ pool, const float* __restrict__ scale, __half* dst,
                        const int* __restrict__ page_table, int max_pages, const int* __restrict__ n_rows, int kv_head_num,
                        int dst_seq, int d, int max_rows) {
    const int bh = blockIdx.y, b = bh / kv_head_num, kvh = bh - b * kv_head_num;
    const int rows = min(n_rows[b], max_rows);
    for (int i = blockIdx.x * blockDim.x + threadIdx.x; i < rows * d; i += gridDim.x * blockDim.x) {
        const int row = i / d, e = i - row * d;
        const size_t r = kv_row_index(page_table, max_pages, b, kvh, kv_head_num, 0, row);
        dst[((size_t)bh * dst_seq + row) * d + e] = __float2half_rn(fp8_to_f(pool[r * d + e]) * scale[r]);
    }
}

// synthetic cache rows for the benches: uniform(-amp, amp) values, scale = amp / 448 per row
__global__ void __launch_bounds__(256)
kv_fill_fp8_kernel(unsigned char* cache, float* scale, size_t rows_total, int max_seq, int len, int d, uint64_t seed, float amp) {
    const size_t n = rows_total * (size_t)len * d;
    const float sc = amp > 0.f ? amp * (1.f / 448.f) : 1.f;
    for (size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (size_t)gridDim.x * blockDim.x) {
        const size_t r = i / ((size_t)len * d);
        const size_t rem = i - r * (size_t)len * d;
        uint64_t z = seed + 0x9E3779B97F4A7C15ull * (i + 1);
        z = (z ^ (z >> 30)) * 0xBF58476D1CE4E5B9ull;
        z = (z ^ (z >> 27)) * 0x94D049BB133111EBull;
        z ^= z >> 31;
        const float u = (uint32_t)(z >> 40) * (1.0f / 16777216.0f);
        cache[r * (size_t)max_seq * d + rem] = f_to_fp8((2.f * u - 1.f) * 448.f);
        if (rem % d == 0) scale[r * (size_t)max_seq + rem / d] = sc;
    }
}

}  // namespace b200

namespace b200 {
const unsigned char* attn_finished_ptr();   // decode_attention.cu
}
using namespace b200;

/* the new token's k, v (qkv [batch, h+2hk, d], after RoPE) -> row step-1 of the fp8 cache layer slice, quantised per row */
static int kv_append_fp8_impl(const void* qkv, void* k_cache8, void* v_cache8, float* k_scale, float* v_scale, int batch,
                              int head_num, int kv_head_num, int head_size, int max_seq_len, int step,
                              const int* step_dev, int step_stride, const int* page_table, int max_pages, void* stream) {
    B200_RETURN_IF(!qkv || !k_cache8 || !v_cache8 || !k_scale || !v_scale, B200_EINVAL);
    B200_RETURN_IF(batch <= 0 || head_num <= 0 || kv_head_num <= 0 || head_size <= 0 || max_seq_len <= 0, B200_EINVAL);
    B200_RETURN_IF(step < 1 || step > max_seq_len, B200_EINVAL);
    cudaError_t e = launch_k(kv_append_fp8_kernel, dim3(batch * kv_head_num), dim3(64), 0, (cudaStream_t)stream, true,
                             (const __half*)qkv, (unsigned char*)k_cache8, (unsigned char*)v_cache8, k_scale, v_scale,
                             head_num, kv_head_num, head_size, max_seq_len, step, step_dev, step_stride, attn_finished_ptr(),
                             page_table, max_pages);
    count_launch();
    return (int)e;
}

extern "C" int b200_kv_append_fp8(const void* qkv, void* k_cache8, void* v_cache8, float* k_scale, float* v_scale, int batch,
                                  int head_num, int kv_head_num, int head_size, int max_seq_len, int step,
                                  const int* step_dev, int step_stride, void* stream) {
    return kv_append_fp8_impl(qkv, k_cache8, v_cache8, k_scale, v_scale, batch, head_num, kv_head_num, head_size, max_seq_len,
                              step, step_dev, step_stride, nullptr, 0, stream);
}

/* the same into a paged fp8 cache: pools [n_pages, kv_heads, 128, d] (+ scale pools [n_pages, kv_heads, 128]) and the
 * sequences' page lists (DEVICE int [batch, max_pages]); the page of row step-1 must already be in the list */
extern "C" int b200_kv_append_fp8_paged(const void* qkv, void* k_pool8, void* v_pool8, float* k_scale_pool, float* v_scale_pool,
                                        const int* page_table, int max_pages, int batch, int head_num, int kv_head_num,
                                        int head_size, int step, const int* step_dev, int step_stride, void* stream) {
    B200_RETURN_IF(!page_table || max_pages <= 0, B200_EINVAL);
    return kv_append_fp8_impl(qkv, k_pool8, v_pool8, k_scale_pool, v_scale_pool, batch, head_num, kv_head_num, head_size,
                              max_pages * B200_KV_PAGE_ROWS, step, step_dev, step_stride, page_table, max_pages, stream);
}

/* rows t < n_rows[b] of an fp16 slice src [batch, kv_heads, src_seq, d] -> rows dst_row0[b] + t of the fp8 slice
 * [batch, kv_heads, max_seq_len, d] (n_rows / dst_row0: DEVICE int [batch]; max_rows >= every n_rows[b]) */
static int kv_quantize_fp8_impl(const void* src, void* dst8, float* scale, const int* n_rows, const int* dst_row0,
                                int batch, int kv_head_num, int src_seq, int max_seq_len, int head_size, int max_rows,
                                const int* page_table, int max_pages, void* stream) {
    B200_RETURN_IF(!src || !dst8 || !scale || !n_rows || !dst_row0, B200_EINVAL);
    B200_RETURN_IF(batch <= 0 || kv_head_num <= 0 || src_seq <= 0 || max_seq_len <= 0 || head_size <= 0 || max_rows < 0, B200_EINVAL);
    if (max_rows == 0) return 0;
    const long long warps = (long long)batch * kv_head_num * max_rows;
    const long long blocks = (warps * 32 + 255) / 256;
    B200_RETURN_IF(blocks >= (1ll << 31), B200_EUNSUPPORTED);
    kv_quantize_rows_kernel<<<(unsigned)blocks, 256, 0, (cudaStream_t)stream>>>(
        (const __half*)src, (unsigned char*)dst8, scale, batch * kv_head_num, src_seq, max_seq_len, head_size, n_rows, dst_row0,
        kv_head_num, max_rows, page_table, max_pages);
    count_launch();
    return (int)cudaGetLastError();
}

extern "C" int b200_kv_quantize_fp8(const void* src, void* dst8, float* scale, const int* n_rows, const int* dst_row0,
                                    int batch, int kv_head_num, int src_seq, int max_seq_len, int head_size, int max_rows,
                                    void* stream) {
    return kv_quantize_fp8_impl(src, dst8, scale, n_rows, dst_row0, batch, kv_head_num, src_seq, max_seq_len, head_size, max_rows,
                                nullptr, 0, stream);
}

/* rows t < n_rows[b] of an fp16 slice [batch, kv_heads, src_seq, d] -> rows dst_row0[b] + t of sequence b's pages, quantised */
extern "C" int b200_kv_quantize_fp8_paged(const void* src, void* pool8, float* scale_pool, const int* page_table, int max_pages,
                                          const int* n_rows, const int* dst_row0, int batch, int kv_head_num, int src_seq,
                                          int head_size, int max_rows, void* stream) {
    B200_RETURN_IF(!page_table || max_pages <= 0, B200_EINVAL);
    return kv_quantize_fp8_impl(src, pool8, scale_pool, n_rows, dst_row0, batch, kv_head_num, src_seq,
                                max_pages * B200_KV_PAGE_ROWS, head_size, max_rows, page_table, max_pages, stream);
}

/* rows [0, n_rows[b]) of sequence b's fp8 pages -> dense fp16 dst [batch, kv_heads, dst_seq, d] (tests) */
extern "C" int b200_kv_dequant_fp8_paged(const void* pool8, const float* scale_pool, void* dst, const int* page_table,
                                         int max_pages, const int* n_rows, int batch, int kv_head_num, int dst_seq,
                                         int head_size, int max_rows, void* stream) {
    B200_RETURN_IF(!pool8 || !scale_pool || !dst || !page_table || !n_rows, B200_EINVAL);
    B200_RETURN_IF(batch <= 0 || kv_head_num <= 0 || dst_seq <= 0 || head_size <= 0 || max_pages <= 0 || max_rows < 0, B200_EINVAL);
    B200_RETURN_IF((long long)batch * kv_head_num > 65535, B200_EUNSUPPORTED);
    if (max_rows == 0) return 0;
    long long bx = ((long long)max_rows * head_size + 255) / 256;
    if (bx > 1024) bx = 1024;
    kv_dequant_pages_kernel<<<dim3((unsigned)bx, (unsigned)(batch * kv_head_num)), 256, 0, (cudaStream_t)stream>>>(
        (const unsigned char*)pool8, scale_pool, (__half*)dst, page_table, max_pages, n_rows, kv_head_num, dst_seq, head_size, max_rows);
    count_launch();
    return (int)cudaGetLastError();
}

/* whole slice back to fp16: dst [rows, d] = scale[row] * e4m3 (tests / debugging) */
extern "C" int b200_kv_dequant_fp8(const void* src8, const float* scale, void* dst, size_t rows, int head_size, void* stream) {
    B200_RETURN_IF(!src8 || !scale || !dst || head_size <= 0, B200_EINVAL);
    if (rows == 0) return 0;
    kv_dequant_kernel<<<148 * 8, 256, 0, (cudaStream_t)stream>>>((const unsigned char*)src8, scale, (__half*)dst, rows, head_size);
    count_launch();
    return (int)cudaGetLastError();
}

namespace b200 {
int kv_fill_fp8(void* cache8, float* scale, size_t rows_total, int max_seq, int len, int d, uint64_t seed, float amp,
                cudaStream_t s) {
    kv_fill_fp8_kernel<<<148 * 32, 256, 0, s>>>((unsigned char*)cache8, scale, rows_total, max_seq, len, d, seed, amp);
    return (int)cudaGetLastError();
}
}  // namespace b200
