// Prefill scale-mask-softmax over the last axis:
//   s = scale*qk + (1-mask)*(-10000);  p = exp(s - max(s)) / (sum + 1e-6)
// Reference: src/kernels/attn_softmax_kernel.cu:80-173 (fp32), :179-290 (half), :337-372.
// Deviations: correct two-stage block reduce, running max starts at -inf, no
// K < 4096 limit. The additive -10000 mask and the +1e-6 denominator are kept.
// One CTA per (b, head, q) row; the row lives in registers between the three passes, so qk and
// mask are read once and the score written once. May run in place (qk == score).
#include "common.cuh"
#include "launch_count.h"

namespace b200 {

constexpr int kSmCache = 8;  // V-element chunks per thread kept in registers

template <typename T, int V> struct Chunk;
template <typename T> struct Chunk<T, 1> {
    static __device__ __forceinline__ void load(const T* p, float (&f)[1]) { f[0] = DT<T>::to_f(*p); }
    static __device__ __forceinline__ void store(T* p, const float (&f)[1]) { *p = DT<T>::from_f(f[0]); }
};
template <> struct Chunk<float, 4> {
    static __device__ __forceinline__ void load(const float* p, float (&f)[4]) { Unpack<float>::get(ld16(p), f); }
    static __device__ __forceinline__ void store(float* p, const float (&f)[4]) { st16(p, Unpack<float>::put(f)); }
};
template <> struct Chunk<__half, 8> {
    static __device__ __forceinline__ void load(const __half* p, float (&f)[8]) { Unpack<__half>::get(ld16(p), f); }
    static __device__ __forceinline__ void store(__half* p, const float (&f)[8]) { st16(p, Unpack<__half>::put(f)); }
};

template <typename T, int V>
__global__ void __launch_bounds__(256)
scale_mask_softmax_kernel(const T* qk, const T* __restrict__ mask, T* score,
                          int head_num, int q_len, int k_len, float scale) {
    __shared__ float scratch[33];
    pdl_wait();
    const size_t row = blockIdx.x;  // over batch*head*q
    const int q = (int)(row % q_len);
    const int b = (int)(row / ((size_t)q_len * head_num));
    const T* qrow = qk + row * k_len;
    const T* mrow = mask + ((size_t)b * q_len + q) * k_len;
    T* orow = score + row * k_len;
    const int chunks = k_len / V;

    float vals[kSmCache][V];
    float mx = -INFINITY;
#pragma unroll
    for (int c = 0; c < kSmCache; ++c) {
        const int p = threadIdx.x + c * blockDim.x;
        if (p < chunks) {
            float m[V];
            Chunk<T, V>::load(qrow + (size_t)p * V, vals[c]);
            Chunk<T, V>::load(mrow + (size_t)p * V, m);
#pragma unroll
            for (int i = 0; i < V; ++i) {
                vals[c][i] = scale * vals[c][i] + (1.f - m[i]) * (-10000.0f);
                mx = fmaxf(mx, vals[c][i]);
            }
        }
    }
    // rows longer than the register cache: the tail is recomputed from memory in each pass; a
    // thread only ever re-reads chunks it alone writes (in pass 3), so in-place stays safe.
    for (int p = threadIdx.x + kSmCache * blockDim.x; p < chunks; p += blockDim.x) {
        float v[V], m[V];
        Chunk<T, V>::load(qrow + (size_t)p * V, v);
        Chunk<T, V>::load(mrow + (size_t)p * V, m);
#pragma unroll
        for (int i = 0; i < V; ++i) mx = fmaxf(mx, scale * v[i] + (1.f - m[i]) * (-10000.0f));
    }
    mx = block_max(mx, scratch);

    float sum = 0.f;
#pragma unroll
    for (int c = 0; c < kSmCache; ++c) {
        const int p = threadIdx.x + c * blockDim.x;
        if (p < chunks) {
#pragma unroll
            for (int i = 0; i < V; ++i) {
                vals[c][i] = expf(vals[c][i] - mx);
                sum += vals[c][i];
            }
        }
    }
    for (int p = threadIdx.x + kSmCache * blockDim.x; p < chunks; p += blockDim.x) {
        float v[V], m[V];
        Chunk<T, V>::load(qrow + (size_t)p * V, v);
        Chunk<T, V>::load(mrow + (size_t)p * V, m);
#pragma unroll
        for (int i = 0; i < V; ++i) sum += expf(scale * v[i] + (1.f - m[i]) * (-10000.0f) - mx);
    }
    sum = block_sum(sum, scratch);
    const float inv = 1.f / (sum + 1e-6f);
    pdl_launch_dependents();

#pragma unroll
    for (int c = 0; c < kSmCache; ++c) {
        const int p = threadIdx.x + c * blockDim.x;
        if (p < chunks) {
#pragma unroll
            for (int i = 0; i < V; ++i) vals[c][i] *= inv;
            Chunk<T, V>::store(orow + (size_t)p * V, vals[c]);
        }
    }
    for (int p = threadIdx.x + kSmCache * blockDim.x; p < chunks; p += blockDim.x) {
        float v[V], m[V];
        Chunk<T, V>::load(qrow + (size_t)p * V, v);
        Chunk<T, V>::load(mrow + (size_t)p * V, m);
#pragma unroll
        for (int i = 0; i < V; ++i) v[i] = expf(scale * v[i] + (1.f - m[i]) * (-10000.0f) - mx) * inv;
        Chunk<T, V>::store(orow + (size_t)p * V, v);
    }
}

template <typename T>
static int softmax_impl(const void* qk, const void* mask, void* score, int batch, int head_num,
                        int q_len, int k_len, float scale, cudaStream_t s) {
    constexpr int E = DT<T>::kPer16B;
    const size_t rows = (size_t)batch * head_num * q_len;
    if (rows == 0 || k_len == 0) return B200_OK;
    B200_RETURN_IF(rows > 0x7fffffffull, B200_EUNSUPPORTED);
    auto al = [](const void* p) { return (reinterpret_cast<uintptr_t>(p) & 15u) == 0; };
    const bool vec = (k_len % E == 0) && al(qk) && al(mask) && al(score);
    const int chunks = vec ? k_len / E : k_len;
    int block = ((chunks + kSmCache - 1) / kSmCache + 31) / 32 * 32;
    if (block < 32) block = 32;
    if (block > 256) block = 256;
    cudaError_t e;
    if (vec)
        e = launch_k(scale_mask_softmax_kernel<T, E>, dim3((unsigned)rows), dim3(block), 0, s, true,
                     (const T*)qk, (const T*)mask, (T*)score, head_num, q_len, k_len, scale);
    else
        e = launch_k(scale_mask_softmax_kernel<T, 1>, dim3((unsigned)rows), dim3(block), 0, s, true,
                     (const T*)qk, (const T*)mask, (T*)score, head_num, q_len, k_len, scale);
    count_launch();
    return (int)e;
}

}  // namespace b200

extern "C" int b200_scale_mask_softmax(const void* qk, const void* mask, void* score, int batch,
                                       int head_num, int q_len, int k_len, float scale, int dtype,
                                       void* stream) {
    B200_RETURN_IF(!qk || !mask || !score || batch < 0 || head_num < 0 || q_len < 0 || k_len < 0,
                   B200_EINVAL);
    cudaStream_t s = (cudaStream_t)stream;
    if (dtype == B200_F32)
        return b200::softmax_impl<float>(qk, mask, score, batch, head_num, q_len, k_len, scale, s);
    if (dtype == B200_F16)
        return b200::softmax_impl<__half>(qk, mask, score, batch, head_num, q_len, k_len, scale, s);
    return B200_EINVAL;
}
