// Prefill bookkeeping: padding offsets (int32 prefix sums) and the causal mask (0/1 in T).
// Integer / predicate work -> bit-exact.
// Reference: src/kernels/cal_paddingoffset.cu:51-72 (a serial <<<1,1>>> loop there; here one CTA
// does a block scan then a parallel fill), src/kernels/build_causal_mask.cu:5-47.
#include "common.cuh"
#include "launch_count.h"

namespace b200 {

// cum_seqlens[b] = sum_{j<b} len[j]; padding_offset[cum[b]+i] = b*max_q_len - cum[b], i < len[b]
__global__ void __launch_bounds__(1024)
padding_offset_kernel(int* __restrict__ padding_offset, int* __restrict__ cum_seqlens,
                      const int* __restrict__ input_lens, int batch, int max_q_len) {
    __shared__ int warp_tot[32];
    __shared__ int carry_s;
    pdl_wait();
    const int lane = threadIdx.x & 31, wid = threadIdx.x >> 5, nw = blockDim.x >> 5;
    if (threadIdx.x == 0) carry_s = 0;
    __syncthreads();
    // chunked exclusive scan over the batch
    for (int base = 0; base < batch; base += blockDim.x) {
        const int b = base + threadIdx.x;
        const int len = (b < batch) ? input_lens[b] : 0;
        int incl = len;
#pragma unroll
        for (int o = 1; o < 32; o <<= 1) {
            int n = __shfl_up_sync(0xffffffffu, incl, o);
            if (lane >= o) incl += n;
        }
        if (lane == 31) warp_tot[wid] = incl;
        __syncthreads();
        int woff = 0;
        for (int w = 0; w < wid; ++w) woff += warp_tot[w];
        const int carry = carry_s;
        if (b < batch) cum_seqlens[b] = carry + woff + incl - len;
        __syncthreads();
        if (threadIdx.x == blockDim.x - 1) carry_s = carry + woff + incl;
        __syncthreads();
        (void)nw;
    }
    if (threadIdx.x == 0) cum_seqlens[batch] = carry_s;
    __syncthreads();  // cum_seqlens written by this CTA are visible to it after the barrier
    const size_t total = (size_t)batch * max_q_len;
    for (size_t idx = threadIdx.x; idx < total; idx += blockDim.x) {
        const int b = (int)(idx / max_q_len), i = (int)(idx - (size_t)b * max_q_len);
        if (i < input_lens[b]) {
            const int cum = cum_seqlens[b];
            padding_offset[cum + i] = b * max_q_len - cum;
        }
    }
    pdl_launch_dependents();
}

template <typename T>
__device__ __forceinline__ T mask_val(int q, int k, int qlen, int klen, int history_visible) {
    bool one = q < qlen && k < klen && k <= q + (klen - qlen);
    if (!history_visible) one = one && (k >= klen - qlen);
    return DT<T>::from_f(one ? 1.f : 0.f);
}

template <typename T>
__global__ void __launch_bounds__(256)
causal_mask_kernel(T* __restrict__ mask, const int* __restrict__ q_lens,
                   const int* __restrict__ k_lens, int max_q_len, int max_k_len,
                   int history_visible, size_t total) {
    pdl_wait();
    for (size_t idx = (size_t)blockIdx.x * blockDim.x + threadIdx.x; idx < total;
         idx += (size_t)gridDim.x * blockDim.x) {
        const int k = (int)(idx % max_k_len);
        size_t r = idx / max_k_len;
        const int q = (int)(r % max_q_len);
        const int b = (int)(r / max_q_len);
        mask[idx] = mask_val<T>(q, k, q_lens[b], k_lens[b], history_visible);
    }
    pdl_launch_dependents();
}

// max_k_len % E == 0: one 16-byte store per thread
template <typename T>
__global__ void __launch_bounds__(256)
causal_mask_vec_kernel(T* __restrict__ mask, const int* __restrict__ q_lens,
                       const int* __restrict__ k_lens, int max_q_len, int max_k_len,
                       int history_visible, size_t total_packets) {
    constexpr int E = DT<T>::kPer16B;
    pdl_wait();
    const int row_packets = max_k_len / E;
    for (size_t p = (size_t)blockIdx.x * blockDim.x + threadIdx.x; p < total_packets;
         p += (size_t)gridDim.x * blockDim.x) {
        const int k0 = (int)(p % row_packets) * E;
        size_t r = p / row_packets;
        const int q = (int)(r % max_q_len);
        const int b = (int)(r / max_q_len);
        const int qlen = q_lens[b], klen = k_lens[b];
        float f[E];
#pragma unroll
        for (int i = 0; i < E; ++i) {
            bool one = q < qlen && (k0 + i) < klen && (k0 + i) <= q + (klen - qlen);
            if (!history_visible) one = one && ((k0 + i) >= klen - qlen);
            f[i] = one ? 1.f : 0.f;
        }
        st16(mask + p * E, Unpack<T>::put(f));
    }
    pdl_launch_dependents();
}

template <typename T>
static int causal_mask_impl(void* mask, const int* q_lens, const int* k_lens, int batch,
                            int max_q_len, int max_k_len, int history_visible, cudaStream_t s) {
    constexpr int E = DT<T>::kPer16B;
    const size_t total = (size_t)batch * max_q_len * max_k_len;
    if (total == 0) return B200_OK;
    const bool vec = (max_k_len % E == 0) && ((reinterpret_cast<uintptr_t>(mask) & 15u) == 0);
    const size_t work = vec ? total / E : total;
    size_t blocks = (work + 255) / 256;
    const size_t cap = (size_t)sm_count() * 8;
    if (blocks > cap) blocks = cap;
    cudaError_t e;
    if (vec)
        e = launch_k(causal_mask_vec_kernel<T>, dim3((unsigned)blocks), dim3(256), 0, s, true,
                     (T*)mask, q_lens, k_lens, max_q_len, max_k_len, history_visible, work);
    else
        e = launch_k(causal_mask_kernel<T>, dim3((unsigned)blocks), dim3(256), 0, s, true, (T*)mask,
                     q_lens, k_lens, max_q_len, max_k_len, history_visible, total);
    count_launch();
    return (int)e;
}

}  // namespace b200

extern "C" int b200_cal_padding_offset(int* padding_offset, int* cum_seqlens, const int* input_lens,
                                       int batch, int max_q_len, void* stream) {
    B200_RETURN_IF(!padding_offset || !cum_seqlens || !input_lens || batch < 0 || max_q_len < 0,
                   B200_EINVAL);
    cudaError_t e = b200::launch_k(b200::padding_offset_kernel, dim3(1), dim3(1024), 0,
                                   (cudaStream_t)stream, true, padding_offset, cum_seqlens,
                                   input_lens, batch, max_q_len);
    b200::count_launch();
    return (int)e;
}

extern "C" int b200_build_causal_mask(void* mask, const int* q_lens, const int* k_lens, int batch,
                                      int max_q_len, int max_k_len, int history_visible, int dtype,
                                      void* stream) {
    B200_RETURN_IF(!mask || !q_lens || !k_lens || batch < 0 || max_q_len < 0 || max_k_len < 0,
                   B200_EINVAL);
    cudaStream_t s = (cudaStream_t)stream;
    if (dtype == B200_F32)
        return b200::causal_mask_impl<float>(mask, q_lens, k_lens, batch, max_q_len, max_k_len,
                                             history_visible, s);
    if (dtype == B200_F16)
        return b200::causal_mask_impl<__half>(mask, q_lens, k_lens, batch, max_q_len, max_k_len,
                                              history_visible, s);
    return B200_EINVAL;
}
