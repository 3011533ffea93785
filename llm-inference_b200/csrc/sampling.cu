// Ops bracketing a decode step: embedding gather, greedy argmax, top-K.
// Reference: src/kernels/input_embedding.cu:5-44; src/kernels/topK.cu:25-190 + topK.h:7-47
// (two-pass cub::BlockReduce top-K with K=5; greedy = final_topk_id[b,0]).
// Index outputs are exact: larger value wins, equal values -> lower index wins.
#include "common.cuh"
#include "launch_count.h"

namespace b200 {

template <int PB>
__global__ void __launch_bounds__(256)
embedding_kernel(const int* __restrict__ ids, const char* __restrict__ table,
                 char* __restrict__ out, int row_packets, int vocab, size_t total) {
    pdl_wait();
    for (size_t it = (size_t)blockIdx.x * blockDim.x + threadIdx.x; it < total;
         it += (size_t)gridDim.x * blockDim.x) {
        const size_t t = it / row_packets;
        const int p = (int)(it - t * row_packets);
        int id = ids[t];
        id = id < 0 ? 0 : (id >= vocab ? vocab - 1 : id);  // clamp instead of faulting
        const char* s = table + ((size_t)id * row_packets + p) * PB;
        char* d = out + it * PB;
        if (PB == 16) st16(d, ld16(s));
        else if (PB == 4) *reinterpret_cast<uint32_t*>(d) = *reinterpret_cast<const uint32_t*>(s);
        else *reinterpret_cast<uint16_t*>(d) = *reinterpret_cast<const uint16_t*>(s);
    }
    pdl_launch_dependents();
}

struct ValIdx {
    float v;
    int i;
};
__device__ __forceinline__ ValIdx better(ValIdx a, ValIdx b) {
    return (b.v > a.v || (b.v == a.v && b.i < a.i)) ? b : a;
}
__device__ __forceinline__ ValIdx warp_best(ValIdx x) {
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) {
        ValIdx y;
        y.v = __shfl_xor_sync(0xffffffffu, x.v, o);
        y.i = __shfl_xor_sync(0xffffffffu, x.i, o);
        x = better(x, y);
    }
    return x;
}

// one CTA per row; K rounds of block argmax, each excluding the ids already taken.
// NaNs never win (comparisons are false), an all-NaN row yields index INT_MAX-free fallback 0.
template <typename T, int KMAX>
__global__ void __launch_bounds__(1024)
topk_kernel(const T* __restrict__ logits, int* __restrict__ out_ids, T* __restrict__ out_vals,
            int vocab, int k) {
    __shared__ float sv[32];
    __shared__ int si[32];
    __shared__ int taken[KMAX];
    pdl_wait();
    const T* row = logits + (size_t)blockIdx.x * vocab;
    const int lane = threadIdx.x & 31, wid = threadIdx.x >> 5, nw = blockDim.x >> 5;
    for (int r = 0; r < k; ++r) {
        ValIdx best{-INFINITY, 0x7fffffff};
        for (int j = threadIdx.x; j < vocab; j += blockDim.x) {
            bool skip = false;
            for (int q = 0; q < r; ++q) skip |= (taken[q] == j);
            if (!skip) best = better(best, ValIdx{DT<T>::to_f(row[j]), j});
        }
        best = warp_best(best);
        if (lane == 0) { sv[wid] = best.v; si[wid] = best.i; }
        __syncthreads();
        if (wid == 0) {
            ValIdx x = lane < nw ? ValIdx{sv[lane], si[lane]} : ValIdx{-INFINITY, 0x7fffffff};
            x = warp_best(x);
            if (lane == 0) {
                if (x.i == 0x7fffffff) x.i = 0;
                taken[r] = x.i;
                out_ids[(size_t)blockIdx.x * k + r] = x.i;
                if (out_vals) out_vals[(size_t)blockIdx.x * k + r] = row[x.i];
            }
        }
        __syncthreads();
    }
    pdl_launch_dependents();
}

}  // namespace b200

using namespace b200;

extern "C" int b200_embedding(const int* ids, const void* table, void* out, int n_tokens,
                              int hidden, int vocab, int dtype, void* stream) {
    B200_RETURN_IF(!ids || !table || !out || n_tokens < 0 || hidden <= 0 || vocab <= 0, B200_EINVAL);
    B200_RETURN_IF(dtype != B200_F32 && dtype != B200_F16, B200_EINVAL);
    if (n_tokens == 0) return B200_OK;
    const int es = dtype == B200_F16 ? 2 : 4;
    const size_t row_bytes = (size_t)hidden * es;
    const bool vec = row_bytes % 16 == 0 && (reinterpret_cast<uintptr_t>(table) & 15u) == 0 &&
                     (reinterpret_cast<uintptr_t>(out) & 15u) == 0;
    const int pb = vec ? 16 : es;
    const int row_packets = (int)(row_bytes / pb);
    const size_t total = (size_t)n_tokens * row_packets;
    size_t blocks = (total + 255) / 256;
    const size_t cap = (size_t)sm_count() * 8;
    if (blocks > cap) blocks = cap;
    cudaStream_t s = (cudaStream_t)stream;
    cudaError_t e;
#define CALL(PB)                                                                            \
    e = launch_k(embedding_kernel<PB>, dim3((unsigned)blocks), dim3(256), 0, s, true, ids,  \
                 (const char*)table, (char*)out, row_packets, vocab, total)
    if (pb == 16) CALL(16); else if (pb == 4) CALL(4); else CALL(2);
#undef CALL
    count_launch();
    return (int)e;
}

extern "C" int b200_topk(const void* logits, int* out_ids, void* out_vals, int batch, int vocab,
                         int k, int dtype, void* stream) {
    B200_RETURN_IF(!logits || !out_ids || batch < 0 || vocab <= 0 || k < 1 || k > 8 || k > vocab,
                   B200_EINVAL);
    if (batch == 0) return B200_OK;
    cudaStream_t s = (cudaStream_t)stream;
    int block = vocab >= 8192 ? 1024 : (vocab >= 1024 ? 256 : 64);
    cudaError_t e;
    if (dtype == B200_F16)
        e = launch_k(topk_kernel<__half, 8>, dim3(batch), dim3(block), 0, s, true,
                     (const __half*)logits, out_ids, (__half*)out_vals, vocab, k);
    else if (dtype == B200_F32)
        e = launch_k(topk_kernel<float, 8>, dim3(batch), dim3(block), 0, s, true,
                     (const float*)logits, out_ids, (float*)out_vals, vocab, k);
    else
        return B200_EINVAL;
    count_launch();
    return (int)e;
}

extern "C" int b200_argmax(const void* logits, int* out_ids, int batch, int vocab, int dtype,
                           void* stream) {
    return b200_topk(logits, out_ids, nullptr, batch, vocab, 1, dtype, stream);
}
