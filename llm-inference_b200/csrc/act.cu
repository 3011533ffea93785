// SwiGLU: out = silu(gate) * up, silu(x) = x / (1 + exp(-x)).
// Reference math: src/kernels/act_kernel.cu:7-31 (fp32), :34-53 (half, defect D11 resolved).
// gate_up is [n_tokens, 2, inter]: gate block then up block per token. 3*n_tokens*inter*T bytes.
#include "common.cuh"
#include "launch_count.h"

namespace b200 {

__device__ __forceinline__ float silu_f(float x) { return x / (1.f + expf(-x)); }

template <typename T>
__global__ void __launch_bounds__(256)
swiglu_vec_kernel(const T* __restrict__ gate_up, T* __restrict__ out, int inter_packets,
                  size_t total_packets) {
    constexpr int E = DT<T>::kPer16B;
    pdl_wait();
    for (size_t p = (size_t)blockIdx.x * blockDim.x + threadIdx.x; p < total_packets;
         p += (size_t)gridDim.x * blockDim.x) {
        const size_t tok = p / inter_packets;
        const size_t col = p - tok * inter_packets;
        const T* g = gate_up + (tok * 2 * inter_packets + col) * E;
        const T* u = g + (size_t)inter_packets * E;
        float gf[E], uf[E];
        Unpack<T>::get(ld16_stream(g), gf);
        Unpack<T>::get(ld16_stream(u), uf);
#pragma unroll
        for (int i = 0; i < E; ++i) {
            // reference rounds silu(gate) to T before the multiply (act_kernel.cu:10,30)
            float s = DT<T>::to_f(DT<T>::from_f(silu_f(gf[i])));
            gf[i] = s * uf[i];
        }
        st16(out + p * E, Unpack<T>::put(gf));
    }
    pdl_launch_dependents();
}

// any inter (no alignment requirement)
template <typename T>
__global__ void __launch_bounds__(256)
swiglu_scalar_kernel(const T* __restrict__ gate_up, T* __restrict__ out, int inter, size_t total) {
    pdl_wait();
    for (size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x; i < total;
         i += (size_t)gridDim.x * blockDim.x) {
        const size_t tok = i / inter, col = i - tok * inter;
        const float g = DT<T>::to_f(gate_up[tok * 2 * inter + col]);
        const float u = DT<T>::to_f(gate_up[tok * 2 * inter + inter + col]);
        const float s = DT<T>::to_f(DT<T>::from_f(silu_f(g)));
        out[i] = DT<T>::from_f(s * u);
    }
    pdl_launch_dependents();
}

template <typename T>
static int swiglu_impl(const void* gate_up, void* out, int n_tokens, int inter, cudaStream_t s) {
    constexpr int E = DT<T>::kPer16B;
    const bool vec = (inter % E == 0) && ((reinterpret_cast<uintptr_t>(gate_up) & 15u) == 0) &&
                     ((reinterpret_cast<uintptr_t>(out) & 15u) == 0);
    const size_t total = (size_t)n_tokens * inter;
    if (total == 0) return B200_OK;
    const size_t work = vec ? total / E : total;
    size_t blocks = (work + 255) / 256;
    const size_t cap = (size_t)sm_count() * 8;
    if (blocks > cap) blocks = cap;
    cudaError_t e;
    if (vec)
        e = launch_k(swiglu_vec_kernel<T>, dim3((unsigned)blocks), dim3(256), 0, s, true,
                     (const T*)gate_up, (T*)out, inter / E, work);
    else
        e = launch_k(swiglu_scalar_kernel<T>, dim3((unsigned)blocks), dim3(256), 0, s, true,
                     (const T*)gate_up, (T*)out, inter, total);
    count_launch();
    return (int)e;
}

}  // namespace b200

extern "C" int b200_swiglu(const void* gate_up, void* out, int n_tokens, int inter, int dtype,
                           void* stream) {
    B200_RETURN_IF(!gate_up || !out || n_tokens < 0 || inter <= 0, B200_EINVAL);
    cudaStream_t s = (cudaStream_t)stream;
    if (dtype == B200_F32) return b200::swiglu_impl<float>(gate_up, out, n_tokens, inter, s);
    if (dtype == B200_F16) return b200::swiglu_impl<__half>(gate_up, out, n_tokens, inter, s);
    return B200_EINVAL;
}
