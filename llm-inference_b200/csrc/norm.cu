// RMSNorm family: b200_rmsnorm, b200_fused_add_bias_residual_rmsnorm, b200_add_residual.
// Reference math: src/kernels/rmsnorm_kernel.cu:63-115, fused_addresidual_norm.cu:62-128,
// add_residual.cu:9-31 (with SURVEY defects D7-D10 resolved: residual is really saved, the
// block reduce is a correct two-stage reduce, <=1024 threads, half adds out += residual).
//
// Layout: one CTA per token row; each thread keeps up to kCache 16-byte packets of the row in
// registers, so a row is read from HBM exactly once (algorithmic bytes: 2 row reads/writes + gamma).
#include "common.cuh"
#include "launch_count.h"

namespace b200 {

constexpr int kCache = 4;

template <typename T, bool kSaveResidual>
__global__ void __launch_bounds__(1024)
rmsnorm_kernel(T* __restrict__ x, T* __restrict__ residual_out, const T* __restrict__ gamma,
               int hidden, float eps) {
    constexpr int E = DT<T>::kPer16B;
    __shared__ float scratch[33];
    pdl_wait();
    const int packets = hidden / E;
    T* row = x + (size_t)blockIdx.x * hidden;
    T* rrow = kSaveResidual ? residual_out + (size_t)blockIdx.x * hidden : nullptr;

    Pack16 cache[kCache];
    float ss = 0.f;
#pragma unroll
    for (int c = 0; c < kCache; ++c) {
        const int p = threadIdx.x + c * blockDim.x;
        if (p < packets) {
            cache[c] = ld16(row + (size_t)p * E);
            if (kSaveResidual) st16(rrow + (size_t)p * E, cache[c]);
            float f[E];
            Unpack<T>::get(cache[c], f);
#pragma unroll
            for (int i = 0; i < E; ++i) ss = fmaf(f[i], f[i], ss);
        }
    }
    // rows wider than kCache*blockDim packets: stream the remainder (re-read in pass 2)
    for (int p = threadIdx.x + kCache * blockDim.x; p < packets; p += blockDim.x) {
        Pack16 v = ld16(row + (size_t)p * E);
        if (kSaveResidual) st16(rrow + (size_t)p * E, v);
        float f[E];
        Unpack<T>::get(v, f);
#pragma unroll
        for (int i = 0; i < E; ++i) ss = fmaf(f[i], f[i], ss);
    }
    ss = block_sum(ss, scratch);
    const float inv = rsqrtf(ss / (float)hidden + eps);
    pdl_launch_dependents();

#pragma unroll
    for (int c = 0; c < kCache; ++c) {
        const int p = threadIdx.x + c * blockDim.x;
        if (p < packets) {
            float f[E], g[E];
            Unpack<T>::get(cache[c], f);
            Unpack<T>::get(ld16(gamma + (size_t)p * E), g);
#pragma unroll
            for (int i = 0; i < E; ++i) f[i] = (f[i] * inv) * g[i];
            st16(row + (size_t)p * E, Unpack<T>::put(f));
        }
    }
    for (int p = threadIdx.x + kCache * blockDim.x; p < packets; p += blockDim.x) {
        float f[E], g[E];
        Unpack<T>::get(ld16(row + (size_t)p * E), f);
        Unpack<T>::get(ld16(gamma + (size_t)p * E), g);
#pragma unroll
        for (int i = 0; i < E; ++i) f[i] = (f[i] * inv) * g[i];
        st16(row + (size_t)p * E, Unpack<T>::put(f));
    }
}

// residual <- residual + out (+bias), rounded to T; out <- gamma * residual * inv_rms
template <typename T, bool kBias>
__global__ void __launch_bounds__(1024)
fused_add_norm_kernel(T* __restrict__ residual, T* __restrict__ out, const T* __restrict__ bias,
                      const T* __restrict__ gamma, int hidden, float eps) {
    constexpr int E = DT<T>::kPer16B;
    __shared__ float scratch[33];
    pdl_wait();
    const int packets = hidden / E;
    T* rrow = residual + (size_t)blockIdx.x * hidden;
    T* orow = out + (size_t)blockIdx.x * hidden;

    Pack16 cache[kCache];
    float ss = 0.f;
    auto add_row = [&](int p) -> Pack16 {
        float a[E], r[E];
        Unpack<T>::get(ld16(orow + (size_t)p * E), a);
        Unpack<T>::get(ld16(rrow + (size_t)p * E), r);
        if (kBias) {
            float b[E];
            Unpack<T>::get(ld16(bias + (size_t)p * E), b);
#pragma unroll
            for (int i = 0; i < E; ++i) a[i] += b[i];
        }
#pragma unroll
        for (int i = 0; i < E; ++i) r[i] = a[i] + r[i];
        Pack16 packed = Unpack<T>::put(r);  // rounding point: the stored residual
        st16(rrow + (size_t)p * E, packed);
        Unpack<T>::get(packed, r);  // the norm sees the stored (rounded) residual
#pragma unroll
        for (int i = 0; i < E; ++i) ss = fmaf(r[i], r[i], ss);
        return packed;
    };
#pragma unroll
    for (int c = 0; c < kCache; ++c) {
        const int p = threadIdx.x + c * blockDim.x;
        if (p < packets) cache[c] = add_row(p);
    }
    for (int p = threadIdx.x + kCache * blockDim.x; p < packets; p += blockDim.x) add_row(p);

    ss = block_sum(ss, scratch);
    const float inv = rsqrtf(ss / (float)hidden + eps);
    pdl_launch_dependents();

    auto norm_row = [&](int p, const Pack16& rp) {
        float r[E], g[E];
        Unpack<T>::get(rp, r);
        Unpack<T>::get(ld16(gamma + (size_t)p * E), g);
#pragma unroll
        for (int i = 0; i < E; ++i) r[i] = (g[i] * r[i]) * inv;
        st16(orow + (size_t)p * E, Unpack<T>::put(r));
    };
#pragma unroll
    for (int c = 0; c < kCache; ++c) {
        const int p = threadIdx.x + c * blockDim.x;
        if (p < packets) norm_row(p, cache[c]);
    }
    for (int p = threadIdx.x + kCache * blockDim.x; p < packets; p += blockDim.x)
        norm_row(p, ld16(rrow + (size_t)p * E));
}

// out += residual, flat over n_tokens*hidden, grid-stride over packets
template <typename T>
__global__ void __launch_bounds__(256)
add_residual_kernel(const T* __restrict__ residual, T* __restrict__ out, size_t packets) {
    constexpr int E = DT<T>::kPer16B;
    pdl_wait();
    for (size_t p = (size_t)blockIdx.x * blockDim.x + threadIdx.x; p < packets;
         p += (size_t)gridDim.x * blockDim.x) {
        float a[E], r[E];
        Unpack<T>::get(ld16(out + p * E), a);
        Unpack<T>::get(ld16(residual + p * E), r);
#pragma unroll
        for (int i = 0; i < E; ++i) a[i] += r[i];
        st16(out + p * E, Unpack<T>::put(a));
    }
    pdl_launch_dependents();
}

static inline int pick_block(int packets) {
    int t = (packets + kCache - 1) / kCache;
    t = ((t + 31) / 32) * 32;
    if (t < 64) t = 64;
    if (t > 1024) t = 1024;
    return t;
}

static inline bool aligned16(const void* p) { return (reinterpret_cast<uintptr_t>(p) & 15u) == 0; }

template <typename T>
static int rmsnorm_impl(void* x, void* residual_out, const void* gamma, int n_tokens, int hidden,
                        float eps, cudaStream_t s) {
    constexpr int E = DT<T>::kPer16B;
    B200_RETURN_IF(hidden % E != 0, B200_EUNSUPPORTED);
    const int block = pick_block(hidden / E);
    cudaError_t e;
    if (residual_out)
        e = launch_k(rmsnorm_kernel<T, true>, dim3(n_tokens), dim3(block), 0, s, true, (T*)x,
                     (T*)residual_out, (const T*)gamma, hidden, eps);
    else
        e = launch_k(rmsnorm_kernel<T, false>, dim3(n_tokens), dim3(block), 0, s, true, (T*)x,
                     (T*)nullptr, (const T*)gamma, hidden, eps);
    count_launch();
    return (int)e;
}

template <typename T>
static int fused_add_norm_impl(void* residual, void* out, const void* bias, const void* gamma,
                               int n_tokens, int hidden, float eps, cudaStream_t s) {
    constexpr int E = DT<T>::kPer16B;
    B200_RETURN_IF(hidden % E != 0, B200_EUNSUPPORTED);
    const int block = pick_block(hidden / E);
    cudaError_t e;
    if (bias)
        e = launch_k(fused_add_norm_kernel<T, true>, dim3(n_tokens), dim3(block), 0, s, true,
                     (T*)residual, (T*)out, (const T*)bias, (const T*)gamma, hidden, eps);
    else
        e = launch_k(fused_add_norm_kernel<T, false>, dim3(n_tokens), dim3(block), 0, s, true,
                     (T*)residual, (T*)out, (const T*)nullptr, (const T*)gamma, hidden, eps);
    count_launch();
    return (int)e;
}

template <typename T>
static int add_residual_impl(const void* residual, void* out, int n_tokens, int hidden,
                             cudaStream_t s) {
    constexpr int E = DT<T>::kPer16B;
    const size_t total = (size_t)n_tokens * hidden;
    B200_RETURN_IF(total % E != 0, B200_EUNSUPPORTED);
    const size_t packets = total / E;
    size_t blocks = (packets + 255) / 256;
    const size_t cap = (size_t)sm_count() * 8;
    if (blocks > cap) blocks = cap;
    if (blocks == 0) return B200_OK;
    cudaError_t e = launch_k(add_residual_kernel<T>, dim3((unsigned)blocks), dim3(256), 0, s, true,
                             (const T*)residual, (T*)out, packets);
    count_launch();
    return (int)e;
}

}  // namespace b200

using namespace b200;

extern "C" int b200_rmsnorm(void* x, void* residual_out, const void* gamma, int n_tokens,
                            int hidden, float eps, int dtype, void* stream) {
    B200_RETURN_IF(!x || !gamma || n_tokens < 0 || hidden <= 0, B200_EINVAL);
    B200_RETURN_IF(!aligned16(x) || !aligned16(gamma) || !aligned16(residual_out), B200_EINVAL);
    if (n_tokens == 0) return B200_OK;
    cudaStream_t s = (cudaStream_t)stream;
    if (dtype == B200_F32) return rmsnorm_impl<float>(x, residual_out, gamma, n_tokens, hidden, eps, s);
    if (dtype == B200_F16) return rmsnorm_impl<__half>(x, residual_out, gamma, n_tokens, hidden, eps, s);
    return B200_EINVAL;
}

extern "C" int b200_fused_add_bias_residual_rmsnorm(void* residual, void* out, const void* bias,
                                                    const void* gamma, int n_tokens, int hidden,
                                                    float eps, int dtype, void* stream) {
    B200_RETURN_IF(!residual || !out || !gamma || n_tokens < 0 || hidden <= 0, B200_EINVAL);
    B200_RETURN_IF(!aligned16(residual) || !aligned16(out) || !aligned16(gamma) || !aligned16(bias),
                   B200_EINVAL);
    if (n_tokens == 0) return B200_OK;
    cudaStream_t s = (cudaStream_t)stream;
    if (dtype == B200_F32)
        return fused_add_norm_impl<float>(residual, out, bias, gamma, n_tokens, hidden, eps, s);
    if (dtype == B200_F16)
        return fused_add_norm_impl<__half>(residual, out, bias, gamma, n_tokens, hidden, eps, s);
    return B200_EINVAL;
}

extern "C" int b200_add_residual(const void* residual, void* out, int n_tokens, int hidden,
                                 int dtype, void* stream) {
    B200_RETURN_IF(!residual || !out || n_tokens < 0 || hidden <= 0, B200_EINVAL);
    B200_RETURN_IF(!aligned16(residual) || !aligned16(out), B200_EINVAL);
    cudaStream_t s = (cudaStream_t)stream;
    if (dtype == B200_F32) return add_residual_impl<float>(residual, out, n_tokens, hidden, s);
    if (dtype == B200_F16) return add_residual_impl<__half>(residual, out, n_tokens, hidden, s);
    return B200_EINVAL;
}
