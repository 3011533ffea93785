// b200_linear / b200_batched_gemm: entry points behind launchLinearGemm and
// launchLinearStridedBatchGemm (src/kernels/linear.cu:39-104, :127-229).
//
// Dispatch of b200_linear:
//   n_tokens <= 8, W [N,K]           -> weight-streaming GEMV              (gemv.cu)
//   fp16, W [N,K], tensor-core shape -> tcgen05 / TMEM GEMM                (gemm_tcgen05.cu)
//   anything else (fp32, W [K,N], odd sizes) -> the SIMT tiled GEMM below (fp32 FMA, exact
//   fp32 semantics of the reference's CUBLAS_DEFAULT_MATH fp32 path; not a performance path).
#include "common.cuh"
#include "launch_count.h"

namespace b200 {

int gemv_plain(const void* x, const void* W, const void* bias, void* y, int n_tokens, int N, int K,
               int dtype, cudaStream_t s);
// tcgen05 path (gemm_tcgen05.cu); returns B200_EUNSUPPORTED when the shape does not fit
int gemm_tcgen05_f16(const void* x, const void* W, const void* bias, void* y, int M, int N, int K,
                     void* workspace, size_t workspace_bytes, cudaStream_t s);
size_t gemm_tcgen05_workspace_bytes(int M, int N, int K);
int gemm_min_tokens();
int gemm_tcgen05_f16_swiglu(const void* x, const void* W, void* out, int M, int inter, int K, cudaStream_t s);
int gemm_tcgen05_f16_residual(const void* x, const void* W, const void* residual, void* y, int M, int N, int K, cudaStream_t s);

constexpr int BM = 64, BN = 64, BK = 16;

// C[M,N] = A[M,K] * B ; B is [N,K] (TRANS_B) or [K,N]. blockIdx.z = batch index.
template <typename T, bool TRANS_B>
__global__ void __launch_bounds__(256)
gemm_simt_kernel(const T* __restrict__ A, const T* __restrict__ B, const T* __restrict__ bias,
                 T* __restrict__ C, int M, int N, int K, long long sA, long long sB, long long sC) {
    __shared__ float As[BK][BM + 1];
    __shared__ float Bs[BK][BN + 1];
    pdl_wait();
    A += (size_t)blockIdx.z * sA;
    B += (size_t)blockIdx.z * sB;
    C += (size_t)blockIdx.z * sC;
    const int m0 = blockIdx.y * BM, n0 = blockIdx.x * BN;
    const int tx = threadIdx.x % 16, ty = threadIdx.x / 16;  // 16 x 16 threads, 4x4 outputs each
    float acc[4][4];
#pragma unroll
    for (int i = 0; i < 4; ++i)
#pragma unroll
        for (int j = 0; j < 4; ++j) acc[i][j] = 0.f;

    for (int k0 = 0; k0 < K; k0 += BK) {
        // A tile: 64 rows x 16 k  (1024 elements, 4 per thread), k fastest in memory
        for (int i = threadIdx.x; i < BM * BK; i += 256) {
            const int r = i / BK, kk = i - r * BK;
            const int gm = m0 + r, gk = k0 + kk;
            As[kk][r] = (gm < M && gk < K) ? DT<T>::to_f(A[(size_t)gm * K + gk]) : 0.f;
        }
        if (TRANS_B) {
            for (int i = threadIdx.x; i < BN * BK; i += 256) {
                const int c = i / BK, kk = i - c * BK;
                const int gn = n0 + c, gk = k0 + kk;
                Bs[kk][c] = (gn < N && gk < K) ? DT<T>::to_f(B[(size_t)gn * K + gk]) : 0.f;
            }
        } else {
            for (int i = threadIdx.x; i < BN * BK; i += 256) {
                const int kk = i / BN, c = i - kk * BN;
                const int gn = n0 + c, gk = k0 + kk;
                Bs[kk][c] = (gn < N && gk < K) ? DT<T>::to_f(B[(size_t)gk * N + gn]) : 0.f;
            }
        }
        __syncthreads();
#pragma unroll
        for (int kk = 0; kk < BK; ++kk) {
            float a[4], b[4];
#pragma unroll
            for (int i = 0; i < 4; ++i) a[i] = As[kk][ty * 4 + i];
#pragma unroll
            for (int j = 0; j < 4; ++j) b[j] = Bs[kk][tx * 4 + j];
#pragma unroll
            for (int i = 0; i < 4; ++i)
#pragma unroll
                for (int j = 0; j < 4; ++j) acc[i][j] = fmaf(a[i], b[j], acc[i][j]);
        }
        __syncthreads();
    }
    pdl_launch_dependents();
#pragma unroll
    for (int i = 0; i < 4; ++i) {
        const int gm = m0 + ty * 4 + i;
        if (gm >= M) continue;
#pragma unroll
        for (int j = 0; j < 4; ++j) {
            const int gn = n0 + tx * 4 + j;
            if (gn >= N) continue;
            float v = acc[i][j];
            if (bias) v += DT<T>::to_f(bias[gn]);
            C[(size_t)gm * N + gn] = DT<T>::from_f(v);
        }
    }
}

template <typename T>
static int gemm_simt(const void* A, const void* B, const void* bias, void* C, int batch, int M,
                     int N, int K, bool trans_b, cudaStream_t s) {
    if (M == 0 || N == 0 || batch == 0) return B200_OK;
    dim3 grid((N + BN - 1) / BN, (M + BM - 1) / BM, batch);
    B200_RETURN_IF(grid.y > 65535 || grid.z > 65535, B200_EUNSUPPORTED);
    const long long sA = (long long)M * K, sB = (long long)N * K, sC = (long long)M * N;
    cudaError_t e;
    if (trans_b)
        e = launch_k(gemm_simt_kernel<T, true>, grid, dim3(256), 0, s, true, (const T*)A,
                     (const T*)B, (const T*)bias, (T*)C, M, N, K, sA, sB, sC);
    else
        e = launch_k(gemm_simt_kernel<T, false>, grid, dim3(256), 0, s, true, (const T*)A,
                     (const T*)B, (const T*)bias, (T*)C, M, N, K, sA, sB, sC);
    count_launch();
    return (int)e;
}

}  // namespace b200

using namespace b200;

// split-K partials of the tcgen05 weight-streaming mode (fp16, 8 < n_tokens <= 256); 0 otherwise.
// Without a workspace the call still succeeds (single-split schedule, less SM coverage).
extern "C" size_t b200_linear_workspace_bytes(int n_tokens, int N, int K, int trans_b, int dtype) {
    if (!trans_b || dtype != B200_F16) return 0;
    return gemm_tcgen05_workspace_bytes(n_tokens, N, K);
}

extern "C" int b200_linear(const void* x, const void* W, const void* bias, void* y, int n_tokens,
                           int N, int K, int trans_b, int dtype, void* workspace,
                           size_t workspace_bytes, void* stream) {
    B200_RETURN_IF(!x || !W || !y || n_tokens < 0 || N <= 0 || K <= 0, B200_EINVAL);
    B200_RETURN_IF(dtype != B200_F32 && dtype != B200_F16, B200_EINVAL);
    if (n_tokens == 0) return B200_OK;
    cudaStream_t s = (cudaStream_t)stream;
    const int E = dtype == B200_F16 ? 8 : 4;
    auto al = [](const void* p) { return (reinterpret_cast<uintptr_t>(p) & 15u) == 0; };
    // fp16 with 5..8 token rows: at K = 11008 the GEMV's x-rows leave no room for its weight ring (tokens walked in two
    // halves = weights streamed twice); the tcgen05 weight-streaming GEMM reads every weight byte once, so it takes
    // these shapes. Up to 4 rows the GEMV (tensor-core consumer from 2 rows on) is as fast or faster (route_sweep.sh).
    if (trans_b && dtype == B200_F16 && n_tokens >= b200::gemm_min_tokens() && n_tokens <= B200_GEMV_MAX_TOKENS) {
        int rc = gemm_tcgen05_f16(x, W, bias, y, n_tokens, N, K, workspace, workspace_bytes, s);
        if (rc != B200_EUNSUPPORTED) return rc;
    }
    if (trans_b && n_tokens <= B200_GEMV_MAX_TOKENS && K % E == 0 && al(x) && al(W)) {
        int rc = gemv_plain(x, W, bias, y, n_tokens, N, K, dtype, s);
        if (rc != B200_EUNSUPPORTED) return rc;
    }
    if (trans_b && dtype == B200_F16) {
        int rc = gemm_tcgen05_f16(x, W, bias, y, n_tokens, N, K, workspace, workspace_bytes, s);
        if (rc != B200_EUNSUPPORTED) return rc;
    }
    if (dtype == B200_F16)
        return gemm_simt<__half>(x, W, bias, y, 1, n_tokens, N, K, trans_b != 0, s);
    return gemm_simt<float>(x, W, bias, y, 1, n_tokens, N, K, trans_b != 0, s);
}

extern "C" int b200_batched_gemm(const void* A, const void* B, void* C, int batch_count, int m,
                                 int n, int k, int trans_b, int dtype, void* stream) {
    B200_RETURN_IF(!A || !B || !C || batch_count < 0 || m < 0 || n < 0 || k <= 0, B200_EINVAL);
    cudaStream_t s = (cudaStream_t)stream;
    if (dtype == B200_F16)
        return gemm_simt<__half>(A, B, nullptr, C, batch_count, m, n, k, trans_b != 0, s);
    if (dtype == B200_F32)
        return gemm_simt<float>(A, B, nullptr, C, batch_count, m, n, k, trans_b != 0, s);
    return B200_EINVAL;
}

/* out[n_tokens, inter] = silu(x W_gate^T) * (x W_up^T) with W = [2*inter, K] (gate rows first): the gate_up
 * launchLinearGemm (ffn.cpp:72) and launchAct (ffn.cpp:80, act_kernel.cu:57) as ONE tcgen05 GEMM whose epilogue
 * applies SwiGLU; returns B200_EUNSUPPORTED outside fp16 prefill shapes (n_tokens > 256, inter % 128 == 0):
 * the caller then runs b200_linear + b200_swiglu. */
extern "C" int b200_linear_swiglu(const void* x, const void* W, void* out, int n_tokens, int inter, int K, int dtype,
                                  void* stream) {
    if (!x || !W || !out || n_tokens < 1 || inter < 1 || K < 1) return B200_EINVAL;
    if (dtype != B200_F16) return B200_EUNSUPPORTED;
    return b200::gemm_tcgen05_f16_swiglu(x, W, out, n_tokens, inter, K, (cudaStream_t)stream);
}

/* y[n_tokens, N] = residual + x W^T (W [N, K]): launchLinearGemm + launchAddResidual (add_residual.cu:58) in one
 * tcgen05 launch, same rounding points (linear output rounded to fp16, then the sum). y may not alias x;
 * B200_EUNSUPPORTED outside fp16 prefill shapes (n_tokens >= 512): run b200_linear + b200_add_residual then. */
extern "C" int b200_linear_residual(const void* x, const void* W, const void* residual, void* y, int n_tokens, int N,
                                    int K, int dtype, void* stream) {
    if (!x || !W || !residual || !y || n_tokens < 1 || N < 1 || K < 1) return B200_EINVAL;
    if (dtype != B200_F16) return B200_EUNSUPPORTED;
    return b200::gemm_tcgen05_f16_residual(x, W, residual, y, n_tokens, N, K, (cudaStream_t)stream);
}
