// Issue rate of the instructions a half-precision GEMV consumer can be built from, per SM, with 8 resident warps
// (the persistent decode kernel's shape): FFMA, fma.rn.f32.f16 (FHFMA), HFMA2, cvt f16x2->f32 + FFMA.
// build: nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o fma_bench fma_bench.cu
#include <cstdio>
#include <cuda_fp16.h>
#include <cuda_runtime.h>

template <int MODE>
__global__ void __launch_bounds__(256, 1) k(float* out, int iters, unsigned seed) {
    float acc[16];
    unsigned w[8];
#pragma unroll
    for (int i = 0; i < 16; ++i) acc[i] = (float)(threadIdx.x + i);
#pragma unroll
    for (int i = 0; i < 8; ++i) w[i] = seed * (threadIdx.x + 7 * i + 1) | 0x3c003c00u;
    const unsigned short xs = (unsigned short)(0x3c00 + (seed & 15));
    const float xf = 1.0f + (seed & 7);
    for (int it = 0; it < iters; ++it) {
#pragma unroll
        for (int i = 0; i < 8; ++i) {
            if (MODE == 0) {   // FFMA x2
                acc[2 * i] = fmaf(__uint_as_float(w[i]), xf, acc[2 * i]);
                acc[2 * i + 1] = fmaf(__uint_as_float(w[i] ^ 0x1000u), xf, acc[2 * i + 1]);
            } else if (MODE == 1) {   // FHFMA x2
                asm("fma.rn.f32.f16 %0, %1, %2, %0;" : "+f"(acc[2 * i]) : "h"((unsigned short)(w[i] & 0xffff)), "h"(xs));
                asm("fma.rn.f32.f16 %0, %1, %2, %0;" : "+f"(acc[2 * i + 1]) : "h"((unsigned short)(w[i] >> 16)), "h"(xs));
            } else if (MODE == 2) {   // HFMA2 (fp16 accumulate): 2 MACs per instruction
                __half2 a = *reinterpret_cast<__half2*>(&acc[2 * i]);
                a = __hfma2(*reinterpret_cast<__half2*>(&w[i]), __half2half2(__ushort_as_half(xs)), a);
                acc[2 * i] = *reinterpret_cast<float*>(&a);
                __half2 b = *reinterpret_cast<__half2*>(&acc[2 * i + 1]);
                b = __hfma2(*reinterpret_cast<__half2*>(&w[i]), __half2half2(__ushort_as_half(xs)), b);
                acc[2 * i + 1] = *reinterpret_cast<float*>(&b);
            } else {   // cvt pair + 2 FFMA
                const float2 f = __half22float2(*reinterpret_cast<__half2*>(&w[i]));
                acc[2 * i] = fmaf(f.x, xf, acc[2 * i]);
                acc[2 * i + 1] = fmaf(f.y, xf, acc[2 * i + 1]);
            }
            w[i] += 0x00010001u * (MODE == 2 ? 0 : 1) * 0;   // keep operands loop-invariant but opaque
        }
    }
    float s = 0.f;
#pragma unroll
    for (int i = 0; i < 16; ++i) s += acc[i];
    out[blockIdx.x * blockDim.x + threadIdx.x] = s;
}

template <int MODE>
void run(const char* name, float* out) {
    const int iters = 20000;
    cudaEvent_t e0, e1;
    cudaEventCreate(&e0);
    cudaEventCreate(&e1);
    k<MODE><<<148, 256>>>(out, 100, 3);
    cudaEventRecord(e0);
    k<MODE><<<148, 256>>>(out, iters, 3);
    cudaEventRecord(e1);
    cudaEventSynchronize(e1);
    float ms;
    cudaEventElapsedTime(&ms, e0, e1);
    const double macs_per_sm = (double)iters * 16 * 256 * (MODE == 2 ? 2 : 1);
    printf("%-28s %8.3f ms  -> %6.1f MAC lanes per clock per SM (at 1.965 GHz), %5.2f warp-instr/clk/SM\n", name, ms,
           macs_per_sm / (ms * 1e-3) / 1.965e9, (double)iters * 16 * 8 / (ms * 1e-3) / 1.965e9);
}

int main() {
    float* out;
    cudaMalloc(&out, 148 * 256 * sizeof(float));
    run<0>("FFMA", out);
    run<1>("fma.rn.f32.f16 (FHFMA)", out);
    run<2>("HFMA2 (fp16 accumulate)", out);
    run<3>("cvt f16x2->f32x2 + 2 FFMA", out);
    printf("%s\n", cudaGetErrorString(cudaGetLastError()));
    return 0;
}
