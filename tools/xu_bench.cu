// Which pipe bounds the softmax of context_attention.cu? Per-SM throughput of the candidates, in elements per clock:
//   0  ex2.approx.ftz.f32                (MUFU.EX2)
//   1  cvt.rn.f16x2.f32                  (F2FP.F16.F32.PACK_AB)      - per PAIR
//   2  ex2.approx.f16x2                  (MUFU.EX2.F16 packed)       - per PAIR
//   3  2 x ex2.f32 + cvt.f16x2           (today's off-diagonal path) - per PAIR
//   4  cvt.f16x2 + ex2.f16x2             (the dropped f16x2 variant) - per PAIR
//   5  degree-3 polynomial exp2 in fp32 on the FMA / ALU pipes (Cody-Waite split, magic-number floor)
//   6  packed f32x2 polynomial exp2      - per PAIR
//   7  half of the pairs as 3, half as 6 + cvt
// Build: nvcc -O3 -gencode arch=compute_100a,code=sm_100a -o tools/xu_bench tools/xu_bench.cu ; run on the box.
#include <cstdio>
#include <cstdint>
#include <cuda_runtime.h>

__device__ __forceinline__ float ex2f(float x) { float y; asm volatile("ex2.approx.ftz.f32 %0, %1;" : "=f"(y) : "f"(x)); return y; }
__device__ __forceinline__ uint32_t cvt2(float lo, float hi) { uint32_t r; asm volatile("cvt.rn.f16x2.f32 %0, %1, %2;" : "=r"(r) : "f"(hi), "f"(lo)); return r; }
__device__ __forceinline__ uint32_t ex2h2(uint32_t x) { uint32_t y; asm volatile("ex2.approx.f16x2 %0, %1;" : "=r"(y) : "r"(x)); return y; }
__device__ __forceinline__ uint64_t pack2f(float lo, float hi) { uint64_t r; asm("mov.b64 %0, {%1, %2};" : "=l"(r) : "r"(__float_as_uint(lo)), "r"(__float_as_uint(hi))); return r; }
__device__ __forceinline__ void unpack2(uint64_t v, float& lo, float& hi) { uint32_t a, b; asm("mov.b64 {%0, %1}, %2;" : "=r"(a), "=r"(b) : "l"(v)); lo = __uint_as_float(a); hi = __uint_as_float(b); }
__device__ __forceinline__ uint64_t fma2(uint64_t a, uint64_t b, uint64_t c) { uint64_t d; asm volatile("fma.rn.f32x2 %0, %1, %2, %3;" : "=l"(d) : "l"(a), "l"(b), "l"(c)); return d; }
__device__ __forceinline__ uint64_t add2(uint64_t a, uint64_t b) { uint64_t d; asm volatile("add.rn.f32x2 %0, %1, %2;" : "=l"(d) : "l"(a), "l"(b)); return d; }
__device__ __forceinline__ uint64_t add2_rm(uint64_t a, uint64_t b) { uint64_t d; asm volatile("add.rm.f32x2 %0, %1, %2;" : "=l"(d) : "l"(a), "l"(b)); return d; }
__device__ __forceinline__ uint64_t sub2(uint64_t a, uint64_t b) { uint64_t d; asm volatile("sub.rn.f32x2 %0, %1, %2;" : "=l"(d) : "l"(a), "l"(b)); return d; }

// 2^x for x <= 0 (clamped at -126): floor by the 1.5 * 2^23 magic number with round-down, degree-3 minimax of 2^f on [0, 1)
__device__ __forceinline__ float exp2_poly(float x) {
    x = fmaxf(x, -126.f);
    float r;
    asm volatile("add.rm.ftz.f32 %0, %1, %2;" : "=f"(r) : "f"(x), "f"(12582912.f));   // low mantissa bits = floor(x)
    const float fl = r - 12582912.f;
    const float f = x - fl;
    float p = fmaf(f, 0.05550410866f, 0.24022650695f);
    p = fmaf(p, f, 0.69314718056f);
    p = fmaf(p, f, 1.0f);
    return __uint_as_float(__float_as_uint(p) + (__float_as_uint(r) << 23));
}
__device__ __forceinline__ uint64_t exp2_poly2(uint64_t x) {
    const uint64_t magic = pack2f(12582912.f, 12582912.f);
    float x0, x1;
    unpack2(x, x0, x1);
    x = pack2f(fmaxf(x0, -126.f), fmaxf(x1, -126.f));
    const uint64_t r = add2_rm(x, magic);
    const uint64_t f = sub2(x, sub2(r, magic));
    uint64_t p = fma2(f, pack2f(0.05550410866f, 0.05550410866f), pack2f(0.24022650695f, 0.24022650695f));
    p = fma2(p, f, pack2f(0.69314718056f, 0.69314718056f));
    p = fma2(p, f, pack2f(1.f, 1.f));
    uint32_t p0, p1, r0, r1;
    asm("mov.b64 {%0, %1}, %2;" : "=r"(p0), "=r"(p1) : "l"(p));
    asm("mov.b64 {%0, %1}, %2;" : "=r"(r0), "=r"(r1) : "l"(r));
    return pack2f(__uint_as_float(p0 + (r0 << 23)), __uint_as_float(p1 + (r1 << 23)));
}

constexpr int kChains = 8;

template <int MODE>
__global__ void __launch_bounds__(256) bench(float* out, int iters, float seed) {
    float a[kChains], b[kChains];
    uint32_t acc = 0;
#pragma unroll
    for (int c = 0; c < kChains; ++c) { a[c] = -seed * (threadIdx.x + c); b[c] = -seed * (threadIdx.x + 2 * c + 1); }
    for (int it = 0; it < iters; ++it) {
#pragma unroll
        for (int c = 0; c < kChains; ++c) {
            if (MODE == 0) { a[c] = -ex2f(a[c]); b[c] = -ex2f(b[c]); }
            else if (MODE == 1) { acc ^= cvt2(a[c], b[c]); a[c] += seed; b[c] += seed; }
            else if (MODE == 2) { a[c] = __uint_as_float(ex2h2(__float_as_uint(a[c])) ^ 0x80008000u); }
            else if (MODE == 3) { const float p0 = ex2f(a[c]), p1 = ex2f(b[c]); acc ^= cvt2(p0, p1); a[c] -= p0; b[c] -= p1; }
            else if (MODE == 4) { const uint32_t h = ex2h2(cvt2(a[c], b[c])); acc ^= h; a[c] += seed; b[c] += seed; }
            else if (MODE == 5) { a[c] = -exp2_poly(a[c]); b[c] = -exp2_poly(b[c]); }
            else if (MODE == 6) { float p0, p1; unpack2(exp2_poly2(pack2f(a[c], b[c])), p0, p1); a[c] = -p0; b[c] = -p1; }
            else if (MODE == 7) {
                if (c & 1) { const float p0 = ex2f(a[c]), p1 = ex2f(b[c]); acc ^= cvt2(p0, p1); a[c] -= p0; b[c] -= p1; }
                else { float p0, p1; unpack2(exp2_poly2(pack2f(a[c], b[c])), p0, p1); acc ^= cvt2(p0, p1); a[c] -= p0; b[c] -= p1; }
            }
        }
    }
    float s = __uint_as_float(acc & 0x3fffffffu);
#pragma unroll
    for (int c = 0; c < kChains; ++c) s += a[c] + b[c];
    out[blockIdx.x * blockDim.x + threadIdx.x] = s;
}

template <int MODE>
static void run(const char* name, float* out, int sms) {
    const int iters = 4096, blocks = sms * 4;
    bench<MODE><<<blocks, 256>>>(out, 64, 1e-3f);
    cudaEvent_t e0, e1;
    cudaEventCreate(&e0); cudaEventCreate(&e1);
    cudaEventRecord(e0);
    bench<MODE><<<blocks, 256>>>(out, iters, 1e-3f);
    cudaEventRecord(e1);
    cudaEventSynchronize(e1);
    float ms; cudaEventElapsedTime(&ms, e0, e1);
    int khz; cudaDeviceGetAttribute(&khz, cudaDevAttrClockRate, 0);
    const double pairs = (double)blocks * 256 * iters * kChains;   // one PAIR of elements per chain per iteration
    const double clk = ms * 1e-3 * khz * 1e3;
    printf("%-44s %8.3f ms  %7.2f pairs/clk/SM (at the %d MHz nominal clock)\n", name, ms, pairs / clk / sms, khz / 1000);
}

int main() {
    int sms; cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, 0);
    float* out; cudaMalloc(&out, (size_t)sms * 4 * 256 * 4);
    run<0>("0 ex2.f32 x2", out, sms);
    run<1>("1 cvt.f16x2.f32 (+2 FADD)", out, sms);
    run<2>("2 ex2.f16x2", out, sms);
    run<3>("3 2 ex2.f32 + cvt.f16x2 (today)", out, sms);
    run<4>("4 cvt.f16x2 + ex2.f16x2", out, sms);
    run<5>("5 poly exp2 f32 x2", out, sms);
    run<6>("6 poly exp2 f32x2", out, sms);
    run<7>("7 half today / half poly f32x2 + cvt", out, sms);
    printf("%s\n", cudaGetErrorString(cudaDeviceSynchronize()));
    return 0;
}
