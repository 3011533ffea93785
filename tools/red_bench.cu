// How expensive is the L2-side reduction the K-sliced (row-parallel) down / o projections need?
// 148 CTAs x 256 threads each add a private fp32 partial [H = 4096] into ONE global accumulator:
//   mode 0: red.global.add.f32 (scalar), mode 1: red.global.add.v4.f32, mode 2: v4 + per-CTA rotated start row.
// Reports the time from a grid-wide start to the last CTA's reds being visible (kernel time for `reps` rounds).
// build: nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o red_bench red_bench.cu
#include <cstdio>
#include <cstdlib>
#include <cuda_runtime.h>

__device__ __forceinline__ void red_v4(float* p, float a, float b, float c, float d) {
    asm volatile("red.global.add.v4.f32 [%0], {%1,%2,%3,%4};" ::"l"(p), "f"(a), "f"(b), "f"(c), "f"(d) : "memory");
}
__device__ __forceinline__ void red_1(float* p, float a) {
    asm volatile("red.global.add.f32 [%0], %1;" ::"l"(p), "f"(a) : "memory");
}

__global__ void __launch_bounds__(256, 1) red_kernel(float* acc, int H, int mode, int reps, int nacc) {
    const int t = threadIdx.x;
    for (int r = 0; r < reps; ++r) {
        float* a = acc + (size_t)(r % nacc) * H;
        const int rot = mode == 2 ? (blockIdx.x * 28 * 4) % H : 0;
        for (int i = t * 4; i < H; i += 256 * 4) {
            const int j = (i + rot) % H;
            const float v = (float)(blockIdx.x + 1);
            if (mode == 0) { red_1(a + j, v); red_1(a + j + 1, v); red_1(a + j + 2, v); red_1(a + j + 3, v); }
            else red_v4(a + j, v, v, v, v);
        }
        __syncthreads();
        if (t == 0) __threadfence();   // what the grid barrier's release would wait for
        __syncthreads();
    }
}

int main() {
    const int H = 4096, nacc = 8;
    float* acc;
    cudaMalloc(&acc, (size_t)nacc * H * sizeof(float));
    cudaMemset(acc, 0, (size_t)nacc * H * sizeof(float));
    cudaEvent_t e0, e1;
    cudaEventCreate(&e0);
    cudaEventCreate(&e1);
    for (int grid : {1, 37, 148}) {
        for (int mode = 0; mode < 3; ++mode) {
            for (int reps : {1, 64}) {
                red_kernel<<<grid, 256>>>(acc, H, mode, reps, nacc);
                cudaDeviceSynchronize();
                cudaEventRecord(e0);
                red_kernel<<<grid, 256>>>(acc, H, mode, reps, nacc);
                cudaEventRecord(e1);
                cudaEventSynchronize(e1);
                float ms;
                cudaEventElapsedTime(&ms, e0, e1);
                printf("grid %3d mode %d reps %2d: %8.2f us total, %6.2f us per round (%d reds of %d B per CTA)\n", grid, mode,
                       reps, ms * 1e3, ms * 1e3 / reps, mode == 0 ? H : H / 4, mode == 0 ? 4 : 16);
            }
        }
    }
    // correctness of the v4 form: every element must equal the sum over CTAs and rounds
    cudaMemset(acc, 0, (size_t)nacc * H * sizeof(float));
    red_kernel<<<148, 256>>>(acc, H, 2, 1, nacc);
    float h[8];
    cudaMemcpy(h, acc, sizeof(h), cudaMemcpyDeviceToHost);
    printf("check: acc[0..3] = %.0f %.0f %.0f %.0f (want %d)\n", h[0], h[1], h[2], h[3], 148 * 149 / 2);
    printf("%s\n", cudaGetErrorString(cudaGetLastError()));
    return 0;
}
