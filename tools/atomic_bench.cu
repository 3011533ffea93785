// How fast are contended global atomics? 148 CTAs, one thread each, N dependent atomicAdds on
// (a) one address, (b) 8 addresses in different 128-byte lines, (c) one address per CTA.
#include <cstdio>
#include <cuda_runtime.h>
__global__ void k(unsigned* ctr, int stride_mask, int n, unsigned long long* dur) {
    unsigned* a = ctr + (size_t)(blockIdx.x & stride_mask) * 32;
    unsigned long long t0, t1;
    asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t0));
    unsigned v = 0;
    for (int i = 0; i < n; ++i) v += atomicAdd(a, 1u + (v & 1u));   // dependent chain
    asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t1));
    dur[blockIdx.x] = t1 - t0;
    if (v == 0xffffffffu) ctr[1] = v;
}
int main() {
    unsigned* ctr; unsigned long long* dur;
    cudaMalloc(&ctr, 1 << 20); cudaMemset(ctr, 0, 1 << 20);
    cudaMallocManaged(&dur, 148 * 8);
    for (int grid : {1, 16, 148}) for (int mask : {0, 7, 255}) {
        const int n = 2000;
        k<<<grid, 1>>>(ctr, mask, n, dur); cudaDeviceSynchronize();
        k<<<grid, 1>>>(ctr, mask, n, dur); cudaDeviceSynchronize();
        double mx = 0; for (int i = 0; i < grid; ++i) mx = dur[i] > mx ? dur[i] : mx;
        printf("grid %3d, %3d distinct addresses: %.0f ns per dependent atomic per CTA; aggregate %.1f M atomics/s\n", grid,
               mask == 0 ? 1 : (mask == 7 ? (grid < 8 ? grid : 8) : grid), mx / n, grid * (double)n / mx * 1e3);
    }
    return 0;
}
