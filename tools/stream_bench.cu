// Microbenchmark behind the decode-step ring design: how fast can 148 persistent CTAs stream a
// buffer into shared memory with 1-D bulk copies (cp.async.bulk, 2 KB per lane, 14 lanes per
// stage, S stages in flight) when the buffer is (a) much larger than L2 (HBM-bound) or (b) L2
// resident; and does a cp.async.bulk.prefetch.L2 issued D stages ahead change the HBM case?
// build: nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o stream_bench stream_bench.cu
#include <cstdio>
#include <cstdlib>
#include <cstdint>
#include <cuda_runtime.h>

__device__ __forceinline__ uint32_t smem_u32(const void* p) { return (uint32_t)__cvta_generic_to_shared(p); }
__device__ __forceinline__ void mbar_init(uint64_t* bar, uint32_t count) {
    asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(smem_u32(bar)), "r"(count));
}
__device__ __forceinline__ void mbar_expect_tx(uint64_t* bar, uint32_t bytes) {
    asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(smem_u32(bar)), "r"(bytes) : "memory");
}
__device__ __forceinline__ void mbar_arrive(uint64_t* bar) {
    asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" ::"r"(smem_u32(bar)) : "memory");
}
__device__ __forceinline__ void mbar_wait(uint64_t* bar, uint32_t parity) {
    uint32_t done;
    do {
        asm volatile("{\n.reg .pred p;\nmbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\nselp.u32 %0, 1, 0, p;\n}\n"
                     : "=r"(done) : "r"(smem_u32(bar)), "r"(parity) : "memory");
    } while (!done);
}
__device__ __forceinline__ void bulk_g2s(void* dst, const void* src, uint32_t bytes, uint64_t* bar) {
    asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];" ::
                 "r"(smem_u32(dst)), "l"(src), "r"(bytes), "r"(smem_u32(bar)) : "memory");
}
__device__ __forceinline__ void bulk_pf(const void* src, uint32_t bytes) {
    asm volatile("cp.async.bulk.prefetch.L2.global [%0], %1;" ::"l"(src), "r"(bytes) : "memory");
}

constexpr int kStage = 28672;   // bytes per ring stage

// plain-load variant: every thread streams 16-byte packets with 8 loads in flight (no TMA, no smem)
__global__ void __launch_bounds__(256, 1) ldg_kernel(const unsigned char* buf, long long n_chunks, int per_cta, float* sink) {
    const long long first = (long long)blockIdx.x * per_cta;
    float acc = 0.f;
    for (int s = 0; s < per_cta; ++s) {
        const long long chunk = (first + s) % n_chunks;
        const uint4* src = reinterpret_cast<const uint4*>(buf + chunk * kStage);
        uint4 v[7];
#pragma unroll
        for (int j = 0; j < 7; ++j) v[j] = __ldcg(src + threadIdx.x + 256 * j);   // 7 x 256 x 16 B = 28 KB
#pragma unroll
        for (int j = 0; j < 7; ++j) acc += __uint_as_float(v[j].x ^ v[j].y ^ v[j].z ^ v[j].w);
    }
    if (acc == 12345.678f) sink[0] = acc;
}

// each CTA streams `per_cta` stages; stage s of CTA c = buffer chunk ((c * per_cta + s) % n_chunks)
__global__ void __launch_bounds__(256, 1) stream_kernel(const unsigned char* buf, long long n_chunks, int per_cta,
                                                       int stages, int pf_ahead, int consume, float* sink, int kLanes,
                                                       int kSeg) {
    extern __shared__ __align__(128) unsigned char smem[];
    uint64_t* full = reinterpret_cast<uint64_t*>(smem + (size_t)stages * kStage);
    uint64_t* empty = full + 16;
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    if (threadIdx.x == 0) {
        for (int s = 0; s < stages; ++s) { mbar_init(&full[s], 1); mbar_init(&empty[s], 7); }
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    }
    __syncthreads();
    const long long first = (long long)blockIdx.x * per_cta;
    if (warp == 7) {
        int slot = 0, pass = 0;
        for (int s = 0; s < per_cta; ++s) {
            if (pass > 0) mbar_wait(&empty[slot], (pass - 1) & 1);
            const long long chunk = (first + s) % n_chunks;
            if (lane == 0) mbar_expect_tx(&full[slot], kLanes * kSeg);
            if (lane < kLanes) bulk_g2s(smem + (size_t)slot * kStage + lane * kSeg, buf + chunk * kStage + lane * kSeg, kSeg, &full[slot]);
            if (pf_ahead > 0 && s + pf_ahead < per_cta && lane < kLanes) {
                const long long pc = (first + s + pf_ahead) % n_chunks;
                bulk_pf(buf + pc * kStage + lane * kSeg, kSeg);
            }
            if (++slot == stages) { slot = 0; ++pass; }
        }
        return;
    }
    int slot = 0;
    uint32_t phase = 0;
    float acc = 0.f;
    for (int s = 0; s < per_cta; ++s) {
        mbar_wait(&full[slot], phase);
        if (consume) {
            const float4* src = reinterpret_cast<const float4*>(smem + (size_t)slot * kStage + warp * 4096);
#pragma unroll
            for (int j = 0; j < 8; ++j) {
                const float4 v = src[lane + 32 * j];
                acc += v.x + v.y + v.z + v.w;
            }
        }
        __syncwarp();
        if (lane == 0) mbar_arrive(&empty[slot]);
        if (++slot == stages) { slot = 0; phase ^= 1; }
    }
    if (acc == 12345.678f) sink[0] = acc;
}

// Does a bulk L2 prefetch make the later bulk copy of the same bytes faster? Each CTA owns
// `n` consecutive stages of a region never touched before; phase 1 (optional) prefetches them,
// everybody sleeps `wait_us`, phase 2 copies them through the ring and records its duration.
__global__ void __launch_bounds__(256, 1) pf_then_copy_kernel(const unsigned char* buf, int n, int stages, int do_pf,
                                                             int pf_bytes, int wait_us, int kLanes, int kSeg,
                                                             unsigned long long* dur) {
    extern __shared__ __align__(128) unsigned char smem[];
    uint64_t* full = reinterpret_cast<uint64_t*>(smem + (size_t)stages * kStage);
    uint64_t* empty = full + 16;
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    if (threadIdx.x == 0) {
        for (int s = 0; s < stages; ++s) { mbar_init(&full[s], 1); mbar_init(&empty[s], 7); }
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    }
    __syncthreads();
    const unsigned char* mine = buf + (size_t)blockIdx.x * n * kStage;
    if (do_pf && warp == 7) {
        const int total = n * kStage;
        for (int off = lane * pf_bytes; off < total; off += 32 * pf_bytes) bulk_pf(mine + off, pf_bytes);
    }
    unsigned long long t0;
    asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t0));
    unsigned long long t = t0;
    while (t - t0 < (unsigned long long)wait_us * 1000ull) asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t));
    __syncthreads();
    asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t0));
    if (warp == 7) {
        int slot = 0, pass = 0;
        for (int s = 0; s < n; ++s) {
            if (pass > 0) mbar_wait(&empty[slot], (pass - 1) & 1);
            if (lane == 0) mbar_expect_tx(&full[slot], kStage);
            if (lane < kLanes) bulk_g2s(smem + (size_t)slot * kStage + lane * kSeg, mine + (size_t)s * kStage + lane * kSeg, kSeg, &full[slot]);
            if (++slot == stages) { slot = 0; ++pass; }
        }
        return;
    }
    int slot = 0;
    uint32_t phase = 0;
    for (int s = 0; s < n; ++s) {
        mbar_wait(&full[slot], phase);
        __syncwarp();
        if (lane == 0) mbar_arrive(&empty[slot]);
        if (++slot == stages) { slot = 0; phase ^= 1; }
    }
    if (threadIdx.x == 0) {
        asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t));
        dur[blockIdx.x] = t - t0;
    }
}

int main(int argc, char** argv) {
    int dev = 0, sms = 0;
    cudaGetDevice(&dev);
    cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev);
    const size_t big = (size_t)6 << 30;
    unsigned char* buf;
    float* sink;
    cudaMalloc(&buf, big);
    cudaMalloc(&sink, 4);
    cudaMemset(buf, 1, big);
    cudaFuncSetAttribute(stream_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, 226 * 1024);
    cudaEvent_t e0, e1;
    cudaEventCreate(&e0);
    cudaEventCreate(&e1);
    printf("SMs %d; stage %d B\n", sms, kStage);
    auto run = [&](long long ws, int grid, int stages, int pf, int consume, int lanes, int seg, bool ldg) {
        const long long n_chunks = ws / kStage;
        const int per_cta = 4096;
        const size_t smem = (size_t)stages * kStage + 256;
        float best = 1e9f;
        for (int rep = 0; rep < 3; ++rep) {
            cudaEventRecord(e0);
            if (ldg) ldg_kernel<<<grid, 256>>>(buf, n_chunks, per_cta, sink);
            else stream_kernel<<<grid, 256, smem>>>(buf, n_chunks, per_cta, stages, pf, consume, sink, lanes, seg);
            cudaEventRecord(e1);
            cudaEventSynchronize(e1);
            float ms;
            cudaEventElapsedTime(&ms, e0, e1);
            if (ms < best) best = ms;
        }
        cudaError_t err = cudaGetLastError();
        const double bytes = (double)grid * per_cta * (ldg ? kStage : lanes * seg);
        printf("%s ws %5lld MB grid %3d stages %d copies %2d x %5d B pf %d consume %d : %7.3f ms %6.0f GB/s (%.1f GB/s per SM) %s\n",
               ldg ? "LDG" : "TMA", ws >> 20, grid, stages, lanes, seg, pf, consume, best, bytes / best / 1e6,
               bytes / best / 1e6 / grid, err ? cudaGetErrorString(err) : "");
    };
    const long long L2 = 48ll << 20, HBM = 4ll << 30;
    for (long long ws : {L2, HBM}) {
        for (int grid : {148, 74, 37, 8}) run(ws, grid, 7, 0, 1, 14, 2048, false);
        run(ws, 148, 7, 0, 1, 1, 28672, false);
        run(ws, 148, 7, 0, 1, 7, 4096, false);
        run(ws, 148, 7, 0, 1, 28, 1024, false);
        run(ws, 148, 7, 0, 1, 2, 8192, false);
        run(ws, 148, 7, 0, 0, 2, 8192, false);
        run(ws, 148, 4, 0, 1, 2, 8192, false);
        for (int grid : {148, 74, 37}) run(ws, grid, 0, 0, 1, 0, 0, true);
    }
    {
        cudaFuncSetAttribute(pf_then_copy_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, 226 * 1024);
        unsigned long long* dur;
        cudaMallocManaged(&dur, 148 * 8);
        size_t region = 0;
        for (int n : {8, 16}) {
            for (int lanes_seg : {0, 1}) {
                const int lanes = lanes_seg ? 7 : 14, seg = lanes_seg ? 4096 : 2048;
                for (int do_pf : {0, 1, 2}) {
                    const int pf_bytes = do_pf == 2 ? 16384 : 2048;
                    double mx = 0, mean = 0;
                    const size_t bytes_per_run = (size_t)sms * n * kStage;
                    region += bytes_per_run;
                    if (region + bytes_per_run > big) region = 0;
                    pf_then_copy_kernel<<<sms, 256, 7 * kStage + 256>>>(buf + region, n, 7, do_pf ? 1 : 0, pf_bytes, 30, lanes, seg, dur);
                    cudaDeviceSynchronize();
                    for (int i = 0; i < sms; ++i) { mx = dur[i] > mx ? dur[i] : mx; mean += dur[i]; }
                    mean /= sms;
                    printf("pf-then-copy: %2d stages/CTA (%.1f MB total) copies %2d x %d B prefetch %s : copy phase mean %.2f us max %.2f us -> %.0f GB/s %s\n",
                           n, bytes_per_run / 1e6, lanes, seg, do_pf == 0 ? "none " : (do_pf == 1 ? "2 KB ops" : "16 KB ops"), mean / 1e3, mx / 1e3,
                           bytes_per_run / mx, cudaGetErrorString(cudaGetLastError()));
                }
            }
        }
    }
    return 0;
}
