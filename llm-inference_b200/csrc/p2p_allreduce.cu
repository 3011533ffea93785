// One-shot NVLink allreduce fused with the residual add, for tensor-parallel decode:
// after the row-parallel o_proj / down_proj every rank holds a partial [B, H]; the reference has no
// multi-GPU path, the baseline is ncclAllReduce + launchAddResidual (add_residual.cu:58). Messages
// are 16 KB (B=1, 70B) .. 512 KB (B=32): latency-bound. Every rank PUSHES its partial straight into a
// slot of every peer's receive buffer with peer stores over NVLink (cudaIpc-mapped memory) and sums
// the `world` slots locally in rank order (all ranks form bit-identical sums), adding the result to
// `hidden` in the same kernel: one launch, one NVLink traversal, no ring, no second kernel.
//
// Wire format ("low latency" lines): a 16-byte store carries 8 bytes of payload and the call number
// twice, {data0, call, data1, call}. The receiver spins on the line itself until both tags equal the
// current call, so there is no separate flag, no system-scope fence and no barrier on the critical
// path: a line is valid as soon as it is visible. (First version, measured on 2 x B200: payload
// stores + __threadfence_system + release flag + acquire spin = 11.3 us per 16 KB call, no better
// than NCCL's 11.8 us; profiles/r01_allreduce_bench_n2.json.)
//
// Layout of a rank's IPC allocation: [header: call counter, ticket][receive buffer 0: world slots]
// [receive buffer 1: world slots]; a slot holds 2 x max_message_bytes of lines. The two receive
// buffers alternate per call: a peer may already push call e+1 while this rank still sums call e; it
// cannot reach e+2 before it has received this rank's e+1 lines, which are sent only after this
// rank's kernel e has finished. The call counter is device resident (CUDA-graph replay), read by
// every CTA at its start and advanced by the last CTA to finish.
#include <vector>

#include "b200_engine.h"
#include "common.cuh"
#include "launch_count.h"
#include "p2p_internal.h"

namespace b200 {
using P2PParams = P2PWire;

__device__ __forceinline__ void st16_volatile(void* p, uint32_t a, uint32_t b, uint32_t c, uint32_t d) {
    asm volatile("st.volatile.global.v4.u32 [%0], {%1,%2,%3,%4};" ::"l"(p), "r"(a), "r"(b), "r"(c), "r"(d) : "memory");
}
__device__ __forceinline__ Pack16 ld16_volatile(const void* p) {
    Pack16 r;
    asm volatile("ld.volatile.global.v4.u32 {%0,%1,%2,%3}, [%4];"
                 : "=r"(r.w[0]), "=r"(r.w[1]), "=r"(r.w[2]), "=r"(r.w[3])
                 : "l"(p)
                 : "memory");
    return r;
}
__device__ __forceinline__ uint2 ld8(const void* p) { return *reinterpret_cast<const uint2*>(p); }

template <typename T> struct Line;   // 8 payload bytes <-> fp32 lanes
template <> struct Line<float> {
    static constexpr int N = 2;
    static __device__ __forceinline__ void get(uint32_t a, uint32_t b, float (&f)[2]) { f[0] = __uint_as_float(a); f[1] = __uint_as_float(b); }
    static __device__ __forceinline__ uint2 put(const float (&f)[2]) { return make_uint2(__float_as_uint(f[0]), __float_as_uint(f[1])); }
    static __device__ __forceinline__ float round(float v) { return v; }
};
template <> struct Line<__half> {
    static constexpr int N = 4;
    static __device__ __forceinline__ void get(uint32_t a, uint32_t b, float (&f)[4]) {
        const float2 x = __half22float2(*reinterpret_cast<const __half2*>(&a));
        const float2 y = __half22float2(*reinterpret_cast<const __half2*>(&b));
        f[0] = x.x; f[1] = x.y; f[2] = y.x; f[3] = y.y;
    }
    static __device__ __forceinline__ uint2 put(const float (&f)[4]) {
        const __half2 x = __floats2half2_rn(f[0], f[1]), y = __floats2half2_rn(f[2], f[3]);
        return make_uint2(*reinterpret_cast<const uint32_t*>(&x), *reinterpret_cast<const uint32_t*>(&y));
    }
    static __device__ __forceinline__ float round(float v) { return __half2float(__float2half_rn(v)); }
};

// ADD: hidden[i] <- round(hidden[i] + round(sum_r partial_r[i])): the two roundings of "allreduce, then
// launchAddResidual" (oracle/model.py _ffn) are kept so TP results match the unsharded oracle.
// !ADD: plain in-place allreduce, hidden (== partial) <- round(sum).
// !PUSH: the lines of this call were already pushed by the producing GEMV's epilogue (gemv.cu, B200_GEMV_P2P).
template <typename T, bool ADD, bool PUSH>
__global__ void __launch_bounds__(kP2PThreads)
p2p_allreduce_add_kernel(const P2PParams p, const T* partial, T* hidden, size_t lines) {
    const int me = p.rank;
    unsigned* hdr = reinterpret_cast<unsigned*>(p.peer[me]);
    pdl_launch_dependents();
    pdl_wait();   // `partial` is the preceding GEMV's output; a preceding allreduce has advanced the counter
    const unsigned call = *reinterpret_cast<volatile unsigned*>(hdr) + 1u;   // same value in every CTA of this launch
    const size_t buf_off = kP2PHeader + (size_t)(call & 1u) * p.world * p.slot_bytes;
    const size_t tid = (size_t)blockIdx.x * kP2PThreads + threadIdx.x, nthr = (size_t)gridDim.x * kP2PThreads;
    // ---- push: line i of this rank's partial into slot `me` of every rank (its own included)
    for (size_t i = tid; PUSH && i < lines; i += nthr) {
        const uint2 v = ld8(reinterpret_cast<const unsigned char*>(partial) + i * 8);
#pragma unroll 1
        for (int r = 0; r < p.world; ++r)
            st16_volatile(p.peer[r] + buf_off + (size_t)me * p.slot_bytes + i * 16, v.x, call, v.y, call);
    }
    // ---- receive + local sum in rank order, fused residual add
    constexpr int N = Line<T>::N;
    for (size_t i = tid; i < lines; i += nthr) {
        float acc[N];
#pragma unroll
        for (int e = 0; e < N; ++e) acc[e] = 0.f;
        for (int r = 0; r < p.world; ++r) {
            const unsigned char* src = p.peer[me] + buf_off + (size_t)r * p.slot_bytes + i * 16;
            Pack16 v = ld16_volatile(src);
            unsigned spins = 0;
            unsigned long long t0 = 0;
            while (v.w[1] != call || v.w[3] != call) {
                if ((++spins & 0x3FFu) == 0) {
                    unsigned long long now;
                    asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(now));
                    if (t0 == 0) t0 = now;
                    else if (now - t0 > 10000000000ull) __trap();   // a peer never arrived: fail instead of hanging
                }
                v = ld16_volatile(src);
            }
            float f[N];
            Line<T>::get(v.w[0], v.w[2], f);
#pragma unroll
            for (int e = 0; e < N; ++e) acc[e] += f[e];
        }
        if (ADD) {
            const uint2 hv = ld8(reinterpret_cast<const unsigned char*>(hidden) + i * 8);
            float h[N];
            Line<T>::get(hv.x, hv.y, h);
#pragma unroll
            for (int e = 0; e < N; ++e) acc[e] = h[e] + Line<T>::round(acc[e]);
        }
        *reinterpret_cast<uint2*>(reinterpret_cast<unsigned char*>(hidden) + i * 8) = Line<T>::put(acc);
    }
    // ---- the last CTA to finish advances the call counter (every CTA has read it by then)
    __syncthreads();
    if (threadIdx.x == 0) {
        const unsigned t = atomicAdd(hdr + 1, 1u);
        if (t == gridDim.x - 1) {
            hdr[1] = 0;
            __threadfence();
            *reinterpret_cast<volatile unsigned*>(hdr) = call;
        }
    }
}
}  // namespace b200

using namespace b200;

extern "C" size_t b200_p2p_handle_bytes(void) { return sizeof(cudaIpcMemHandle_t); }

extern "C" b200_p2p* b200_p2p_create(int world, int rank, size_t max_message_bytes) {
    if (world < 1 || world > kP2PMaxWorld || rank < 0 || rank >= world || max_message_bytes == 0) return nullptr;
    b200_p2p* c = new b200_p2p();
    c->world = world;
    c->rank = rank;
    c->slot_bytes = (2 * max_message_bytes + 255) & ~(size_t)255;   // 16-byte lines carry 8 payload bytes
    c->total = kP2PHeader + 2 * (size_t)world * c->slot_bytes;
    if (cudaMalloc(&c->local, c->total) != cudaSuccess || cudaMemset(c->local, 0, c->total) != cudaSuccess ||
        cudaDeviceSynchronize() != cudaSuccess) {
        cudaFree(c->local);
        delete c;
        return nullptr;
    }
    c->peers.assign(world, nullptr);
    c->opened.assign(world, false);
    c->peers[rank] = c->local;
    return c;
}

extern "C" int b200_p2p_local_handle(b200_p2p* c, void* out_handle) {
    if (!c || !out_handle) return B200_EINVAL;
    cudaIpcMemHandle_t h;
    cudaError_t e = cudaIpcGetMemHandle(&h, c->local);
    if (e != cudaSuccess) return (int)e;
    memcpy(out_handle, &h, sizeof(h));
    return 0;
}

/* all_handles: world x b200_p2p_handle_bytes(), rank-major (every rank's b200_p2p_local_handle) */
extern "C" int b200_p2p_connect(b200_p2p* c, const void* all_handles) {
    if (!c || !all_handles) return B200_EINVAL;
    for (int r = 0; r < c->world; ++r) {
        if (r == c->rank || c->opened[r]) continue;
        cudaIpcMemHandle_t h;
        memcpy(&h, (const char*)all_handles + (size_t)r * sizeof(h), sizeof(h));
        void* ptr = nullptr;
        cudaError_t e = cudaIpcOpenMemHandle(&ptr, h, cudaIpcMemLazyEnablePeerAccess);
        if (e != cudaSuccess) return (int)e;
        c->peers[r] = (unsigned char*)ptr;
        c->opened[r] = true;
    }
    c->connected = true;
    return 0;
}

extern "C" void b200_p2p_destroy(b200_p2p* c) {
    if (!c) return;
    cudaDeviceSynchronize();
    for (int r = 0; r < c->world; ++r)
        if (c->opened[r]) cudaIpcCloseMemHandle(c->peers[r]);
    cudaFree(c->local);
    delete c;
}

/* hidden[count] += sum over ranks of partial[count]; every rank must make the same sequence of calls */
static int p2p_launch(b200_p2p* c, const void* partial, void* hidden, size_t count, int dtype, void* stream, bool add,
                      bool push = true) {
    if (!c || !partial || !hidden) return B200_EINVAL;
    if (!c->connected && c->world > 1) return B200_EINVAL;
    const size_t es = dtype == B200_F16 ? 2 : 4;
    const size_t bytes = count * es;
    if (bytes == 0) return 0;
    if (2 * bytes > c->slot_bytes || (bytes & 7) || (reinterpret_cast<uintptr_t>(partial) & 7) ||
        (reinterpret_cast<uintptr_t>(hidden) & 7))
        return B200_EUNSUPPORTED;
    const size_t lines = bytes / 8;
    size_t ctas = (lines + kP2PThreads - 1) / kP2PThreads;
    if (ctas > (size_t)kP2PMaxCtas) ctas = kP2PMaxCtas;
    const P2PParams p = p2p_wire(c);
    cudaError_t e;
    const dim3 g((unsigned)ctas), b(kP2PThreads);
    cudaStream_t s = (cudaStream_t)stream;
    if (dtype == B200_F16) {
        if (!push) e = launch_k(p2p_allreduce_add_kernel<__half, true, false>, g, b, 0, s, true, p, (const __half*)partial, (__half*)hidden, lines);
        else if (add) e = launch_k(p2p_allreduce_add_kernel<__half, true, true>, g, b, 0, s, true, p, (const __half*)partial, (__half*)hidden, lines);
        else e = launch_k(p2p_allreduce_add_kernel<__half, false, true>, g, b, 0, s, true, p, (const __half*)partial, (__half*)hidden, lines);
    } else {
        if (!push) return B200_EUNSUPPORTED;
        if (add) e = launch_k(p2p_allreduce_add_kernel<float, true, true>, g, b, 0, s, true, p, (const float*)partial, (float*)hidden, lines);
        else e = launch_k(p2p_allreduce_add_kernel<float, false, true>, g, b, 0, s, true, p, (const float*)partial, (float*)hidden, lines);
    }
    count_launch();
    return (int)e;
}
extern "C" int b200_p2p_allreduce_add(b200_p2p* c, const void* partial, void* hidden, size_t count, int dtype,
                                      void* stream) {
    return p2p_launch(c, partial, hidden, count, dtype, stream, true);
}
/* second half of the fused GEMV + allreduce: the partials of this call were pushed by b200_gemv_fused with
 * args->p2p set (fp16); hidden[count] += sum over ranks */
extern "C" int b200_p2p_reduce_add(b200_p2p* c, void* hidden, size_t count, int dtype, void* stream) {
    return p2p_launch(c, hidden, hidden, count, dtype, stream, true, false);
}
/* buf[count] <- sum over ranks of buf[count], in place */
extern "C" int b200_p2p_allreduce(b200_p2p* c, void* buf, size_t count, int dtype, void* stream) {
    return p2p_launch(c, buf, buf, count, dtype, stream, false);
}
extern "C" size_t b200_p2p_max_message_bytes(const b200_p2p* c) { return c ? c->slot_bytes / 2 : 0; }
