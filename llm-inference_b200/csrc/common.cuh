// Shared device helpers for the sm_100a LLaMA decoder kernels.
// Everything here is header-only and included by each kernel translation unit.
#pragma once
#include <cuda.h>
#include <cuda_fp16.h>
#include <cuda_runtime.h>
#include <stdint.h>
#include <mutex>
#include <unordered_set>

#include "b200_llama.h"

namespace b200 {

constexpr int kWarp = 32;

// ---------------------------------------------------------------------------
// dtype traits: T is float or __half; all math is carried out in fp32.
// ---------------------------------------------------------------------------
template <typename T> struct DT;
template <> struct DT<float> {
    static constexpr int kPer16B = 4;
    static __device__ __forceinline__ float to_f(float v) { return v; }
    static __device__ __forceinline__ float from_f(float v) { return v; }
};
template <> struct DT<__half> {
    static constexpr int kPer16B = 8;
    static __device__ __forceinline__ float to_f(__half v) { return __half2float(v); }
    static __device__ __forceinline__ __half from_f(float v) { return __float2half_rn(v); }
};

// 16-byte packet: the unit every bandwidth kernel moves per thread per access.
struct __align__(16) Pack16 {
    uint32_t w[4];
};

__device__ __forceinline__ Pack16 ld16(const void* p) {
    return *reinterpret_cast<const Pack16*>(p);
}
__device__ __forceinline__ void st16(void* p, const Pack16& v) {
    *reinterpret_cast<Pack16*>(p) = v;
}
// streaming (read-once) global load that does not allocate in L1. Read-only (.nc) path: only for
// data that no thread of the running kernel writes.
__device__ __forceinline__ Pack16 ld16_stream(const void* p) {
    Pack16 r;
    asm volatile("ld.global.nc.L1::no_allocate.v4.u32 {%0,%1,%2,%3}, [%4];"
                 : "=r"(r.w[0]), "=r"(r.w[1]), "=r"(r.w[2]), "=r"(r.w[3])
                 : "l"(p));
    return r;
}

// unpack a 16B packet to fp32 lanes (4 for float, 8 for half)
template <typename T> struct Unpack;
template <> struct Unpack<float> {
    static constexpr int N = 4;
    static __device__ __forceinline__ void get(const Pack16& p, float (&f)[4]) {
#pragma unroll
        for (int i = 0; i < 4; ++i) f[i] = __uint_as_float(p.w[i]);
    }
    static __device__ __forceinline__ Pack16 put(const float (&f)[4]) {
        Pack16 p;
#pragma unroll
        for (int i = 0; i < 4; ++i) p.w[i] = __float_as_uint(f[i]);
        return p;
    }
};
template <> struct Unpack<__half> {
    static constexpr int N = 8;
    static __device__ __forceinline__ void get(const Pack16& p, float (&f)[8]) {
#pragma unroll
        for (int i = 0; i < 4; ++i) {
            __half2 h = *reinterpret_cast<const __half2*>(&p.w[i]);
            float2 v = __half22float2(h);
            f[2 * i] = v.x;
            f[2 * i + 1] = v.y;
        }
    }
    static __device__ __forceinline__ Pack16 put(const float (&f)[8]) {
        Pack16 p;
#pragma unroll
        for (int i = 0; i < 4; ++i) {
            __half2 h = __floats2half2_rn(f[2 * i], f[2 * i + 1]);
            p.w[i] = *reinterpret_cast<uint32_t*>(&h);
        }
        return p;
    }
};

// dot of two 16B packets accumulated into fp32.
// half: mixed-precision fma.rn.f32.f16 (SASS FHFMA) -> products are exact in fp32,
// bit-identical to convert-then-FFMA but without the conversions.
__device__ __forceinline__ float fma_h(uint16_t a, uint16_t b, float c) {
#ifdef B200_NO_FHFMA
    return fmaf(__half2float(__ushort_as_half(a)), __half2float(__ushort_as_half(b)), c);
#else
    asm("fma.rn.f32.f16 %0, %1, %2, %0;" : "+f"(c) : "h"(a), "h"(b));
    return c;
#endif
}
template <typename T>
__device__ __forceinline__ float dot16(const Pack16& a, const Pack16& b, float acc);
template <>
__device__ __forceinline__ float dot16<float>(const Pack16& a, const Pack16& b, float acc) {
#pragma unroll
    for (int i = 0; i < 4; ++i) acc = fmaf(__uint_as_float(a.w[i]), __uint_as_float(b.w[i]), acc);
    return acc;
}
template <>
__device__ __forceinline__ float dot16<__half>(const Pack16& a, const Pack16& b, float acc) {
#ifdef B200_NO_FHFMA
#pragma unroll
    for (int i = 0; i < 4; ++i) {
        const float2 x = __half22float2(*reinterpret_cast<const __half2*>(&a.w[i]));
        const float2 y = __half22float2(*reinterpret_cast<const __half2*>(&b.w[i]));
        acc = fmaf(x.x, y.x, acc);
        acc = fmaf(x.y, y.y, acc);
    }
    return acc;
#endif
#pragma unroll
    for (int i = 0; i < 4; ++i) {
        acc = fma_h((uint16_t)(a.w[i] & 0xffffu), (uint16_t)(b.w[i] & 0xffffu), acc);
        acc = fma_h((uint16_t)(a.w[i] >> 16), (uint16_t)(b.w[i] >> 16), acc);
    }
    return acc;
}

// ---------------------------------------------------------------------------
// RoPE angle: pos / base^(2i/rot_dim) (reference qkv_bias_and_RoPE.cu:29-36). The power is
// evaluated in double and rounded once, so it is the correctly rounded float power on every
// implementation (device powf is only ~2 ulp, which at position 4095 already moves the angle by
// 1e-3 rad). The fp32 division and cosf/sinf then see bit-identical angles as the CPU oracle.
// ---------------------------------------------------------------------------
__device__ __forceinline__ float rope_angle(int pair, int rot_dim, float base, float pos) {
    const float ex = (float)(2 * pair) / (float)rot_dim;
    const float den = (float)pow((double)base, (double)ex);
    return pos / den;
}

// ---------------------------------------------------------------------------
// reductions
// ---------------------------------------------------------------------------
__device__ __forceinline__ float warp_sum(float v) {
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
    return v;
}
__device__ __forceinline__ float warp_max(float v) {
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) v = fmaxf(v, __shfl_xor_sync(0xffffffffu, v, o));
    return v;
}
// block-wide sum; `scratch` must hold >= 33 floats. Every thread gets the result.
__device__ __forceinline__ float block_sum(float v, float* scratch) {
    const int lane = threadIdx.x & 31, wid = threadIdx.x >> 5;
    const int nw = (blockDim.x + 31) >> 5;
    v = warp_sum(v);
    __syncthreads();  // protect scratch reuse across calls
    if (lane == 0) scratch[wid] = v;
    __syncthreads();
    float t = (lane < nw) ? scratch[lane] : 0.f;
    t = warp_sum(t);
    return t;
}
__device__ __forceinline__ float block_max(float v, float* scratch) {
    const int lane = threadIdx.x & 31, wid = threadIdx.x >> 5;
    const int nw = (blockDim.x + 31) >> 5;
    v = warp_max(v);
    __syncthreads();
    if (lane == 0) scratch[wid] = v;
    __syncthreads();
    float t = (lane < nw) ? scratch[lane] : -INFINITY;
    t = warp_max(t);
    return t;
}

// ---------------------------------------------------------------------------
// programmatic dependent launch (PDL): a kernel may start while its
// predecessor drains; everything that reads the predecessor's output must sit
// after pdl_wait().
// ---------------------------------------------------------------------------
__device__ __forceinline__ void pdl_wait() { asm volatile("griddepcontrol.wait;" ::: "memory"); }
__device__ __forceinline__ void pdl_launch_dependents() {
    asm volatile("griddepcontrol.launch_dependents;" ::: "memory");
}

// ---------------------------------------------------------------------------
// mbarrier + 1-D bulk async copy (TMA engine, SASS UBLKCP)
// ---------------------------------------------------------------------------
__device__ __forceinline__ uint32_t smem_u32(const void* p) {
    return (uint32_t)__cvta_generic_to_shared(p);
}
__device__ __forceinline__ void mbar_init(uint64_t* bar, uint32_t count) {
    asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(smem_u32(bar)), "r"(count));
}
__device__ __forceinline__ void mbar_fence_init() {
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
}
__device__ __forceinline__ void mbar_expect_tx(uint64_t* bar, uint32_t bytes) {
    asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(smem_u32(bar)),
                 "r"(bytes)
                 : "memory");
}
__device__ __forceinline__ void mbar_arrive(uint64_t* bar) {
    asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" ::"r"(smem_u32(bar)) : "memory");
}
__device__ __forceinline__ void mbar_wait(uint64_t* bar, uint32_t parity) {
    uint32_t done;
    do {
        asm volatile(
            "{\n"
            ".reg .pred p;\n"
            "mbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\n"
            "selp.u32 %0, 1, 0, p;\n"
            "}\n"
            : "=r"(done)
            : "r"(smem_u32(bar)), "r"(parity)
            : "memory");
    } while (!done);
}
// global -> shared bulk copy; bytes % 16 == 0, both addresses 16B aligned.
__device__ __forceinline__ void bulk_g2s(void* smem_dst, const void* gsrc, uint32_t bytes,
                                         uint64_t* bar) {
    asm volatile(
        "cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];" ::
            "r"(smem_u32(smem_dst)),
        "l"(gsrc), "r"(bytes), "r"(smem_u32(bar))
        : "memory");
}

// ---------------------------------------------------------------------------
// host-side helpers
// ---------------------------------------------------------------------------
// Per-DEVICE caches: function attributes and the SM count belong to the device that is current at
// the call, not to the process (a process may drive engines on several GPUs).
constexpr int kMaxDevices = 64;
inline int current_device() {
    int dev = 0;
    cudaGetDevice(&dev);
    return (dev < 0 || dev >= kMaxDevices) ? 0 : dev;
}
inline int sm_count() {
    static int n[kMaxDevices] = {};
    const int dev = current_device();
    if (n[dev] == 0) {
        int v = 0;
        cudaDeviceGetAttribute(&v, cudaDevAttrMultiProcessorCount, dev);
        n[dev] = v > 0 ? v : 148;
    }
    return n[dev];
}
// cudaFuncAttributeMaxDynamicSharedMemorySize once per (kernel instantiation, device). `done` is the
// caller's function-local `static unsigned long long` bit mask (one per instantiation).
inline cudaError_t ensure_max_dynamic_smem(const void* func, int bytes, unsigned long long& done) {
    const int dev = current_device();
    if (done & (1ull << dev)) return cudaSuccess;
    const cudaError_t e = cudaFuncSetAttribute(func, cudaFuncAttributeMaxDynamicSharedMemorySize, bytes);
    if (e == cudaSuccess) done |= 1ull << dev;
    return e;
}

inline bool pdl_enabled() {
    static int v = -1;
    if (v < 0) {
        const char* e = getenv("B200_PDL");
        v = (e && e[0] == '0') ? 0 : 1;
    }
    return v == 1;
}

// Every kernel of the decode path asks for the maximum shared-memory carveout: an SM keeps one
// carveout configuration at a time, so kernels with different preferences cannot be co-resident
// and a PDL-launched successor would have to wait for the SM to drain.
inline void prefer_max_smem_carveout(const void* func) {
    static std::mutex mu;
    static std::unordered_set<const void*> done[kMaxDevices];
    const int dev = current_device();
    std::lock_guard<std::mutex> g(mu);
    if (done[dev].insert(func).second)
        cudaFuncSetAttribute(func, cudaFuncAttributePreferredSharedMemoryCarveout,
                             cudaSharedmemCarveoutMaxShared);
}

// launch with optional programmatic-stream-serialization attribute
template <typename... KArgs, typename... Args>
inline cudaError_t launch_k(void (*kernel)(KArgs...), dim3 grid, dim3 block, size_t smem,
                            cudaStream_t stream, bool pdl, Args... args) {
    prefer_max_smem_carveout(reinterpret_cast<const void*>(kernel));
    cudaLaunchConfig_t cfg{};
    cfg.gridDim = grid;
    cfg.blockDim = block;
    cfg.dynamicSmemBytes = smem;
    cfg.stream = stream;
    cudaLaunchAttribute attr[1];
    int na = 0;
    if (pdl && pdl_enabled()) {
        attr[na].id = cudaLaunchAttributeProgrammaticStreamSerialization;
        attr[na].val.programmaticStreamSerializationAllowed = 1;
        ++na;
    }
    cfg.attrs = attr;
    cfg.numAttrs = na;
    return cudaLaunchKernelEx(&cfg, kernel, static_cast<KArgs>(args)...);
}

#define B200_RETURN_IF(cond, code) \
    do {                           \
        if (cond) return (code);   \
    } while (0)

inline int last_err() { return (int)cudaGetLastError(); }

}  // namespace b200
