// CudaAllocator: caching device allocator behind BaseAllocator.
// Same role as reference src/memory/allocator/cuda_allocator.h:40-316 (per-device pools of
// re-usable blocks, mutex-guarded, host path = malloc), redesigned for 180 GB parts:
//  * sizes are size_t end to end and rounded to 256 B (reference: int size_32, D21);
//  * ONE pool of size-bucketed free lists (power-of-two classes up to 1 MiB, then 2 MiB-aligned
//    exact sizes) instead of two linear-scan vectors; Malloc is O(log n), Free never calls
//    cudaFree (no implicit device sync on the hot path) until the pool is trimmed;
//  * the decode engine does not use it at all on the hot path (persistent arena); it exists for
//    the reference-style layer classes, which allocate scratch per forward call.
#pragma once
#include <cstdlib>
#include <cstring>
#include <map>
#include <mutex>
#include <unordered_map>

#include "src/memory/allocator/base_allocator.h"
#include "src/utils/macro.h"

class CudaAllocator : public BaseAllocator {
    struct DevicePool {
        std::multimap<size_t, void*> free_blocks;      // capacity -> block
        std::unordered_map<void*, size_t> live;        // block -> capacity
        size_t reserved = 0;
        size_t cached = 0;                             // bytes sitting in free_blocks
    };
    std::map<int, DevicePool> pools_;
    std::recursive_mutex mu_;
    size_t trim_threshold_;

    static size_t bucket(size_t n) {
        if (n == 0) n = 1;
        if (n <= (size_t(1) << 20)) {  // small: next power of two >= 256
            size_t b = 256;
            while (b < n) b <<= 1;
            return b;
        }
        const size_t g = size_t(2) << 20;  // big: 2 MiB granules (one TLB page)
        return (n + g - 1) / g * g;
    }

public:
    explicit CudaAllocator(size_t trim_threshold_bytes = size_t(8) << 30) : trim_threshold_(trim_threshold_bytes) {}
    ~CudaAllocator() override { releaseAll(); }

    void* UnifyMalloc(void* ptr, size_t size, bool is_host = false) override {
        (void)ptr;
        if (is_host) {
            void* p = std::malloc(size ? size : 1);
            if (p) std::memset(p, 0, size);
            return p;
        }
        std::lock_guard<std::recursive_mutex> g(mu_);
        int dev = 0;
        CHECK(cudaGetDevice(&dev));
        DevicePool& pool = pools_[dev];
        const size_t cap = bucket(size);
        auto it = pool.free_blocks.lower_bound(cap);
        // accept a cached block only if it wastes < 2x (keeps fragmentation bounded)
        if (it != pool.free_blocks.end() && it->first <= 2 * cap) {
            void* p = it->second;
            pool.live[p] = it->first;
            pool.cached -= it->first;
            pool.free_blocks.erase(it);
            return p;
        }
        void* p = nullptr;
        cudaError_t e = cudaMalloc(&p, cap);
        if (e != cudaSuccess) {  // out of memory: give the cache back and retry once
            cudaGetLastError();
            trim(dev);
            CHECK(cudaMalloc(&p, cap));
        }
        pool.live[p] = cap;
        pool.reserved += cap;
        return p;
    }

    void UnifyFree(void* ptr, bool is_host = false) override {
        if (ptr == nullptr) return;
        if (is_host) { std::free(ptr); return; }
        std::lock_guard<std::recursive_mutex> g(mu_);
        for (auto& kv : pools_) {
            auto it = kv.second.live.find(ptr);
            if (it == kv.second.live.end()) continue;
            kv.second.free_blocks.emplace(it->second, ptr);
            kv.second.cached += it->second;   // running total: no walk over the free list per call
            kv.second.live.erase(it);
            if (kv.second.cached > trim_threshold_) trim(kv.first);
            return;
        }
        CHECK(cudaFree(ptr));  // not ours: behave like the reference's fallback (cuda_allocator.h:300-305)
    }

    void trim(int dev) {
        auto pit = pools_.find(dev);
        if (pit == pools_.end()) return;
        int cur = 0;
        cudaGetDevice(&cur);
        cudaSetDevice(dev);
        for (auto& fb : pit->second.free_blocks) { cudaFree(fb.second); pit->second.reserved -= fb.first; }
        pit->second.free_blocks.clear();
        pit->second.cached = 0;
        cudaSetDevice(cur);
    }
    void releaseAll() {
        std::lock_guard<std::recursive_mutex> g(mu_);
        for (auto& kv : pools_) {
            trim(kv.first);
            for (auto& lv : kv.second.live) cudaFree(lv.first);
            kv.second.live.clear();
        }
    }
    size_t reservedBytes(int dev) { return pools_[dev].reserved; }
};
