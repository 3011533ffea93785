// Common plumbing of the header-only launch* shim: the reference's `launch*` templates
// (src/kernels/*.h) are re-created on top of the C ABI in include/b200_llama.h. These headers
// contain no CUDA syntax, so they compile in plain .cpp translation units (the reference has to
// instantiate its templates in .cu files); linking needs only libb200llama.so.
#pragma once
#include <cuda_fp16.h>
#include <cuda_runtime.h>
#include <type_traits>

#include "b200_llama.h"
#include "src/utils/macro.h"
#include "src/utils/tensor.h"

namespace b200shim {

// The reference launches everything on the legacy default stream (no <<<.., stream>>> anywhere,
// SURVEY §0); that stays the default. A host application that owns a stream sets it once.
inline cudaStream_t& current_stream() {
    static thread_local cudaStream_t s = nullptr;
    return s;
}
inline void set_stream(cudaStream_t s) { current_stream() = s; }
inline void* stream() { return (void*)current_stream(); }

template <typename T>
constexpr int dtype_of() {
    static_assert(std::is_same<T, float>::value || std::is_same<T, half>::value,
                  "the decoder path is instantiated for float and half only");
    return std::is_same<T, half>::value ? B200_F16 : B200_F32;
}

// one persistent scratch buffer per thread for launchers whose C entry point needs workspace
// (the reference's launchers take none); grown on demand, never shrunk.
inline void* workspace(size_t bytes) {
    struct WS {
        void* p = nullptr;
        size_t n = 0;
        ~WS() { if (p) cudaFree(p); }
    };
    static thread_local WS ws;
    if (bytes > ws.n) {
        if (ws.p) CHECK(cudaFree(ws.p));
        CHECK(cudaMalloc(&ws.p, bytes));
        ws.n = bytes;
    }
    return ws.p;
}

}  // namespace b200shim
