// Error-handling conventions of the reference (src/utils/macro.h), kept source-compatible:
//   CHECK(call)                      raw CUDA API failure -> print + exit(1)      (:18-32)
//   CHECK_CUBLAS(val) / check()      non-zero status -> std::runtime_error        (:66-92)
//   DeviceSyncAndCheckCudaError()    device sync + last error -> runtime_error    (:98-109)
//   LLM_CHECK / LLM_CHECK_WITH_INFO  "[oneLLM][ERROR] <info> Assertion fail: file:line"  (:113-133)
// plus B200_CHECK_RC for the int codes the b200_* C ABI returns.
#pragma once
#include <cuda.h>
#include <cuda_runtime.h>
#include <cstdio>
#include <cstdlib>
#include <stdexcept>
#include <string>

#define CHECK(call)                                                                  \
    do {                                                                             \
        const cudaError_t b200_ec_ = (call);                                         \
        if (b200_ec_ != cudaSuccess) {                                               \
            std::printf("CUDA Error:\n    File:       %s\n    Line:       %d\n", __FILE__, __LINE__); \
            std::printf("    Error code: %d\n    Error text: %s\n", (int)b200_ec_,   \
                        cudaGetErrorString(b200_ec_));                               \
            std::exit(1);                                                            \
        }                                                                            \
    } while (0)

inline const char* _cudaGetErrorEnum(cudaError_t error) { return cudaGetErrorString(error); }

template <typename T>
void check(T result, char const* const func, const char* const file, int const line) {
    (void)func;
    if (result) {
        throw std::runtime_error(std::string("[TM][ERROR] CUDA runtime error: ") + _cudaGetErrorEnum(result) +
                                 " " + file + ":" + std::to_string(line) + " \n");
    }
}
#define CHECK_CUBLAS(val) check((val), #val, __FILE__, __LINE__)

inline void syncAndCheck(const char* const file, int const line) {
    cudaDeviceSynchronize();
    const cudaError_t result = cudaGetLastError();
    if (result) {
        throw std::runtime_error(std::string("[TM][ERROR] CUDA runtime error: ") + cudaGetErrorString(result) +
                                 " " + file + ":" + std::to_string(line) + " \n");
    }
}
#define DeviceSyncAndCheckCudaError() syncAndCheck(__FILE__, __LINE__)

[[noreturn]] inline void throwRuntimeError(const char* const file, int const line, std::string const& info = "") {
    throw std::runtime_error(std::string("[oneLLM][ERROR] ") + info + " Assertion fail: " + file + ":" +
                             std::to_string(line) + " \n");
}
inline void llmAssert(bool result, const char* const file, int const line, std::string const& info = "") {
    if (!result) throwRuntimeError(file, line, info);
}
#define LLM_CHECK(val) llmAssert(val, __FILE__, __LINE__)
#define LLM_CHECK_WITH_INFO(val, info)                                   \
    do {                                                                 \
        const bool b200_ok_ = (val);                                     \
        if (!b200_ok_) llmAssert(b200_ok_, __FILE__, __LINE__, (info));  \
    } while (0)

// b200_* C-ABI return code -> the same exception type the reference launchers raise
inline void b200CheckRc(int rc, const char* what, const char* file, int line) {
    if (rc == 0) return;
    std::string info = std::string(what) + (rc > 0 ? std::string(": CUDA error ") + cudaGetErrorString((cudaError_t)rc)
                                            : rc == -1 ? ": invalid argument" : rc == -2 ? ": unsupported shape"
                                                                                         : ": workspace too small");
    throwRuntimeError(file, line, info);
}
#define B200_CHECK_RC(call) b200CheckRc((call), #call, __FILE__, __LINE__)
