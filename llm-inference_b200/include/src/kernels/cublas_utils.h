// cublasWrapper — reference src/kernels/cublas_utils.h. Kept as a type so that layer code written
// against the reference (`cublasWrapper* cublas_wrapper` in every layer constructor) compiles
// unchanged; the B200 linears are hand-written kernels, so the wrapper carries only the chosen
// precision configuration and never calls cuBLAS (the library is not linked).
#pragma once
#include <cuda_runtime.h>

#include "src/utils/macro.h"

// opaque stand-ins so the reference's constructor call `new cublasWrapper(handle, lt_handle)`
// still type-checks without <cublas_v2.h>
#ifndef CUBLAS_V2_H_
typedef struct cublasContext* cublasHandle_t;
#endif
#ifndef CUBLASLT_H_
typedef struct cublasLtContext* cublasLtHandle_t;
#endif

class cublasWrapper {
private:
    cublasHandle_t cublas_handle_ = nullptr;
    cublasLtHandle_t cublaslt_handle_ = nullptr;
    bool fp16_ = false;

public:
    cublasWrapper() = default;
    cublasWrapper(cublasHandle_t cublas_handle, cublasLtHandle_t cublaslt_handle)
        : cublas_handle_(cublas_handle), cublaslt_handle_(cublaslt_handle) {}
    ~cublasWrapper() = default;
    void setFP32GemmConfig() { fp16_ = false; }   // cublas_utils.cc:14-20
    void setFP16GemmConfig() { fp16_ = true; }    // cublas_utils.cc:22-28 (ours accumulates in fp32)
    bool isFP16() const { return fp16_; }
};
