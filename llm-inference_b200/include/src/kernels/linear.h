// launchLinearGemm / launchLinearStridedBatchGemm — reference src/kernels/linear.h,
// linear.cu:39-104 and :127-229 (cublasWrapper::Gemm / stridedBatchedGemm, cublas_utils.cc:32,75).
#pragma once
#include "src/kernels/b200_shim.h"
#include "src/kernels/cublas_utils.h"
#include "src/weights/base_weights.h"

// y[num_tokens, N] = x[num_tokens, K] * op(W).  trans_b: weight.shape = [N, K] (every engine call
// site); otherwise [K, N] (tests/unittests/test_linear.cu). 3-D input / output are flattened over
// their last two dims exactly as linear.cu:57-58 does. trans_a is not used anywhere on the decoder
// path and is rejected. The cublasWrapper is accepted and ignored.
template <typename T>
void launchLinearGemm(TensorWrapper<T>* input, BaseWeight<T>& weight, TensorWrapper<T>* output,
                      cublasWrapper* cublas_wrapper, bool trans_a = false, bool trans_b = false) {
    (void)cublas_wrapper;
    LLM_CHECK_WITH_INFO(!trans_a, "launchLinearGemm: trans_a is not supported on the decoder path");
    const int Bn = input->shape[0];
    const int Bk = input->shape.size() == 3 ? input->shape[1] * input->shape[2] : input->shape[1];
    const int Cm = output->shape.size() == 3 ? output->shape[1] * output->shape[2] : output->shape[1];
    const int Wk = trans_b ? weight.shape[1] : weight.shape[0];
    const int Wn = trans_b ? weight.shape[0] : weight.shape[1];
    LLM_CHECK_WITH_INFO(Wk == Bk, "2nd dim of input MUST = 1st dim of weight");
    LLM_CHECK_WITH_INFO(Wn == Cm && output->shape[0] == Bn, "output shape MUST be [num_tokens, weight out dim]");
    const int dt = b200shim::dtype_of<T>();
    const size_t ws_bytes = b200_linear_workspace_bytes(Bn, Wn, Wk, trans_b ? 1 : 0, dt);
    void* ws = ws_bytes ? b200shim::workspace(ws_bytes) : nullptr;
    B200_CHECK_RC(b200_linear(input->data, weight.data, nullptr, output->data, Bn, Wn, Wk, trans_b ? 1 : 0, dt,
                              ws, ws_bytes, b200shim::stream()));
}

// input1 [bs, heads, m, k]; input2 [bs, heads, n, k] with trans_b (Q.K^T) or [bs, heads, k, n]
// (P.V); output [bs, heads, m, n]. Strides follow the shapes (the reference's strideC = Cm*Cm,
// linear.cu:187, is D14 and is not reproduced).
template <typename T>
void launchLinearStridedBatchGemm(TensorWrapper<T>* input1, TensorWrapper<T>* input2, TensorWrapper<T>* output,
                                  cublasWrapper* cublas_wrapper, bool trans_a = false, bool trans_b = false) {
    (void)cublas_wrapper;
    LLM_CHECK_WITH_INFO(!trans_a, "launchLinearStridedBatchGemm: trans_a is not supported");
    const int batch_count = input1->shape[0] * input1->shape[1];
    const int m = input1->shape[2], k = input1->shape[3];
    const int n = trans_b ? input2->shape[2] : input2->shape[3];
    const int k2 = trans_b ? input2->shape[3] : input2->shape[2];
    LLM_CHECK_WITH_INFO(k == k2, "2nd dim of input MUST = 1st dim of weight");
    LLM_CHECK_WITH_INFO(output->shape[2] == m && output->shape[3] == n, "output MUST be [bs, heads, m, n]");
    B200_CHECK_RC(b200_batched_gemm(input1->data, input2->data, output->data, batch_count, m, n, k,
                                    trans_b ? 1 : 0, b200shim::dtype_of<T>(), b200shim::stream()));
}
