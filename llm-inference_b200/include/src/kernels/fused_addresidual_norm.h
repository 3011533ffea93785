// launchFusedAddBiasResidualRMSNorm — reference src/kernels/fused_addresidual_norm.h,
// fused_addresidual_norm.cu:193-225: residual += decoder_out (+ norm.bias);
// decoder_out = RMSNorm(residual) * scale.
#pragma once
#include "src/kernels/b200_shim.h"
#include "src/utils/vectorize_utils.h"
#include "src/weights/base_weights.h"
#include "src/weights/llama/norm_weights.h"

template <typename T>
void launchFusedAddBiasResidualRMSNorm(TensorWrapper<T>* residual, TensorWrapper<T>* decoder_out,
                                       BaseWeight<T>& norm, T* scale, float eps) {
    const int num_tokens = decoder_out->shape[0];
    const int hidden_units = decoder_out->shape[1];
    B200_CHECK_RC(b200_fused_add_bias_residual_rmsnorm(residual->data, decoder_out->data, norm.bias, scale,
                                                       num_tokens, hidden_units, eps,
                                                       b200shim::dtype_of<T>(), b200shim::stream()));
}
