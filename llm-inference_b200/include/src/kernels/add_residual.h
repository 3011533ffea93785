// launchAddResidual — reference src/kernels/add_residual.h, add_residual.cu:58-77:
// decoder_out += residual, both [num_tokens, hidden].
#pragma once
#include "src/kernels/b200_shim.h"
#include "src/utils/vectorize_utils.h"

template <typename T>
void launchAddResidual(TensorWrapper<T>* residual, TensorWrapper<T>* decoder_out, bool is_print = false) {
    (void)is_print;
    const int num_tokens = decoder_out->shape[0];
    const int hidden_units = decoder_out->shape[1];
    LLM_CHECK_WITH_INFO(residual->shape[0] == num_tokens && residual->shape[1] == hidden_units,
                        "residual and decoder_out shapes must match");
    B200_CHECK_RC(b200_add_residual(residual->data, decoder_out->data, num_tokens, hidden_units,
                                    b200shim::dtype_of<T>(), b200shim::stream()));
}
