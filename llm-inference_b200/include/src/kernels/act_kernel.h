// launchAct (SwiGLU) — reference src/kernels/act_kernel.h, act_kernel.cu:57-75:
// input [num_tokens, 2, inter] (gate block then up block) -> out [num_tokens, inter].
#pragma once
#include "src/kernels/b200_shim.h"
#include "src/utils/vectorize_utils.h"

template <typename T>
void launchAct(TensorWrapper<T>* input, TensorWrapper<T>* out) {
    const int batch_size = input->shape[0];
    LLM_CHECK(input->shape[1] == 2);   // act_kernel.cu:62
    const int intermedia_size = input->shape[2];
    B200_CHECK_RC(b200_swiglu(input->data, out->data, batch_size, intermedia_size, b200shim::dtype_of<T>(),
                              b200shim::stream()));
}
