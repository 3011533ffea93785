// launchScaleMaskAndSoftmax — reference src/kernels/attn_softmax_kernel.h,
// attn_softmax_kernel.cu:337-366: softmax_row(scale*qk + (1-mask)*(-10000)), denominator + 1e-6.
#pragma once
#include "src/kernels/b200_shim.h"
#include "src/utils/vectorize_utils.h"

// qk / attn_score [B, h, q_len, k_len] (may alias: context_attention.cpp:150); mask [B, q_len, k_len].
// No k_len < 4096 limit (attn_softmax_kernel.cu:293-334 has one).
template <typename T>
void launchScaleMaskAndSoftmax(TensorWrapper<T>* qk, TensorWrapper<T>* mask, TensorWrapper<T>* attn_score,
                               float scale) {
    const int batch_size = qk->shape[0];
    const int head_nums = qk->shape[1];
    const int q_length = qk->shape[2];
    const int k_length = qk->shape[3];
    B200_CHECK_RC(b200_scale_mask_softmax(qk->data, mask->data, attn_score->data, batch_size, head_nums, q_length,
                                          k_length, scale, b200shim::dtype_of<T>(), b200shim::stream()));
}
