// launchBuildCausalMasks — reference src/kernels/build_causal_mask.h, build_causal_mask.cu:50-66.
#pragma once
#include "src/kernels/b200_shim.h"

// mask [B, max_q_len, max_k_len] of 0/1 in T: q < qlen && k < klen && k <= q + (klen - qlen)
// (the unit test's CPU predicate, test_casual_mask.cu:12-32: history keys visible; D17).
// history_visible = false reproduces the GPU kernel's extra `k >= klen - qlen` clause (:29).
template <typename T>
void launchBuildCausalMasks(TensorWrapper<T>* mask, TensorWrapper<int>* q_lens, TensorWrapper<int>* k_lens,
                            bool history_visible = true) {
    const int batch_size = mask->shape[0];
    const int max_q_len = mask->shape[1];
    const int max_k_len = mask->shape[2];
    B200_CHECK_RC(b200_build_causal_mask(mask->data, q_lens->data, k_lens->data, batch_size, max_q_len, max_k_len,
                                         history_visible ? 1 : 0, b200shim::dtype_of<T>(), b200shim::stream()));
}
