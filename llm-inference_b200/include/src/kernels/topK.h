// launchTopKforBeamSearch — reference src/kernels/topK.h, topK.cu:147-190 (K = 5, two rounds of
// cub::BlockReduce). One pass here; indices are bit-exact with ties resolved to the lower index.
#pragma once
#include "src/kernels/b200_shim.h"

// probs [bs*beam, vocab]; final_topk_ids / final_topk_values [bs*beam, K] with K = last dim
// (<= 8). topk_ids / topK_values (the reference's round-1 scratch) are accepted and unused.
template <typename T>
void launchTopKforBeamSearch(TensorWrapper<T>* probs, TensorWrapper<int>* topk_ids, TensorWrapper<T>* topK_values,
                             TensorWrapper<int>* final_topk_ids, TensorWrapper<T>* final_topk_values) {
    (void)topk_ids; (void)topK_values;
    const int bsxbm = probs->shape[0];
    const int vocab_size = probs->shape[1];
    const int K = final_topk_ids->shape.back();
    B200_CHECK_RC(b200_topk(probs->data, final_topk_ids->data, final_topk_values->data, bsxbm, vocab_size, K,
                            b200shim::dtype_of<T>(), b200shim::stream()));
}
