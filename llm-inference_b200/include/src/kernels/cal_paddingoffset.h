// launchCalPaddingoffset — reference src/kernels/cal_paddingoffset.h, cal_paddingoffset.cu:74-86.
#pragma once
#include "src/kernels/b200_shim.h"

// padding_offset [B, max_q_len] (first sum(len) entries written), cum_seqlens [B+1],
// input_lens [B]; int32, bit-exact (golden vector: tests/unittests/test_cal_paddingoffset.cu:54-56).
inline void launchCalPaddingoffset(TensorWrapper<int>* padding_offset, TensorWrapper<int>* cum_seqlens,
                                   TensorWrapper<int>* input_lens) {
    const int batch_size = padding_offset->shape[0];
    const int max_q_len = padding_offset->shape[1];
    LLM_CHECK_WITH_INFO(batch_size == input_lens->shape[0],
                        "input lengths numbers should equal to padding offset bs dim!");
    LLM_CHECK_WITH_INFO(batch_size == cum_seqlens->shape[0] - 1,
                        "cum seqlen numbers should equal to padding offset bs dim + 1!");
    B200_CHECK_RC(b200_cal_padding_offset(padding_offset->data, cum_seqlens->data, input_lens->data, batch_size,
                                          max_q_len, b200shim::stream()));
}
