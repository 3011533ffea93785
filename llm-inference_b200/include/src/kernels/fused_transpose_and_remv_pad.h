// launchTransposeOutRemovePadding — reference src/kernels/fused_transpose_and_remv_pad.h,
// fused_transpose_and_remv_pad.cu:51-78: [B, h, S, d] -> [num_tokens, h*d], pads dropped.
#pragma once
#include "src/kernels/b200_shim.h"

template <typename T>
void launchTransposeOutRemovePadding(TensorWrapper<T>* qkv_buf_w_pad, TensorWrapper<int>* padding_offset,
                                     TensorWrapper<T>* qkv_buf_wo_pad_1) {
    const int batch_size = qkv_buf_w_pad->shape[0];
    const int head_num = qkv_buf_w_pad->shape[1];
    const int seq_len = qkv_buf_w_pad->shape[2];
    const int head_size = qkv_buf_w_pad->shape[3];
    const int num_tokens = qkv_buf_wo_pad_1->shape[0];
    B200_CHECK_RC(b200_transpose_remove_padding(qkv_buf_w_pad->data, qkv_buf_wo_pad_1->data, padding_offset->data,
                                                num_tokens, batch_size, seq_len, head_num, head_size,
                                                b200shim::dtype_of<T>(), b200shim::stream()));
}
