// launchConcatKVCache — reference src/kernels/concat_past_kv.h, concat_past_kv.cu:103-145.
#pragma once
#include "src/kernels/b200_shim.h"

// k_src/v_src [B, hk, max_q_len, d] -> k_dst/v_dst [layers, B, hk, max_seq, d] at rows
// history_length[b] + t, t < cur_query_length[b]. Bit-exact copy.
template <typename T>
void launchConcatKVCache(TensorWrapper<T>* k_src, TensorWrapper<T>* v_src, TensorWrapper<int>* layer_id,
                         TensorWrapper<int>* cur_query_length, TensorWrapper<int>* history_length,
                         TensorWrapper<T>* k_dst, TensorWrapper<T>* v_dst) {
    const int batch_size = k_src->shape[0];
    const int kv_head_num = k_src->shape[1];
    const int max_q_len = k_src->shape[2];
    const int head_size = k_src->shape[3];
    const int max_seq_len = k_dst->shape[3];
    const int layer = layer_id->getVal();
    const size_t layer_offset = (size_t)layer * batch_size * kv_head_num * max_seq_len * head_size;
    B200_CHECK_RC(b200_concat_kv_cache(k_src->data, v_src->data, k_dst->data, v_dst->data, layer_offset,
                                       cur_query_length->data, history_length->data, batch_size, kv_head_num,
                                       max_q_len, max_seq_len, head_size, b200shim::dtype_of<T>(),
                                       b200shim::stream()));
}
