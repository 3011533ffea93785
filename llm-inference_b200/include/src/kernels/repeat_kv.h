// launchRepeatKVCache — reference src/kernels/repeat_kv.h, repeat_kv.cu:76-120.
#pragma once
#include "src/kernels/b200_shim.h"

// k_cache_src [layers, B, hk, max_seq, d] -> k_cache_dst [B, h, max_k_len, d], head j reads kv
// head j / (h/hk), rows < context_length[b] only. Bit-exact copy.
template <typename T>
void launchRepeatKVCache(TensorWrapper<T>* k_cache_src, TensorWrapper<T>* v_cache_src,
                         TensorWrapper<int>* context_length, TensorWrapper<int>* layer_id,
                         TensorWrapper<T>* k_cache_dst, TensorWrapper<T>* v_cache_dst) {
    const int batch_size = context_length->shape[0];
    const int kv_head_num = k_cache_src->shape[2];
    const int max_seq_len = k_cache_src->shape[3];
    const int head_num = k_cache_dst->shape[1];
    const int max_k_len = k_cache_dst->shape[2];
    const int head_size = k_cache_dst->shape[3];
    const int layer = layer_id->getVal();
    const size_t layer_offset = (size_t)layer * batch_size * kv_head_num * max_seq_len * head_size;
    B200_CHECK_RC(b200_repeat_kv_cache(k_cache_src->data, v_cache_src->data, k_cache_dst->data, v_cache_dst->data,
                                       layer_offset, context_length->data, batch_size, head_num, kv_head_num,
                                       max_seq_len, max_k_len, head_size, b200shim::dtype_of<T>(),
                                       b200shim::stream()));
}
