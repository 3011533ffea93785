// launchDecoderMaskedMHA — reference src/kernels/fused_decoder_self_attention.h:10-19,
// fused_decoder_self_attention.cu:317-365. Split-KV (flash-decoding) kernel behind it.
#pragma once
#include "src/kernels/b200_shim.h"
#include "src/models/llama/llama_params.h"
#include "src/utils/vectorize_utils.h"
#include "src/weights/base_weights.h"

// qkv_buf [B, h+2hk, d] token-major (D5); k_cache/v_cache [layers, B, hk, max_seq, d] (whole
// cache; the layer slice is taken with the kv_head_num factor the reference omits, D3);
// step / layer_id host-resident; appends k,v at row step-1 then attends over [0, step);
// mha_output [B, h*d]. qkv.bias, finished and static_params are accepted and unused, as in the
// reference kernel (:341-344).
template <typename T>
void launchDecoderMaskedMHA(TensorWrapper<T>* qkv_buf, BaseWeight<T>& qkv, TensorWrapper<int>* layer_id,
                            TensorWrapper<T>* k_cache, TensorWrapper<T>* v_cache, TensorWrapper<bool>* finished,
                            TensorWrapper<int>* step, TensorWrapper<T>* mha_output,
                            LLaMAAttentionStaticParams& static_params) {
    (void)qkv; (void)finished; (void)static_params;
    const int batch_size = qkv_buf->shape[0];
    const int qkv_head_num = qkv_buf->shape[1];
    const int kv_head_num = k_cache->shape[2];
    const int head_num = qkv_head_num - 2 * kv_head_num;
    const int max_seq_len = k_cache->shape[3];
    const int head_size = qkv_buf->shape[2];
    const int cur_step = step->getVal();
    const int layer = layer_id->getVal();
    const size_t layer_offset = (size_t)layer * batch_size * kv_head_num * max_seq_len * head_size;
    const size_t ws_bytes = b200_decode_attention_workspace_bytes(batch_size, head_num, head_size, max_seq_len);
    void* ws = b200shim::workspace(ws_bytes);
    B200_CHECK_RC(b200_decode_attention(qkv_buf->data, k_cache->data + layer_offset, v_cache->data + layer_offset,
                                        mha_output->data, ws, ws_bytes, batch_size, head_num, kv_head_num,
                                        head_size, max_seq_len, cur_step, /*append_kv=*/1,
                                        b200shim::dtype_of<T>(), b200shim::stream()));
}
