// launchAddFusedQKVBiasTransposeAndRoPE / launchRoPE — reference src/kernels/qkv_bias_and_RoPE.h,
// qkv_bias_and_RoPE.cu:248-290 and :410-445.
#pragma once
#include "src/kernels/b200_shim.h"
#include "src/models/llama/llama_params.h"
#include "src/utils/vectorize_utils.h"
#include "src/weights/base_weights.h"

// prefill: QKV [token_num, h+2hk, d] (no padding) -> q [B,h,S,d], k/v [B,hk,S,d] (re-padded),
// + qkv.bias if present, RoPE on q,k at history_length[b] + local token index (D12 resolved).
template <typename T>
void launchAddFusedQKVBiasTransposeAndRoPE(TensorWrapper<T>* q_buf, TensorWrapper<T>* k_buf, TensorWrapper<T>* v_buf,
                                           TensorWrapper<T>* QKV, BaseWeight<T>& qkv,
                                           TensorWrapper<int>* padding_offset, TensorWrapper<int>* history_length,
                                           TensorWrapper<int>* input_length, LLaMAAttentionStaticParams& params) {
    const int token_num = QKV->shape[0];
    const int qkv_head_num = QKV->shape[1];
    const int head_size = QKV->shape[2];
    const int batch_size = q_buf->shape[0];
    const int head_num = q_buf->shape[1];
    const int seq_len = q_buf->shape[2];
    const int kv_head_num = k_buf->shape[1];
    LLM_CHECK_WITH_INFO(qkv_head_num == head_num + 2 * kv_head_num, "QKV head count MUST be head_num + 2*kv_head_num");
    B200_CHECK_RC(b200_qkv_bias_rope_transpose(q_buf->data, k_buf->data, v_buf->data, QKV->data, qkv.bias,
                                               padding_offset->data, history_length->data, input_length->data,
                                               batch_size, seq_len, token_num, head_num, kv_head_num, head_size,
                                               params.rotary_embedding_dim, params.rotary_embedding_base,
                                               b200shim::dtype_of<T>(), b200shim::stream()));
}

// decode: in-place RoPE on q,k of the newest token (position step-1), qkv_buf [B, h+2hk, d]
// token-major (D5). The reference infers head_num = shape[1]/3 (MHA only, :416); pass
// kv_head_num explicitly for GQA.
template <typename T>
void launchRoPE(TensorWrapper<T>* qkv_buf, TensorWrapper<int>* step, LLaMAAttentionStaticParams& static_params,
                int kv_head_num = -1) {
    const int batch_size = qkv_buf->shape[0];
    const int qkv_head_num = qkv_buf->shape[1];
    const int head_size = qkv_buf->shape[2];
    if (kv_head_num < 0) kv_head_num = qkv_head_num / 3;
    const int head_num = qkv_head_num - 2 * kv_head_num;
    const int cur_step = step->getVal();
    B200_CHECK_RC(b200_rope_decode(qkv_buf->data, batch_size, head_num, kv_head_num, head_size, cur_step,
                                   static_params.rotary_embedding_dim, static_params.rotary_embedding_base,
                                   b200shim::dtype_of<T>(), b200shim::stream()));
}
