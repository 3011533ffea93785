// LLaMASelfAttentionLayer<T> — decode-time attention layer with the reference's interface
// (src/layers/attention/masked_self_attention.{h,cpp}): Forward(inputs, outputs, weights, params)
// with TensorMap keys attention_input / step / finished / layer_id -> attention_output /
// all_k_cache / all_v_cache (masked_self_attention.cpp:55-92). Call order kept:
//   qkv GEMM -> RoPE (q,k at position step-1) -> fused masked MHA (KV append + split-KV
//   attention) -> output GEMM.
// Scratch persists between calls; nothing synchronises the device (reference: Malloc/Free of two
// buffers and four device syncs per layer per token, :27-52,63-85).
#pragma once
#include "src/kernels/cublas_utils.h"
#include "src/kernels/fused_decoder_self_attention.h"
#include "src/kernels/linear.h"
#include "src/kernels/qkv_bias_and_RoPE.h"
#include "src/memory/allocator/cuda_allocator.h"
#include "src/models/llama/llama_params.h"
#include "src/utils/macro.h"
#include "src/utils/tensor.h"
#include "src/weights/llama/attention_weights.h"

template <typename T>
class LLaMASelfAttentionLayer {
private:
    const int head_num;
    const int kv_head_num;
    const int head_size;
    const int hidden_size;
    const int q_head_per_kv;
    float scale;
    LLaMAAttentionStaticParams attn_static_params;
    cudaStream_t stream;
    BaseAllocator* allocator;
    cublasWrapper* cublas_wrapper;
    TensorWrapper<T>* qkv_buf = nullptr;     // [B, h + 2hk, d]
    TensorWrapper<T>* mha_output = nullptr;  // [B, h*d]
    int capacity_batch = 0;

public:
    LLaMASelfAttentionLayer(int head_num, int kv_head_num, int head_size, LLaMAAttentionStaticParams attn_params,
                            cudaStream_t stream, cublasWrapper* cublas_wrapper, BaseAllocator* allocator)
        : head_num(head_num), kv_head_num(kv_head_num), head_size(head_size), hidden_size(head_num * head_size),
          q_head_per_kv(head_num / kv_head_num), scale(1.0f / sqrtf((float)head_size)),
          attn_static_params(attn_params), stream(stream), allocator(allocator), cublas_wrapper(cublas_wrapper) {}
    ~LLaMASelfAttentionLayer() { freeBuf(); }
    LLaMASelfAttentionLayer(const LLaMASelfAttentionLayer&) = delete;
    LLaMASelfAttentionLayer& operator=(const LLaMASelfAttentionLayer&) = delete;

    LLaMAAttentionStaticParams& GetAttnStaticParams() { return attn_static_params; }

    void allocForForward(LLaMAAttentionDynParams& params) {
        const int batch_size = params.batch_size;
        const int qkv_head_num = head_num + 2 * kv_head_num;
        if (batch_size <= capacity_batch) {
            qkv_buf->shape = {batch_size, qkv_head_num, head_size};
            mha_output->shape = {batch_size, hidden_size};
            return;
        }
        freeBuf();
        const DataType type = getTensorType<T>();
        qkv_buf = new TensorWrapper<T>(Device::GPU, type, {batch_size, qkv_head_num, head_size});
        mha_output = new TensorWrapper<T>(Device::GPU, type, {batch_size, hidden_size});
        qkv_buf->data = allocator->Malloc(qkv_buf->data, sizeof(T) * (size_t)batch_size * qkv_head_num * head_size, false);
        mha_output->data = allocator->Malloc(mha_output->data, sizeof(T) * (size_t)batch_size * hidden_size, false);
        capacity_batch = batch_size;
    }
    void freeBuf() {
        if (qkv_buf) { allocator->Free(qkv_buf->data); delete qkv_buf; qkv_buf = nullptr; }
        if (mha_output) { allocator->Free(mha_output->data); delete mha_output; mha_output = nullptr; }
        capacity_batch = 0;
    }

    void Forward(TensorMap& inputs, TensorMap& outputs, LLaMAattentionWeights<T>& weights,
                 LLaMAAttentionDynParams& params) {
        allocForForward(params);
        Tensor* attention_input = inputs["attention_input"];
        Tensor* step = inputs["step"];
        Tensor* finished = inputs["finished"];
        Tensor* layer_id = inputs["layer_id"];
        Tensor* attention_output = outputs["attention_output"];
        Tensor* key_cache = outputs["all_k_cache"];
        Tensor* value_cache = outputs["all_v_cache"];
        launchLinearGemm(attention_input->as<T>(), weights.qkv, qkv_buf, cublas_wrapper, false, true);
        launchRoPE(qkv_buf, step->as<int>(), attn_static_params, kv_head_num);
        launchDecoderMaskedMHA<T>(qkv_buf, weights.qkv, layer_id->as<int>(), key_cache->as<T>(), value_cache->as<T>(),
                                  finished->as<bool>(), step->as<int>(), mha_output, attn_static_params);
        launchLinearGemm(mha_output, weights.output, attention_output->as<T>(), cublas_wrapper, false, true);
    }
};
