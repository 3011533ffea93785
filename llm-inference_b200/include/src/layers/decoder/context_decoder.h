// LlamaContextDecoder<T> — the prefill layer loop with the reference's interface
// (src/layers/decoder/context_decoder.{h,cpp}; forward :48-140). TensorMap keys:
//   inputs  decoder_input [num_tokens, H] (normalised in place), input_length [B],
//           history_length [B], context_length [B] (device int32), layer_id (CPU int)
//   outputs decoder_output [num_tokens, H], all_k_cache, all_v_cache
// Setup per call: padding offsets + cumulative lengths, causal mask (built only when the
// attention layer runs in reference order; the fused flash kernel needs no dense mask).
#pragma once
#include <vector>

#include "src/kernels/add_residual.h"
#include "src/kernels/build_causal_mask.h"
#include "src/kernels/cal_paddingoffset.h"
#include "src/kernels/fused_addresidual_norm.h"
#include "src/kernels/rmsnorm_kernel.h"
#include "src/layers/attention/context_attention.h"
#include "src/layers/ffn/ffn.h"
#include "src/utils/tensor.h"
#include "src/weights/llama/layer_weights.h"

template <typename T>
class LlamaContextDecoder {
private:
    int head_num;
    int kv_head_num;
    int head_size;
    int inter_size;
    int num_layer;
    int hidden_units;
    float rmsnorm_eps;
    DataType data_type;
    cudaStream_t stream;
    cublasWrapper* cublas_wrapper;
    BaseAllocator* allocator;
    TensorWrapper<T>* attention_mask = nullptr;  // [B, max_q, max_k]
    TensorWrapper<int>* padding_offset = nullptr;  // [B, max_q]
    TensorWrapper<int>* cum_seqlens = nullptr;     // [B + 1]
    TensorWrapper<T>* decoder_residual = nullptr;  // [num_tokens, H]
    LLaMAContextAttentionLayer<T>* ctxAttn;
    LLaMAFFNLayer<T>* ffn;
    bool dense_mask = true;   // build the [B,Sq,Sk] mask tensor (reference behaviour)

public:
    LlamaContextDecoder(int head_num, int kv_head_num, int head_size, int inter_size, int num_layer,
                        const LLaMAAttentionStaticParams& attn_params, float rmsnorm_eps, cudaStream_t stream,
                        cublasWrapper* cublas_wrapper, BaseAllocator* allocator)
        : head_num(head_num), kv_head_num(kv_head_num), head_size(head_size), inter_size(inter_size),
          num_layer(num_layer), hidden_units(head_num * head_size), rmsnorm_eps(rmsnorm_eps),
          data_type(getTensorType<T>()), stream(stream), cublas_wrapper(cublas_wrapper), allocator(allocator) {
        ctxAttn = new LLaMAContextAttentionLayer<T>(head_num, kv_head_num, head_size, attn_params, stream,
                                                    cublas_wrapper, allocator);
        ffn = new LLaMAFFNLayer<T>(head_num, head_size, inter_size, stream, cublas_wrapper, allocator);
    }
    ~LlamaContextDecoder() {
        freeBuf();
        delete ctxAttn;
        delete ffn;
    }
    LlamaContextDecoder(const LlamaContextDecoder&) = delete;
    LlamaContextDecoder& operator=(const LlamaContextDecoder&) = delete;

    // reference order (RepeatKV + batched GEMMs + dense mask) instead of the fused flash kernel
    void setFusedAttention(bool on) { ctxAttn->setFused(on); }
    void setDenseMask(bool on) { dense_mask = on; }

    void allocForForward(LLaMAAttentionDynParams& dyn_params) {
        freeBuf();
        const int num_tokens = dyn_params.num_tokens, batch_size = dyn_params.batch_size;
        const int max_q_len = dyn_params.max_q_len, max_k_len = dyn_params.max_k_len;
        const DataType type_int = getTensorType<int>();
        decoder_residual = new TensorWrapper<T>(Device::GPU, data_type, {num_tokens, hidden_units});
        attention_mask = new TensorWrapper<T>(Device::GPU, data_type, {batch_size, max_q_len, max_k_len});
        padding_offset = new TensorWrapper<int>(Device::GPU, type_int, {batch_size, max_q_len});
        cum_seqlens = new TensorWrapper<int>(Device::GPU, type_int, {batch_size + 1});
        decoder_residual->data = allocator->Malloc(decoder_residual->data, sizeof(T) * (size_t)num_tokens * hidden_units, false);
        attention_mask->data = allocator->Malloc(attention_mask->data, sizeof(T) * (size_t)batch_size * max_q_len * max_k_len, false);
        padding_offset->data = allocator->Malloc(padding_offset->data, sizeof(int) * (size_t)batch_size * max_q_len, false);
        cum_seqlens->data = allocator->Malloc(cum_seqlens->data, sizeof(int) * (size_t)(batch_size + 1), false);
    }
    void freeBuf() {
        if (decoder_residual) { allocator->Free(decoder_residual->data); delete decoder_residual; decoder_residual = nullptr; }
        if (attention_mask) { allocator->Free(attention_mask->data); delete attention_mask; attention_mask = nullptr; }
        if (padding_offset) { allocator->Free(padding_offset->data); delete padding_offset; padding_offset = nullptr; }
        if (cum_seqlens) { allocator->Free(cum_seqlens->data); delete cum_seqlens; cum_seqlens = nullptr; }
    }

    void forward(TensorMap& input_tensors, const std::vector<LlamaLayerWeight<T>*>& layerWeights,
                 TensorMap& output_tensors, LLaMAAttentionDynParams& dyn_params) {
        allocForForward(dyn_params);
        Tensor* seq_lens = input_tensors["input_length"];
        Tensor* context_length = input_tensors["context_length"];
        Tensor* history_length = input_tensors["history_length"];
        Tensor* decoder_output = output_tensors["decoder_output"];
        Tensor* all_k_cache = output_tensors["all_k_cache"];
        Tensor* all_v_cache = output_tensors["all_v_cache"];
        Tensor* layer_id0 = input_tensors["layer_id"];
        Tensor* decoder_input = input_tensors["decoder_input"];
        LLM_CHECK_WITH_INFO(decoder_input->as<T>()->data != nullptr, "the data ptr of tensor inserted into TensorMap is nullptr!");
        LLM_CHECK_WITH_INFO(history_length->as<int>()->data != nullptr, "the data ptr of tensor inserted into TensorMap is nullptr!");
        LLM_CHECK_WITH_INFO((int)layerWeights.size() >= num_layer, "one LlamaLayerWeight per layer is required");

        launchCalPaddingoffset(padding_offset, cum_seqlens, seq_lens->as<int>());
        if (dense_mask) launchBuildCausalMasks<T>(attention_mask, seq_lens->as<int>(), context_length->as<int>());

        int cur_layer = layer_id0->as<int>()->getVal();
        TensorWrapper<int> layer(Device::CPU, getTensorType<int>(), {1}, &cur_layer);
        TensorMap ctx_attn_inputs{{"attention_input", decoder_input}, {"padding_offset", padding_offset},
                                  {"history_length", history_length}, {"input_length", seq_lens},
                                  {"layer_id", &layer}, {"context_length", context_length},
                                  {"attention_mask", attention_mask}};
        TensorMap ctx_attn_output{{"attention_output", decoder_output}, {"all_k_cache", all_k_cache},
                                  {"all_v_cache", all_v_cache}};
        for (int layer_id = 0; layer_id < num_layer; layer_id++) {
            cur_layer = layer_id;
            decoder_input = ctx_attn_inputs["attention_input"];
            launchRMSNorm(decoder_input->as<T>(), decoder_residual, layerWeights[layer_id]->attn_norm_weight, rmsnorm_eps);
            ctxAttn->forward(ctx_attn_inputs, ctx_attn_output, layerWeights[layer_id]->self_attn_weight, dyn_params,
                             ctxAttn->GetAttnStaticParams());
            launchFusedAddBiasResidualRMSNorm(decoder_residual, decoder_output->as<T>(),
                                              layerWeights[layer_id]->self_attn_weight.output,
                                              layerWeights[layer_id]->ffn_norm_weight.gamma, rmsnorm_eps);
            TensorMap ffn_inputs{{"ffn_input", decoder_output}};
            TensorMap ffn_outputs{{"ffn_output", decoder_output}};
            dyn_params.is_ctx = true;
            ffn->forward(ffn_inputs, ffn_outputs, layerWeights[layer_id]->ffn_weight, dyn_params);
            launchAddResidual(decoder_residual, decoder_output->as<T>());
            ctx_attn_inputs.insert("attention_input", decoder_output);
        }
    }
};
