// LlamaSelfDecoder<T> — the per-token layer loop with the reference's interface
// (src/layers/decoder/self_decoder.{h,cpp}; forward :24-85). TensorMap keys:
//   inputs  decoder_input [B, H] (normalised in place, as the reference does), step (CPU int),
//           finished [B] bool, layer_id (CPU int)
//   outputs decoder_output [B, H], all_k_cache, all_v_cache [layers, B, hk, max_seq, d]
// Per layer: RMSNorm(+residual save) -> self attention -> fused add-residual+RMSNorm -> FFN ->
// add residual. D18 (FFN TensorMap keys) and the never-initialised per-layer layer_id tensor of
// the reference (:55-57) are fixed; the residual buffer persists between calls.
#pragma once
#include <vector>

#include "src/kernels/add_residual.h"
#include "src/kernels/fused_addresidual_norm.h"
#include "src/kernels/fused_decoder_self_attention.h"
#include "src/kernels/rmsnorm_kernel.h"
#include "src/layers/attention/masked_self_attention.h"
#include "src/layers/ffn/ffn.h"
#include "src/utils/tensor.h"
#include "src/weights/llama/layer_weights.h"

template <typename T>
class LlamaSelfDecoder {
private:
    int head_num;
    int kv_head_num;
    int head_size;
    int inter_size;
    int num_layer;
    int hidden_units;
    float rmsnorm_eps;
    cudaStream_t stream;
    cublasWrapper* cublas_wrapper;
    BaseAllocator* allocator;
    TensorWrapper<T>* decoder_residual = nullptr;
    LLaMASelfAttentionLayer<T>* selfAttn;
    LLaMAFFNLayer<T>* ffn;
    DataType data_type;
    int capacity_batch = 0;

public:
    LlamaSelfDecoder(int head_num, int kv_head_num, int head_size, int inter_size, int num_layer,
                     const LLaMAAttentionStaticParams& attn_params, float rmsnorm_eps, cudaStream_t stream,
                     cublasWrapper* cublas_wrapper, BaseAllocator* allocator)
        : head_num(head_num), kv_head_num(kv_head_num), head_size(head_size), inter_size(inter_size),
          num_layer(num_layer), hidden_units(head_num * head_size), rmsnorm_eps(rmsnorm_eps), stream(stream),
          cublas_wrapper(cublas_wrapper), allocator(allocator), data_type(getTensorType<T>()) {
        selfAttn = new LLaMASelfAttentionLayer<T>(head_num, kv_head_num, head_size, attn_params, stream,
                                                  cublas_wrapper, allocator);
        ffn = new LLaMAFFNLayer<T>(head_num, head_size, inter_size, stream, cublas_wrapper, allocator);
    }
    ~LlamaSelfDecoder() {
        freeBuf();
        delete selfAttn;
        delete ffn;
    }
    LlamaSelfDecoder(const LlamaSelfDecoder&) = delete;
    LlamaSelfDecoder& operator=(const LlamaSelfDecoder&) = delete;

    void allocForForward(LLaMAAttentionDynParams& dyn_params) {
        const int batch_size = dyn_params.batch_size;
        if (batch_size <= capacity_batch) {
            decoder_residual->shape = {batch_size, hidden_units};
            return;
        }
        freeBuf();
        decoder_residual = new TensorWrapper<T>(Device::GPU, data_type, {batch_size, hidden_units});
        decoder_residual->data =
            allocator->Malloc(decoder_residual->data, sizeof(T) * (size_t)batch_size * hidden_units, false);
        capacity_batch = batch_size;
    }
    void freeBuf() {
        if (decoder_residual) { allocator->Free(decoder_residual->data); delete decoder_residual; decoder_residual = nullptr; }
        capacity_batch = 0;
    }

    void forward(TensorMap& input_tensors, const std::vector<LlamaLayerWeight<T>*>& layerWeights,
                 TensorMap& output_tensors, LLaMAAttentionDynParams& dyn_params) {
        allocForForward(dyn_params);
        Tensor* decoder_input = input_tensors["decoder_input"];
        Tensor* decoder_output = output_tensors["decoder_output"];
        Tensor* step = input_tensors["step"];
        Tensor* finished = input_tensors["finished"];
        Tensor* all_k_cache = output_tensors["all_k_cache"];
        Tensor* all_v_cache = output_tensors["all_v_cache"];
        Tensor* layer_id0 = input_tensors["layer_id"];
        LLM_CHECK_WITH_INFO(decoder_input->as<T>()->data != nullptr, "the data ptr of tensor inserted into TensorMap is nullptr!");
        LLM_CHECK_WITH_INFO(decoder_output->as<T>()->data != nullptr, "the data ptr of tensor inserted into TensorMap is nullptr!");
        LLM_CHECK_WITH_INFO(step->as<int>()->data != nullptr, "the data ptr of tensor inserted into TensorMap is nullptr!");
        LLM_CHECK_WITH_INFO((int)layerWeights.size() >= num_layer, "one LlamaLayerWeight per layer is required");

        int cur_layer = layer_id0->as<int>()->getVal();
        TensorWrapper<int> layer(Device::CPU, getTensorType<int>(), {1}, &cur_layer);
        TensorMap self_attn_input{{"attention_input", decoder_input}, {"finished", finished}, {"step", step},
                                  {"layer_id", &layer}};
        TensorMap self_attn_output{{"attention_output", decoder_output}, {"all_k_cache", all_k_cache},
                                   {"all_v_cache", all_v_cache}};
        LLaMAAttentionDynParams ffn_params = dyn_params;
        ffn_params.num_tokens = 0;  // decode: the FFN sizes its scratch by batch (ffn.cpp:55-57)
        for (int layer_id = 0; layer_id < num_layer; layer_id++) {
            cur_layer = layer_id;
            decoder_input = self_attn_input["attention_input"];
            launchRMSNorm(decoder_input->as<T>(), decoder_residual, layerWeights[layer_id]->attn_norm_weight,
                          rmsnorm_eps, false);
            selfAttn->Forward(self_attn_input, self_attn_output, layerWeights[layer_id]->self_attn_weight, dyn_params);
            launchFusedAddBiasResidualRMSNorm(decoder_residual, decoder_output->as<T>(),
                                              layerWeights[layer_id]->self_attn_weight.output,
                                              layerWeights[layer_id]->ffn_norm_weight.gamma, rmsnorm_eps);
            TensorMap ffn_inputs{{"ffn_input", decoder_output}};
            TensorMap ffn_outputs{{"ffn_output", decoder_output}};
            ffn->forward(ffn_inputs, ffn_outputs, layerWeights[layer_id]->ffn_weight, ffn_params);
            launchAddResidual(decoder_residual, decoder_output->as<T>(), false);
            self_attn_input.insert("attention_input", decoder_output);
        }
    }
};
