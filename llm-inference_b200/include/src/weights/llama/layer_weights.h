// LlamaLayerWeight<T>: one decoder layer's weights (reference src/weights/llama/layer_weights.h:9-43,
// layer_weights.cc:18-177). Differences, all listed in SURVEY §2.3: down is allocated H*I elements
// (D19), sizes are exact (D20); everything lives in ONE 256-byte-aligned device arena per layer
// so the GEMV's bulk copies and the GEMM's TMA maps see naturally aligned row-major [N,K] tiles.
// The on-disk format is the reference's: raw little-endian fp32 .bin per tensor named
// model.layers.<i>.{input_layernorm,post_attention_layernorm}.weight.bin,
// model.layers.<i>.self_attn.{qkv,o_proj}.weight.bin, model.layers.<i>.mlp.{gate_up_proj,down_proj}.weight.bin
#pragma once
#include <string>

#include "src/weights/llama/attention_weights.h"
#include "src/weights/llama/ffn_weights.h"
#include "src/weights/llama/norm_weights.h"

template <typename T>
class LlamaLayerWeight {
private:
    int head_num;
    int kv_head_num;
    int head_size;
    int hidden_units;
    int inter_size;
    WeightType weight_type;
    bool attn_bias;
    int bit_size;
    char* arena_ = nullptr;  // single device allocation backing every tensor of the layer

public:
    LlamaLayerWeight() = delete;
    LlamaLayerWeight(int head_num, int kv_head_num, int head_size, int inter_size, WeightType weight_type,
                     bool attn_bias);
    ~LlamaLayerWeight();
    LlamaLayerWeight(const LlamaLayerWeight&) = delete;
    LlamaLayerWeight& operator=(const LlamaLayerWeight&) = delete;

    void loadWeights(std::string weight_path, WeightType weight_type);  // <path> + tensor name + ".bin"
    void loadWeights();                                                 // deterministic dummy fill
    void loadWeights(unsigned long long seed, float std);               // N(0, std^2) on the host, H2D

    LayerNormWeight<T> attn_norm_weight;
    LayerNormWeight<T> ffn_norm_weight;
    LLaMAattentionWeights<T> self_attn_weight;
    LLaMAFFNWeights<T> ffn_weight;
};

#include "src/weights/llama/layer_weights_impl.h"
