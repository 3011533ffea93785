// Implementation of LlamaLayerWeight<T> (included from layer_weights.h; header-only, plain C++).
// Stands in for reference src/weights/llama/layer_weights.cc:9-177 + src/utils/weight_utils.cu:91-181
// (loadWeightFromBin: size-checked raw little-endian fp32 read, cast to T).
#pragma once
#include <cuda_runtime.h>

#include <cstdint>
#include <cstdio>
#include <fstream>
#include <random>
#include <vector>

#include "src/utils/macro.h"

namespace b200weights {

inline size_t align256(size_t n) { return (n + 255) & ~(size_t)255; }

template <typename T> inline T from_float(float v);
template <> inline float from_float<float>(float v) { return v; }
template <> inline half from_float<half>(float v) { return __float2half(v); }

// host fp32 -> device T
template <typename T>
inline void upload(T* dst, const std::vector<float>& src) {
    std::vector<T> tmp(src.size());
    for (size_t i = 0; i < src.size(); ++i) tmp[i] = from_float<T>(src[i]);
    CHECK(cudaMemcpy(dst, tmp.data(), sizeof(T) * tmp.size(), cudaMemcpyHostToDevice));
}

// reference weight_utils.cu:91-142: the file must hold exactly `elems` fp32 values
template <typename T>
inline void load_bin(T* dst, size_t elems, const std::string& file) {
    std::ifstream in(file, std::ios::in | std::ios::binary);
    LLM_CHECK_WITH_INFO(in.is_open(), "file " + file + " cannot be opened, loading model fails!");
    in.seekg(0, in.end);
    const size_t bytes = (size_t)in.tellg();
    in.seekg(0, in.beg);
    LLM_CHECK_WITH_INFO(bytes == elems * sizeof(float),
                        "file " + file + " has " + std::to_string(bytes) + " bytes, expected " +
                            std::to_string(elems * sizeof(float)));
    std::vector<float> host(elems);
    in.read(reinterpret_cast<char*>(host.data()), (std::streamsize)bytes);
    upload(dst, host);
}

}  // namespace b200weights

template <typename T>
LlamaLayerWeight<T>::LlamaLayerWeight(int head_num, int kv_head_num, int head_size, int inter_size,
                                      WeightType weight_type, bool attn_bias)
    : head_num(head_num), kv_head_num(kv_head_num), head_size(head_size), hidden_units(head_num * head_size),
      inter_size(inter_size), weight_type(weight_type), attn_bias(attn_bias), bit_size(8 * (int)sizeof(T)) {
    using b200weights::align256;
    const size_t H = hidden_units, QN = (size_t)(head_num + 2 * kv_head_num) * head_size, I = inter_size;
    const size_t sz[8] = {H, H, QN * H, H * H, 2 * I * H, H * I, attn_bias ? QN : 0, attn_bias ? H : 0};
    size_t off[8], total = 0;
    for (int i = 0; i < 8; ++i) { off[i] = total; total += align256(sz[i] * sizeof(T)); }
    CHECK(cudaMalloc((void**)&arena_, total));
    CHECK(cudaMemset(arena_, 0, total));
    auto at = [&](int i) { return sz[i] ? reinterpret_cast<T*>(arena_ + off[i]) : nullptr; };
    attn_norm_weight.gamma = at(0);
    ffn_norm_weight.gamma = at(1);
    self_attn_weight.qkv.type = weight_type;
    self_attn_weight.qkv.shape = {(int)QN, (int)H};
    self_attn_weight.qkv.data = at(2);
    self_attn_weight.output.type = weight_type;
    self_attn_weight.output.shape = {(int)H, (int)H};
    self_attn_weight.output.data = at(3);
    ffn_weight.gateAndup.type = weight_type;
    ffn_weight.gateAndup.shape = {(int)(2 * I), (int)H};
    ffn_weight.gateAndup.data = at(4);
    ffn_weight.down.type = weight_type;
    ffn_weight.down.shape = {(int)H, (int)I};   // D19: the reference allocates H*H here
    ffn_weight.down.data = at(5);
    self_attn_weight.qkv.bias = at(6);
    self_attn_weight.output.bias = at(7);
}

template <typename T>
LlamaLayerWeight<T>::~LlamaLayerWeight() {
    if (arena_) cudaFree(arena_);
}

template <typename T>
void LlamaLayerWeight<T>::loadWeights(std::string weight_path, WeightType weight_type_) {
    (void)weight_type_;
    using b200weights::load_bin;
    const size_t H = hidden_units, QN = (size_t)(head_num + 2 * kv_head_num) * head_size, I = inter_size;
    load_bin(attn_norm_weight.gamma, H, weight_path + ".input_layernorm.weight.bin");
    load_bin(ffn_norm_weight.gamma, H, weight_path + ".post_attention_layernorm.weight.bin");
    load_bin(self_attn_weight.qkv.data, QN * H, weight_path + ".self_attn.qkv.weight.bin");
    load_bin(self_attn_weight.output.data, H * H, weight_path + ".self_attn.o_proj.weight.bin");
    load_bin(ffn_weight.gateAndup.data, 2 * I * H, weight_path + ".mlp.gate_up_proj.weight.bin");
    load_bin(ffn_weight.down.data, H * I, weight_path + ".mlp.down_proj.weight.bin");
    if (attn_bias) {
        load_bin(self_attn_weight.qkv.bias, QN, weight_path + ".self_attn.qkv.bias.bin");
        load_bin(self_attn_weight.output.bias, H, weight_path + ".self_attn.o_proj.bias.bin");
    }
}

// dummy weights in the reference's range (layer_weights.cc:70-120: rand()%100 / 1e5), but from a
// fixed-seed generator so that runs are reproducible
template <typename T>
void LlamaLayerWeight<T>::loadWeights() {
    const size_t H = hidden_units, QN = (size_t)(head_num + 2 * kv_head_num) * head_size, I = inter_size;
    uint64_t state = 0x9E3779B97F4A7C15ull;
    auto fill = [&](T* dst, size_t n) {
        if (!dst) return;
        std::vector<float> h(n);
        for (size_t i = 0; i < n; ++i) {
            state = state * 6364136223846793005ull + 1442695040888963407ull;
            h[i] = (float)((state >> 33) % 100) / 100000.0f;
        }
        b200weights::upload(dst, h);
    };
    fill(attn_norm_weight.gamma, H);
    fill(ffn_norm_weight.gamma, H);
    fill(self_attn_weight.qkv.data, QN * H);
    fill(self_attn_weight.output.data, H * H);
    fill(ffn_weight.gateAndup.data, 2 * I * H);
    fill(ffn_weight.down.data, H * I);
    fill(self_attn_weight.qkv.bias, QN);
    fill(self_attn_weight.output.bias, H);
}

template <typename T>
void LlamaLayerWeight<T>::loadWeights(unsigned long long seed, float std) {
    const size_t H = hidden_units, QN = (size_t)(head_num + 2 * kv_head_num) * head_size, I = inter_size;
    std::mt19937_64 gen(seed);
    auto fill = [&](T* dst, size_t n, float mean) {
        if (!dst) return;
        std::normal_distribution<float> nd(mean, std);
        std::vector<float> h(n);
        for (auto& v : h) v = nd(gen);
        b200weights::upload(dst, h);
    };
    fill(attn_norm_weight.gamma, H, 1.f);
    fill(ffn_norm_weight.gamma, H, 1.f);
    fill(self_attn_weight.qkv.data, QN * H, 0.f);
    fill(self_attn_weight.output.data, H * H, 0.f);
    fill(ffn_weight.gateAndup.data, 2 * I * H, 0.f);
    fill(ffn_weight.down.data, H * I, 0.f);
    fill(self_attn_weight.qkv.bias, QN, 0.f);
    fill(self_attn_weight.output.bias, H, 0.f);
}
