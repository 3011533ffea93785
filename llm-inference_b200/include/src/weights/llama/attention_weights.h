#pragma once
#include "src/weights/base_weights.h"
// reference src/weights/llama/attention_weights.h:5-11. qkv is the fused [ (h+2hk)*d, H ] matrix,
// output is o_proj [H, h*d]; both row-major [N, K] (used with trans_b = true).
template <typename T>
struct LLaMAattentionWeights {
    BaseWeight<T> q;
    BaseWeight<T> k;
    BaseWeight<T> v;
    BaseWeight<T> qkv;
    BaseWeight<T> output;
};
