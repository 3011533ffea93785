#pragma once
#include "src/weights/base_weights.h"
// reference src/weights/llama/ffn_weights.h:5-10. gateAndup is [2*I, H] with the gate rows first.
template <typename T>
struct LLaMAFFNWeights {
    BaseWeight<T> gate;
    BaseWeight<T> up;
    BaseWeight<T> down;
    BaseWeight<T> gateAndup;
};
