#pragma once
// reference src/weights/llama/norm_weights.h:7
template <typename T>
struct LayerNormWeight {
    T* gamma = nullptr;
};
