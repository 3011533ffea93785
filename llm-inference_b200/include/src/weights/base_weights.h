// Weight carriers (reference src/weights/base_weights.h:6-38): WeightType, getWeightType<T>,
// BaseWeight<T>{shape, type, data, bias}.
#pragma once
#include <cuda_fp16.h>
#include <cstdint>
#include <type_traits>
#include <vector>

enum class WeightType { FP32_W, FP16_W, INT8_W, UNSUPPORTED_W };

template <typename T>
inline WeightType getWeightType() {
    using U = typename std::remove_const<T>::type;
    if (std::is_same<U, float>::value) return WeightType::FP32_W;
    if (std::is_same<U, half>::value) return WeightType::FP16_W;
    if (std::is_same<U, int8_t>::value) return WeightType::INT8_W;
    return WeightType::UNSUPPORTED_W;
}

template <typename T>
struct BaseWeight {
    std::vector<int> shape;
    WeightType type;
    T* data = nullptr;
    T* bias = nullptr;
};
