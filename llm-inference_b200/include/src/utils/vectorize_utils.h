// Vec<T>: the vector type a kernel moves per access (reference src/utils/vectorize_utils.h:48-66).
// The B200 kernels move 16-byte packets for both types (8 halfs / 4 floats); Vec<> is kept so
// that code written against the reference header still compiles.
#pragma once
#include <cuda_fp16.h>
#include <cuda_runtime.h>

template <typename T> struct Vec { using Type = T; static constexpr int size = 0; };
template <> struct Vec<half> { using Type = half2; static constexpr int size = 2; };
template <> struct Vec<float> { using Type = float4; static constexpr int size = 4; };
