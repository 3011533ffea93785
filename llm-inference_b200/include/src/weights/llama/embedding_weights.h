#pragma once
#include "src/weights/base_weights.h"
// reference src/weights/llama/embedding_weights.h
template <typename T>
struct EmbeddingWeight : public BaseWeight<T> {};
