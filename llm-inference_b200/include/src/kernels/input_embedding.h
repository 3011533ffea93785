// launchInputEmbedding — reference src/kernels/input_embedding.h, input_embedding.cu:30-50.
#pragma once
#include "src/kernels/b200_shim.h"
#include "src/weights/llama/embedding_weights.h"

// input_ids [token_num] -> output [token_num, hidden]; embed_table->data [vocab, hidden]
// (embed_table->shape = {vocab, hidden}; ids outside [0, vocab) are clamped instead of faulting).
template <typename T>
void launchInputEmbedding(TensorWrapper<int>* input_ids, TensorWrapper<T>* output, EmbeddingWeight<T>* embed_table) {
    const int max_context_token_num = output->shape[0];
    const int hidden_size = output->shape[1];
    LLM_CHECK_WITH_INFO(max_context_token_num == input_ids->shape[0],
                        "input ids 1st shape should equal to 1st shape of output.");
    const int vocab = embed_table->shape.empty() ? 0x7fffffff : embed_table->shape[0];
    B200_CHECK_RC(b200_embedding(input_ids->data, embed_table->data, output->data, max_context_token_num,
                                 hidden_size, vocab, b200shim::dtype_of<T>(), b200shim::stream()));
}
