// Static / dynamic attention parameters crossing the kernel boundary.
// Byte-compatible with reference src/models/llama/llama_params.h:3-8, :30-37.
#pragma once
struct LLaMAAttentionStaticParams {
    int rotary_embedding_dim;
    float rotary_embedding_base;
    int max_position_embeddings;
    bool use_dynamic_ntk;  // carried, unused by any kernel (as in the reference)
};
struct LLaMAAttentionDynParams {
    int batch_size;
    int num_tokens;
    int max_q_len;
    int max_k_len;
    int num_layers;
    bool is_ctx = false;
};
