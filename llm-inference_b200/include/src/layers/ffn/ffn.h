// LLaMAFFNLayer<T> — same constructor / forward signature / TensorMap keys as the reference's
// src/layers/ffn/ffn.{h,cpp} ("ffn_input" -> "ffn_output", ffn.cpp:62-63), re-written over the
// B200 launchers. Header-only: the launch* shim has no CUDA syntax, so no .cu instantiation file
// is needed. Differences from the reference, all deliberate:
//   * scratch is kept between calls and re-allocated only when it must grow (the reference
//     Malloc/Frees two buffers per layer per token and device-syncs after each launch,
//     ffn.cpp:40-51,73-90);
//   * launches are stream-ordered on b200shim::current_stream(); no host synchronisation.
#pragma once
#include "src/kernels/act_kernel.h"
#include "src/kernels/cublas_utils.h"
#include "src/kernels/linear.h"
#include "src/memory/allocator/cuda_allocator.h"
#include "src/models/llama/llama_params.h"
#include "src/utils/macro.h"
#include "src/utils/tensor.h"
#include "src/weights/llama/ffn_weights.h"

template <typename T>
class LLaMAFFNLayer {
private:
    const int head_num;
    const int head_size;
    const int inter_size;
    const int hidden_units;
    int count = -1;  // layer index currently being run (reference ffn.h: debug counter)
    cudaStream_t stream;
    cublasWrapper* cublas_wrapper;
    BaseAllocator* allocator;
    TensorWrapper<T>* SwiGLU_input = nullptr;     // [num_tokens, 2, inter_size]
    TensorWrapper<T>* down_proj_input = nullptr;  // [num_tokens, inter_size]
    int capacity_tokens = 0;

    void ensure(int num_tokens) {
        if (num_tokens <= capacity_tokens) {
            SwiGLU_input->shape = {num_tokens, 2, inter_size};
            down_proj_input->shape = {num_tokens, inter_size};
            return;
        }
        freeBuf();
        const DataType type = getTensorType<T>();
        SwiGLU_input = new TensorWrapper<T>(Device::GPU, type, {num_tokens, 2, inter_size});
        down_proj_input = new TensorWrapper<T>(Device::GPU, type, {num_tokens, inter_size});
        SwiGLU_input->data = allocator->Malloc(SwiGLU_input->data, sizeof(T) * (size_t)num_tokens * 2 * inter_size, false);
        down_proj_input->data = allocator->Malloc(down_proj_input->data, sizeof(T) * (size_t)num_tokens * inter_size, false);
        capacity_tokens = num_tokens;
    }

public:
    LLaMAFFNLayer(int head_num, int head_size, int inter_size, cudaStream_t stream, cublasWrapper* cublas_wrapper,
                  BaseAllocator* allocator)
        : head_num(head_num), head_size(head_size), inter_size(inter_size), hidden_units(head_num * head_size),
          stream(stream), cublas_wrapper(cublas_wrapper), allocator(allocator) {}
    ~LLaMAFFNLayer() { freeBuf(); }
    LLaMAFFNLayer(const LLaMAFFNLayer&) = delete;
    LLaMAFFNLayer& operator=(const LLaMAFFNLayer&) = delete;

    void allocForForward(LLaMAAttentionDynParams& params) { ensure(params.num_tokens); }  // prefill
    void allocForForward(int batch_size) { ensure(batch_size); }                          // decode
    void freeBuf() {
        if (SwiGLU_input) { allocator->Free(SwiGLU_input->data); delete SwiGLU_input; SwiGLU_input = nullptr; }
        if (down_proj_input) { allocator->Free(down_proj_input->data); delete down_proj_input; down_proj_input = nullptr; }
        capacity_tokens = 0;
    }

    // inputs["ffn_input"] [num_tokens, hidden] -> outputs["ffn_output"] [num_tokens, hidden]
    // (may alias, as the decoders pass the same buffer): gate_up GEMM -> SwiGLU -> down GEMM.
    void forward(TensorMap& inputs, TensorMap& outputs, LLaMAFFNWeights<T>& weights, LLaMAAttentionDynParams& params) {
        (void)params;  // the reference sizes scratch from params (ffn.cpp:55-60); the input's own row count is exact
        Tensor* ffn_input = inputs["ffn_input"];
        Tensor* ffn_output = outputs["ffn_output"];
        ensure(ffn_input->shape[0]);
        count += 1;
        launchLinearGemm(ffn_input->as<T>(), weights.gateAndup, SwiGLU_input, cublas_wrapper, false, true);
        launchAct(SwiGLU_input, down_proj_input);
        launchLinearGemm(down_proj_input, weights.down, ffn_output->as<T>(), cublas_wrapper, false, true);
    }
};
