/*
 * b200_llama.h — C ABI of the B200-native (sm_100a) LLaMA decoder hot path.
 *
 * Every entry point replaces one `launch*` template of the reference
 * (Mr-wang27/llm-inference, the headers under `src/kernels`); the reference file:line each one
 * stands in for is cited above it. Plain pointers and sizes only: no C++ types,
 * no torch types. The reference's TensorWrapper-based templates are re-created on
 * top of this ABI by the header shim in `llm-inference_b200/include/src/kernels/`.
 *
 * Conventions
 *  - `dtype`: B200_F32 (float) or B200_F16 (IEEE half). All arithmetic accumulates in fp32.
 *  - all data pointers are DEVICE pointers unless the name ends in `_host`.
 *  - `stream` is a cudaStream_t passed as void* (NULL = legacy default stream, which is
 *    what the reference launches on).
 *  - return value: 0 on success; >0 is a cudaError_t from the launch; <0 is one of the
 *    B200_E* argument errors. Nothing allocates; scratch is passed in and sized by the
 *    matching *_workspace_bytes() query.
 *  - there is no CPU fallback anywhere: without a CUDA device every call returns an error.
 */
#ifndef B200_LLAMA_H
#define B200_LLAMA_H

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define B200_F32 0
#define B200_F16 1

#define B200_OK 0
#define B200_EINVAL (-1)       /* bad shape / null pointer / misalignment */
#define B200_EUNSUPPORTED (-2) /* shape outside what the kernel covers */
#define B200_EWORKSPACE (-3)   /* workspace missing or too small */
#define B200_ENOMEM (-4)       /* a fixed-size pool (the KV page pool of a paged engine) is exhausted */

/* library identity; b200_abi_version() bumps when a signature changes */
int b200_abi_version(void);
const char* b200_build_info(void);
/* number of kernel launches issued through this library since load (or last reset) */
uint64_t b200_launch_count(void);
void b200_launch_count_reset(void);

/* ------------------------------------------------------------------------------------
 * a1  launchRMSNorm            src/kernels/rmsnorm_kernel.cu:175 (decl rmsnorm_kernel.h:12-17)
 * in place x <- x * rsqrt(mean(x^2)+eps) * gamma ; residual_out <- x (pre-norm) if non-NULL
 * x, residual_out: [n_tokens, hidden]; gamma: [hidden]
 * ------------------------------------------------------------------------------------ */
int b200_rmsnorm(void* x, void* residual_out, const void* gamma, int n_tokens, int hidden,
                 float eps, int dtype, void* stream);

/* a5  launchFusedAddBiasResidualRMSNorm   src/kernels/fused_addresidual_norm.cu:193
 * residual <- residual + out (+ bias) ; out <- RMSNorm(residual) * gamma.  bias may be NULL. */
int b200_fused_add_bias_residual_rmsnorm(void* residual, void* out, const void* bias,
                                         const void* gamma, int n_tokens, int hidden, float eps,
                                         int dtype, void* stream);

/* a7  launchAddResidual        src/kernels/add_residual.cu:58    out <- out + residual */
int b200_add_residual(const void* residual, void* out, int n_tokens, int hidden, int dtype,
                      void* stream);

/* a6  launchAct (SwiGLU)       src/kernels/act_kernel.cu:57
 * gate_up: [n_tokens, 2, inter] (gate block, then up block); out: [n_tokens, inter]
 * out = gate/(1+exp(-gate)) * up */
int b200_swiglu(const void* gate_up, void* out, int n_tokens, int inter, int dtype, void* stream);

/* a3  launchRoPE               src/kernels/qkv_bias_and_RoPE.cu:410
 * rotate-half RoPE, in place, on q and k of the newest token (position step-1).
 * qkv: token-major [batch, head_num + 2*kv_head_num, head_size]. */
int b200_rope_decode(void* qkv, int batch, int head_num, int kv_head_num, int head_size, int step,
                     int rotary_dim, float rotary_base, int dtype, void* stream);
/* graph-replay form: `step_dev` (DEVICE int, may be NULL) overrides `step` at run time */
int b200_rope_decode_ex(void* qkv, int batch, int head_num, int kv_head_num, int head_size, int step,
                        const int* step_dev, int rotary_dim, float rotary_base, int dtype,
                        void* stream);

/* same with the DEVICE table of b200_rope_table() (rope_den may be NULL) */
int b200_rope_decode_tab(void* qkv, int batch, int head_num, int kv_head_num, int head_size, int step,
                         const int* step_dev, int rotary_dim, float rotary_base, const float* rope_den,
                         int dtype, void* stream);
/* ragged batch: token b sits at position seq_steps_dev[b] - 1 (DEVICE int [batch]) */
int b200_rope_decode_ragged(void* qkv, int batch, int head_num, int kv_head_num, int head_size,
                            const int* seq_steps_dev, int rotary_dim, float rotary_base,
                            const float* rope_den, int dtype, void* stream);

/* a8  launchAddFusedQKVBiasTransposeAndRoPE   src/kernels/qkv_bias_and_RoPE.cu:248
 * QKV [token_num, head_num+2*kv_head_num, head_size] (no padding) ->
 * q [batch, head_num, seq_len, head_size], k,v [batch, kv_head_num, seq_len, head_size];
 * optional bias [(head_num+2*kv_head_num)*head_size]; RoPE on q,k at
 * history_len[b] + local token index; padded rows of q/k/v are left untouched. */
int b200_qkv_bias_rope_transpose(void* q, void* k, void* v, const void* qkv, const void* bias,
                                 const int* padding_offset, const int* history_len,
                                 const int* input_len, int batch, int seq_len, int token_num,
                                 int head_num, int kv_head_num, int head_size, int rotary_dim,
                                 float rotary_base, int dtype, void* stream);

/* a8 + a9 in one pass (engine prefill): q as above; rotated k and v go straight to rows
 * history_len[b] + t of the cache layer slice [batch, kv_head_num, max_seq_len, head_size]. */
int b200_qkv_bias_rope_to_cache(void* q, void* k_cache_layer, void* v_cache_layer, const void* qkv,
                                const void* bias, const int* padding_offset, const int* history_len,
                                int batch, int seq_len, int token_num, int head_num, int kv_head_num,
                                int head_size, int max_seq_len, int rotary_dim, float rotary_base,
                                int dtype, void* stream);

/* a9  launchConcatKVCache      src/kernels/concat_past_kv.cu:103
 * k_src,v_src [batch, kv_head_num, max_q_len, head_size] -> cache layer slice
 * [batch, kv_head_num, max_seq_len, head_size] at row history_len[b]+t for t < cur_len[b].
 * k_cache/v_cache point at the START of the whole cache; layer_offset is in elements. */
int b200_concat_kv_cache(const void* k_src, const void* v_src, void* k_cache, void* v_cache,
                         size_t layer_offset, const int* cur_query_len, const int* history_len,
                         int batch, int kv_head_num, int max_q_len, int max_seq_len, int head_size,
                         int dtype, void* stream);

/* a10 launchRepeatKVCache      src/kernels/repeat_kv.cu:76
 * cache layer slice -> dense [batch, head_num, max_k_len, head_size], head j <- j/(head_num/kv_head_num),
 * rows < context_len[b] only (other rows untouched). */
int b200_repeat_kv_cache(const void* k_cache, const void* v_cache, void* k_dst, void* v_dst,
                         size_t layer_offset, const int* context_len, int batch, int head_num,
                         int kv_head_num, int max_seq_len, int max_k_len, int head_size, int dtype,
                         void* stream);

/* a12 launchTransposeOutRemovePadding   src/kernels/fused_transpose_and_remv_pad.cu:51
 * src [batch, head_num, seq_len, head_size] -> dst [token_num, head_num*head_size] */
int b200_transpose_remove_padding(const void* src, void* dst, const int* padding_offset,
                                  int token_num, int batch, int seq_len, int head_num,
                                  int head_size, int dtype, void* stream);

/* a13 launchCalPaddingoffset   src/kernels/cal_paddingoffset.cu:74   (int32, bit-exact)
 * padding_offset [batch*max_q_len] (first sum(len) entries written), cum_seqlens [batch+1] */
int b200_cal_padding_offset(int* padding_offset, int* cum_seqlens, const int* input_lens,
                            int batch, int max_q_len, void* stream);

/* a13 launchBuildCausalMasks   src/kernels/build_causal_mask.cu:50   (0/1 values, bit-exact)
 * mask [batch, max_q_len, max_k_len]. history_visible=1: the unit test's CPU predicate
 * (tests/unittests/test_casual_mask.cu:12-32: q<qlen && k<klen && k<=q+klen-qlen);
 * history_visible=0: the GPU kernel's extra `k >= klen-qlen` clause (build_causal_mask.cu:29). */
int b200_build_causal_mask(void* mask, const int* q_lens, const int* k_lens, int batch,
                           int max_q_len, int max_k_len, int history_visible, int dtype,
                           void* stream);

/* a11 launchScaleMaskAndSoftmax   src/kernels/attn_softmax_kernel.cu:337
 * score = softmax_row(scale*qk + (1-mask)*(-10000)), denominator + 1e-6; may run in place.
 * qk/score [batch, head_num, q_len, k_len]; mask [batch, q_len, k_len]. */
int b200_scale_mask_softmax(const void* qk, const void* mask, void* score, int batch, int head_num,
                            int q_len, int k_len, float scale, int dtype, void* stream);

/* ------------------------------------------------------------------------------------
 * a2  launchLinearGemm         src/kernels/linear.cu:39  (cublasWrapper::Gemm cublas_utils.cc:32)
 * y[n_tokens, N] = x[n_tokens, K] * op(W) (+ bias[N] if non-NULL)
 *   trans_b = 1: W is [N, K] row-major (the layout of every engine call site)
 *   trans_b = 0: W is [K, N] row-major (tests/unittests/test_linear.cu)
 * Dispatch: n_tokens <= B200_GEMV_MAX_TOKENS and trans_b -> weight-streaming GEMV (bulk-async
 * smem ring); fp16 with large n_tokens -> tcgen05/TMEM GEMM; everything else -> SIMT tiled GEMM.
 * workspace may be NULL when b200_linear_workspace_bytes() returns 0.
 * ------------------------------------------------------------------------------------ */
#define B200_GEMV_MAX_TOKENS 8
size_t b200_linear_workspace_bytes(int n_tokens, int N, int K, int trans_b, int dtype);
int b200_linear(const void* x, const void* W, const void* bias, void* y, int n_tokens, int N, int K,
                int trans_b, int dtype, void* workspace, size_t workspace_bytes, void* stream);

/* gate_up GEMM + SwiGLU in one launch (ffn.cpp:72-80): out[n_tokens, inter] = silu(x Wg^T) * (x Wu^T),
 * W = [2*inter, K], gate rows first. tcgen05 GEMM with the activation as its epilogue (same fp16
 * rounding points as b200_linear followed by b200_swiglu). B200_EUNSUPPORTED outside fp16 prefill
 * shapes (n_tokens > 256, inter % 128 == 0): run b200_linear + b200_swiglu then. */
int b200_linear_swiglu(const void* x, const void* W, void* out, int n_tokens, int inter, int K,
                       int dtype, void* stream);

/* linear + launchAddResidual (add_residual.cu:58) in one launch: y = residual + x W^T, linear output
 * rounded to fp16 before the sum as in the unfused pair. fp16 prefill shapes only, else
 * B200_EUNSUPPORTED (run b200_linear + b200_add_residual). */
int b200_linear_residual(const void* x, const void* W, const void* residual, void* y, int n_tokens,
                         int N, int K, int dtype, void* stream);

/* Fused decode-time linears (engine path; the unfused reference sequence they replace is cited).
 * All take W [N, K] row-major, n_tokens <= B200_GEMV_MAX_TOKENS.
 *
 * flags for b200_gemv_fused: */
#define B200_GEMV_NORM 1u     /* prologue: x <- (x*inv_rms)*gamma in smem (x itself untouched)  */
#define B200_GEMV_RESIDUAL 2u /* epilogue: y <- y_in + W x (y is read-modify-write)             */
#define B200_GEMV_SWIGLU 4u   /* W = [2*N, K] (gate rows, then up rows); y[N] = silu(g)*u        */
#define B200_GEMV_ROPE_KV 8u  /* W = qkv; epilogue rotates q,k (pos step-1), writes qkv to y and */
                              /* k,v also straight into the KV cache row step-1                 */
#define B200_GEMV_NORM_GFIRST 16u /* with NORM: (gamma*x)*inv_rms, the rounding order of        */
                                  /* fused_addresidual_norm.cu:119-124 instead of rmsnorm's     */
typedef struct b200_gemv_args {
    const void* x;     /* [n_tokens, K]                                   */
    const void* W;     /* [N(or 2N), K]                                   */
    const void* bias;  /* [N] or NULL                                     */
    void* y;           /* [n_tokens, N]                                   */
    const void* gamma; /* [K], with B200_GEMV_NORM                        */
    float eps;
    int n_tokens, N, K;
    unsigned flags;
    /* B200_GEMV_ROPE_KV only: N = (head_num+2*kv_head_num)*head_size */
    void* k_cache_layer; /* [batch, kv_head_num, max_seq_len, head_size] */
    void* v_cache_layer;
    int head_num, kv_head_num, head_size, max_seq_len, step, rotary_dim;
    float rotary_base;
    const int* step_dev; /* optional DEVICE int overriding `step` (CUDA-graph replay) */
    const float* rope_den; /* optional DEVICE table [rotary_dim/2] of base^(2i/rotary_dim) from     */
                           /* b200_rope_table(); NULL = computed in the kernel                   */
    void* p2p; /* optional b200_p2p* (b200_engine.h): tensor-parallel row-split linear, fp16, no epilogue */
               /* mode: the partial outputs go straight to every rank's receive buffer over NVLink     */
    int step_stride; /* 0: *step_dev is the step all tokens share; 1: step_dev[token] (ragged batch)          */
               /* (16-byte lines, payload + call tag) instead of y; b200_p2p_reduce_add() finishes     */
} b200_gemv_args;
int b200_gemv_fused(const b200_gemv_args* args, int dtype, void* stream);
/* fills den[i] = base^(2i/rotary_dim), i < rotary_dim/2 (correctly rounded; see common.cuh) */
int b200_rope_table(float* den, int rotary_dim, float rotary_base, void* stream);

/* launchLinearStridedBatchGemm  src/kernels/linear.cu:127 (cublas_utils.cc:75)
 * C[b] = A[b] * op(B[b]) for b < batch_count; A [m,k] row-major; B [n,k] if trans_b else [k,n];
 * C [m,n]; strides in elements (the reference's strideC defect D14 is not reproduced). */
int b200_batched_gemm(const void* A, const void* B, void* C, int batch_count, int m, int n, int k,
                      int trans_b, int dtype, void* stream);

/* ------------------------------------------------------------------------------------
 * a4  launchDecoderMaskedMHA   src/kernels/fused_decoder_self_attention.cu:317
 * One new token per sequence. qkv [batch, head_num+2*kv_head_num, head_size] (token-major),
 * caches point at the layer slice [batch, kv_head_num, max_seq_len, head_size].
 * append_kv=1 (reference semantics): write k,v of the new token to row step-1, then attend over
 * rows [0, step). append_kv=0: the row is already in the cache (fused QKV GEMV wrote it).
 * out [batch, head_num*head_size]. softmax: max-subtracted, denominator + 1e-6 (:214).
 * Split-KV: grid (batch*kv_head_num, n_splits); partials in workspace; combine kernel.
 * ------------------------------------------------------------------------------------ */
size_t b200_decode_attention_workspace_bytes(int batch, int head_num, int head_size, int step);
int b200_decode_attention(const void* qkv, void* k_cache_layer, void* v_cache_layer, void* out,
                          void* workspace, size_t workspace_bytes, int batch, int head_num,
                          int kv_head_num, int head_size, int max_seq_len, int step, int append_kv,
                          int dtype, void* stream);
/* graph-replay form: `step_dev` (DEVICE int, may be NULL) overrides `step` at run time; the
 * launch geometry is sized for `step` (an upper bound of *step_dev within the bucket). */
int b200_decode_attention_ex(const void* qkv, void* k_cache_layer, void* v_cache_layer, void* out,
                             void* workspace, size_t workspace_bytes, int batch, int head_num,
                             int kv_head_num, int head_size, int max_seq_len, int step,
                             const int* step_dev, int append_kv, int dtype, void* stream);
/* ragged batch: sequence b attends over seq_steps_dev[b] keys (DEVICE int [batch]; its new token
 * is row seq_steps_dev[b]-1); max_step >= every entry sizes the split count and the workspace. */
int b200_decode_attention_ragged(const void* qkv, void* k_cache_layer, void* v_cache_layer,
                                 void* out, void* workspace, size_t workspace_bytes, int batch,
                                 int head_num, int kv_head_num, int head_size, int max_seq_len,
                                 int max_step, const int* seq_steps_dev, int append_kv, int dtype,
                                 void* stream);

/* `finished` of launchDecoderMaskedMHA (fused_decoder_self_attention.h:10-19; the reference passes it and never reads
 * it): DEVICE flags [batch] applied to every b200_decode_attention* / b200_kv_append_fp8 launch issued afterwards
 * from this host thread (NULL = off). A set flag retires the sequence: no K/V append, no cache reads, zero output. */
void b200_decode_attention_set_finished(const unsigned char* finished_dev);

/* Allocation-free variant of b200_decode_attention. No counterpart in the reference: added
 * for callers that cannot call cudaMalloc (CUDA-graph capture, tight
 * serving loops, harnesses that forbid it). Same arguments minus the workspace: the split-KV partials
 * live in a file-scope `static __device__` array of B200_DECODE_ATTENTION_STATIC_MAX_BYTES inside the
 * library (every GPU that loads the library holds it until exit); its device address is resolved once
 * per device by cudaGetSymbolAddress on the first call made with that device current (make that call
 * outside stream capture). Calls no cudaMalloc / cudaFree / allocating routine. A shape whose scratch
 * exceeds the maximum returns B200_EWORKSPACE and prints the needed size to stderr - use
 * b200_decode_attention with a caller-owned workspace then. All calls share the array: nothing in it
 * needs clearing between calls (every partial is written before it is read), but two calls that
 * overlap on the GPU (different streams without a dependency) race on it. */
#define B200_DECODE_ATTENTION_STATIC_MAX_BYTES (8u << 20)
int b200_decode_attention_static(const void* qkv, void* k_cache_layer, void* v_cache_layer, void* out,
                                 int batch, int head_num, int kv_head_num, int head_size,
                                 int max_seq_len, int step, int append_kv, int dtype, void* stream);

/* ------------------------------------------------------------------------------------
 * fp8 (e4m3) KV cache, SURVEY 8 f4 - the evolution of the reference's T-typed cache
 * [layers,B,hk,max_seq,d] (src/models/llama/llama.cpp:77-78,102-103): one byte per element + one fp32 scale
 * per cached row ([B, hk, max_seq] per layer), value = scale * e4m3, scale = amax(row) / 448 fixed at append.
 * fp16 models, head size 128, head_num / kv_head_num in {1,2,4}.
 *   b200_kv_append_fp8       the new token's rotated k, v (qkv [batch, h+2hk, d]) -> cache row step-1
 *   b200_decode_attention_fp8 launchDecoderMaskedMHA (fused_decoder_self_attention.cu:317) over rows [0, step) of
 *                            that cache; the new row must already be there; workspace as b200_decode_attention
 *   b200_kv_quantize_fp8     rows t < n_rows[b] of an fp16 slice [batch, hk, src_seq, d] -> cache rows dst_row0[b] + t
 *                            (the prompt pass: launchConcatKVCache concat_past_kv.cu:103 with quantisation)
 *   b200_kv_dequant_fp8      slice back to fp16 (tests)
 * step_dev (DEVICE int, may be NULL) overrides step; step_stride 1 = one entry per sequence (ragged batch).
 * ------------------------------------------------------------------------------------ */
int b200_kv_append_fp8(const void* qkv, void* k_cache8, void* v_cache8, float* k_scale, float* v_scale, int batch,
                       int head_num, int kv_head_num, int head_size, int max_seq_len, int step, const int* step_dev,
                       int step_stride, void* stream);
int b200_decode_attention_fp8(const void* qkv, const void* k_cache8, const void* v_cache8, const float* k_scale,
                              const float* v_scale, void* out, void* workspace, size_t workspace_bytes, int batch,
                              int head_num, int kv_head_num, int head_size, int max_seq_len, int step,
                              const int* step_dev, int step_stride, void* stream);
int b200_kv_quantize_fp8(const void* src, void* dst8, float* scale, const int* n_rows, const int* dst_row0, int batch,
                         int kv_head_num, int src_seq, int max_seq_len, int head_size, int max_rows, void* stream);
int b200_kv_dequant_fp8(const void* src8, const float* scale, void* dst, size_t rows, int head_size, void* stream);

/* ------------------------------------------------------------------------------------
 * Paged KV cache, SURVEY 8 f4 - the other evolution of [layers,B,hk,max_seq,d] (src/models/llama/llama.cpp:77-78):
 * a layer's K (V) is a pool of pages [n_pages, hk, B200_KV_PAGE_ROWS, d]; sequence b owns the pages listed in
 * page_table[b * max_pages + i] (DEVICE int, one list for all layers), cache row r of the sequence = row r % 128 of
 * page r / 128. No reference counterpart: the reference reserves max_seq rows per sequence slot.
 *   b200_decode_attention_paged  launchDecoderMaskedMHA (fused_decoder_self_attention.cu:317) over the pages: same
 *                            kernel, split / combine and semantics as b200_decode_attention_ragged (append_kv writes
 *                            the new row into the page that holds row step-1, which the list must already name);
 *                            fp16, head size 128, head_num / kv_head_num in {1,2,4,8}. seq_steps_dev NULL: every
 *                            sequence is at max_step; step_stride 0: one shared device step, 1: one per sequence.
 *   b200_kv_scatter_pages    rows t < n_rows[b] of a dense slice [batch, hk, src_seq, d] -> rows dst_row0[b] + t of the
 *                            sequence's pages (the prompt pass: launchConcatKVCache concat_past_kv.cu:103 into pages)
 *   b200_kv_gather_pages     rows [0, n_rows[b]) of the pages -> the dense slice [batch, hk, dst_seq, d] (tests)
 * elem_bytes: 2 (fp16), 4 (fp32) or 1; n_rows / dst_row0: DEVICE int [batch]; max_rows >= every n_rows[b].
 * ------------------------------------------------------------------------------------ */
#define B200_KV_PAGE_ROWS 128
int b200_decode_attention_paged(const void* qkv, void* k_pool_layer, void* v_pool_layer, const int* page_table,
                                int max_pages, void* out, void* workspace, size_t workspace_bytes, int batch,
                                int head_num, int kv_head_num, int head_size, int max_step, const int* seq_steps_dev,
                                int step_stride, int append_kv, void* stream);
/* b200_qkv_bias_rope_to_cache into a paged cache: k, v of token t of sequence b -> row history_len[b] + t of its pages */
int b200_qkv_bias_rope_to_pages(void* q, void* k_pool_layer, void* v_pool_layer, const int* page_table, int max_pages,
                                const void* qkv, const void* bias, const int* padding_offset, const int* history_len,
                                int batch, int seq_len, int token_num, int head_num, int kv_head_num, int head_size,
                                int rotary_dim, float rotary_base, int dtype, void* stream);
/* b200_context_attention over a paged cache: a page is exactly one 128-key tile of the kernel (one list lookup per tile in
 * the TMA producer). Every page the context touches must be in the list and hold this pass's rows already. */
int b200_context_attention_paged(const void* q, const void* k_pool_layer, const void* v_pool_layer, const int* page_table,
                                 int max_pages, int n_pages, void* out, const int* padding_offset, const int* input_len,
                                 const int* history_len, int batch, int seq_len, int token_num, int head_num, int kv_head_num,
                                 int head_size, int dtype, void* stream);
int b200_kv_scatter_pages(const void* src, void* pool, const int* page_table, int max_pages, const int* n_rows,
                          const int* dst_row0, int batch, int kv_head_num, int src_seq, int head_size, int elem_bytes,
                          int max_rows, void* stream);
int b200_kv_gather_pages(const void* pool, void* dst, const int* page_table, int max_pages, const int* n_rows, int batch,
                         int kv_head_num, int dst_seq, int head_size, int elem_bytes, int max_rows, void* stream);
/* fp8 pages: the four fp8 entries above over pools [n_pages, hk, 128, d] of e4m3 bytes + scale pools [n_pages, hk, 128]
 * (one fp32 per cached row) and the same page lists; arguments as their dense counterparts, max_seq_len replaced by
 * (page_table, max_pages). b200_kv_dequant_fp8_paged writes rows [0, n_rows[b]) as fp16 into a dense slice (tests). */
int b200_kv_append_fp8_paged(const void* qkv, void* k_pool8, void* v_pool8, float* k_scale_pool, float* v_scale_pool,
                             const int* page_table, int max_pages, int batch, int head_num, int kv_head_num, int head_size,
                             int step, const int* step_dev, int step_stride, void* stream);
int b200_decode_attention_fp8_paged(const void* qkv, const void* k_pool8, const void* v_pool8, const float* k_scale_pool,
                                    const float* v_scale_pool, const int* page_table, int max_pages, void* out,
                                    void* workspace, size_t workspace_bytes, int batch, int head_num, int kv_head_num,
                                    int head_size, int step, const int* step_dev, int step_stride, void* stream);
int b200_kv_quantize_fp8_paged(const void* src, void* pool8, float* scale_pool, const int* page_table, int max_pages,
                               const int* n_rows, const int* dst_row0, int batch, int kv_head_num, int src_seq, int head_size,
                               int max_rows, void* stream);
int b200_kv_dequant_fp8_paged(const void* pool8, const float* scale_pool, void* dst, const int* page_table, int max_pages,
                              const int* n_rows, int batch, int kv_head_num, int dst_seq, int head_size, int max_rows,
                              void* stream);

/* Prefill attention without materialising scores / mask / repeated KV (replaces
 * context_attention.cpp:136-161: RepeatKV + QK^T + ScaleMaskSoftmax + PV + transpose):
 * q [batch, head_num, seq_len, head_size]; cache layer slice as above;
 * out [token_num, head_num*head_size] (padding removed). Same -10000/1e-6 softmax contract. */
int b200_context_attention(const void* q, const void* k_cache_layer, const void* v_cache_layer,
                           void* out, const int* padding_offset, const int* input_len,
                           const int* history_len, int batch, int seq_len, int token_num,
                           int head_num, int kv_head_num, int head_size, int max_seq_len,
                           int dtype, void* stream);

/* debug: %globaltimer stamps of CTA 0 of the following b200_context_attention launches (NULL = off);
 * layout in csrc/context_attention.cu, reader tools/ctx_trace.py */
void b200_context_attention_set_trace(void* dev_buf);

/* ------------------------------------------------------------------------------------
 * next-tier ops bracketing a decode step
 * launchInputEmbedding   src/kernels/input_embedding.cu:30   out[t,:] = table[ids[t],:]
 * greedy pick            src/kernels/topK.cu:147 (K=1 case): lowest index wins ties
 * ------------------------------------------------------------------------------------ */
int b200_embedding(const int* ids, const void* table, void* out, int n_tokens, int hidden,
                   int vocab, int dtype, void* stream);
int b200_argmax(const void* logits, int* out_ids, int batch, int vocab, int dtype, void* stream);
/* top-K (K <= 8) values+ids per row, descending, ties -> lower index first */
int b200_topk(const void* logits, int* out_ids, void* out_vals, int batch, int vocab, int k,
              int dtype, void* stream);

#ifdef __cplusplus
}
#endif
#endif /* B200_LLAMA_H */
