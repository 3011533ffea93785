/*
 * b200_engine.h — C ABI of the decode-step driver built on b200_llama.h.
 *
 * Stands in for the reference's per-token path:
 *   Llama<T>::continueTokenGen            src/models/llama/llama.cpp:319-361
 *   LlamaSelfDecoder<T>::forward          src/layers/decoder/self_decoder.cpp:24-85
 *   LLaMASelfAttentionLayer<T>::Forward   src/layers/attention/masked_self_attention.cpp:55-92
 *   LLaMAFFNLayer<T>::forward             src/layers/ffn/ffn.cpp:53-93
 *   Llama<T>::LMHeadAndTopKSample         src/models/llama/llama.cpp:215-271 (greedy: top-1)
 * with the per-call Malloc/Free and cudaDeviceSynchronize calls of those files replaced by a
 * persistent arena (src/memory change), stream-ordered launches and CUDA-graph replay.
 *
 * Memory layout (one cudaMalloc each, 256-byte aligned sub-allocations):
 *   weights arena : per layer [attn_norm γ | qkv W[(h+2hk)d, H] | o W[H, hd] | ffn_norm γ |
 *                   gate_up W[2I, H] (gate rows first) | down W[H, I]], then embedding [V,H],
 *                   final norm γ, lm_head W[V,H]  — all row-major [N,K] (the reference's
 *                   trans_b=true layout), which is exactly what the GEMV's bulk copies and the
 *                   GEMM's TMA maps read; no repacking.
 *   KV cache      : K and V each [layers, batch, kv_heads, max_seq_len, head_size]
 *                   (reference llama.cpp:77-78,102-103), 64-bit indexed.
 *   activations   : hidden [B,H], qkv [B,(h+2hk)d], attn [B,hd], act [B,I] (fused) or
 *                   gate_up [B,2,I] (unfused), logits [B,V], token ids, device-side step.
 *
 * Tensor parallelism (config 5): tp_size ranks each hold h/tp query heads, hk/tp kv heads and
 * I/tp FFN columns; o_proj and down_proj produce partial sums that the caller-provided
 * allreduce hook sums after each (the only two collective sites).
 */
#ifndef B200_ENGINE_H
#define B200_ENGINE_H

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

typedef struct b200_engine b200_engine;

typedef struct b200_model_config {
    int num_layers;
    int head_num;     /* GLOBAL query heads  */
    int kv_head_num;  /* GLOBAL kv heads     */
    int head_size;
    int inter_size;   /* GLOBAL FFN width    */
    int vocab_size;
    int max_seq_len;
    int max_batch;
    float rms_eps;
    int rotary_dim;
    float rotary_base;
    int dtype;        /* B200_F32 / B200_F16 */
    int tp_size;      /* 1 = no tensor parallelism */
    int tp_rank;
} b200_model_config;

/* weight slots for b200_engine_set_weight / get_weight (layer ignored for the last three) */
enum {
    B200_W_ATTN_NORM = 0,
    B200_W_QKV = 1,
    B200_W_O = 2,
    B200_W_FFN_NORM = 3,
    B200_W_GATE_UP = 4,
    B200_W_DOWN = 5,
    B200_W_EMBED = 6,
    B200_W_FINAL_NORM = 7,
    B200_W_LM_HEAD = 8
};

/* creation flags for b200_engine_create_ex */
#define B200_ENGINE_KV_FP8 1u  /* KV cache in fp8 e4m3 with one fp32 scale per cached row (replaces the T-typed
                                * [layers,B,hk,max_seq,d] caches of llama.cpp:77-78). fp16 models, head size 128,
                                * head_num / kv_head_num in {1,2,4}. A row is rounded once, when it enters the cache; decode
                                * runs the reference launch order (RoPE -> b200_kv_append_fp8 -> b200_decode_attention_fp8);
                                * only the FIRST prompt pass of a sequence is supported. Not bit-compatible with the fp16
                                * cache: tests/test_kv_fp8_gpu.py pins it to the oracle run on the same rounded rows. */

#define B200_ENGINE_KV_PAGED 2u /* KV cache as page pools [layers][n_pages, hk, 128 rows, d] + one page list per sequence
                                * (SURVEY 8 f4 "paged blocks"; b200_llama.h, csrc/kv_paged.cu) instead of max_seq rows
                                * reserved per sequence slot. fp16 models, head size 128, head_num / kv_head_num in
                                * {1,2,4,8}. Bit-identical results to the dense cache on the same launch order
                                * (tests/test_kv_paged_gpu.py). With KV_FP8 as well: fp8 pages + scale pools
                                * (head_num / kv_head_num in {1,2,4}), bit-identical to the dense fp8 cache. Decode runs the reference launch order
                                * with b200_decode_attention_paged; a prompt pass scatters its rows into the pages behind
                                * the history and attends with b200_context_attention_paged (fp8 pages: first prompt
                                * pass of a sequence only). Pages are taken when a sequence grows into them, returned by
                                * reset / b200_engine_release_finished; an empty pool fails the call with B200_ENOMEM. */

/* mode flags for b200_engine_set_mode */
#define B200_ENGINE_FUSED 1u   /* fused GEMV prologues/epilogues (default)                    */
#define B200_ENGINE_GRAPH 2u   /* CUDA-graph replay of the whole step (default)               */
#define B200_ENGINE_PERSISTENT 4u /* with FUSED: ONE persistent cooperative kernel per step     */
/* Default: all three for fp16 batches 1-2 (and fp32 / tensor-parallel engines, where PERSISTENT only applies once a
 * b200_p2p communicator is attached and the mode asks for it); FUSED | GRAPH for fp16 batches >= 3 (the measured best,
 * DESIGN 4c): fused tensor-core GEMVs up to 4 token rows; above, the reference launch order on the tcgen05
 * weight-streaming GEMM. */
                                  /* (fp16, batch <= 4, tp_size 1; default where it applies)   */
/* without FUSED the engine issues the reference's kernel sequence one launch* at a time
 * (rmsnorm, linear, rope, masked MHA, linear, fused add+norm, linear, act, linear, add) */

/* allreduce hook: sum `count` elements of `buf` (dtype as the model) in place across the
 * tensor-parallel group, enqueued on `stream`. NULL when tp_size == 1. */
typedef int (*b200_allreduce_fn)(void* buf, size_t count, int dtype, void* stream, void* user);

/* One-shot NVLink allreduce fused with the residual add (csrc/p2p_allreduce.cu): the product path of
 * tensor-parallel decode; ncclAllReduce through the hook above is the baseline. One communicator per
 * process (= per GPU): create -> exchange the 64-byte IPC handles of all ranks by any means (the
 * Python plumbing uses torch.distributed.all_gather_object) -> connect. */
typedef struct b200_p2p b200_p2p;
size_t b200_p2p_handle_bytes(void);
b200_p2p* b200_p2p_create(int world, int rank, size_t max_message_bytes);
int b200_p2p_local_handle(b200_p2p* c, void* out_handle);
int b200_p2p_connect(b200_p2p* c, const void* all_handles /* world x handle_bytes, rank-major */);
/* hidden[count] += sum over ranks of partial[count] (same rounding points as allreduce + AddResidual) */
int b200_p2p_allreduce_add(b200_p2p* c, const void* partial, void* hidden, size_t count, int dtype, void* stream);
/* fused form: b200_gemv_fused with args->p2p = comm (fp16, no epilogue mode, N % 4 == 0) pushes each
 * finished row pair straight into the peers' receive buffers from the GEMV epilogue (y is not
 * written); this call then waits for the lines of all ranks, sums and adds: hidden[count] += sum. */
int b200_p2p_reduce_add(b200_p2p* c, void* hidden, size_t count, int dtype, void* stream);
/* buf[count] <- sum over ranks of buf[count], in place (the batched / prefill paths) */
int b200_p2p_allreduce(b200_p2p* c, void* buf, size_t count, int dtype, void* stream);
size_t b200_p2p_max_message_bytes(const b200_p2p* c);
/* Frees this rank's receive buffer and unmaps the peers'. ALL ranks must have finished (stream-synchronised) their
 * last call on the communicator and then passed a barrier of the caller's (e.g. torch.distributed.barrier) before any
 * rank destroys it: a lagging peer would otherwise still push into, or read from, freed memory. The library only
 * synchronises the local device. */
void b200_p2p_destroy(b200_p2p* c);

b200_engine* b200_engine_create(const b200_model_config* cfg, int batch);
/* the same with B200_ENGINE_KV_* creation flags (0 = b200_engine_create) */
b200_engine* b200_engine_create_ex(const b200_model_config* cfg, int batch, unsigned flags);
/* Sequence retirement, the reference's `finished` flags (src/kernels/sampling.cu:46,81-83, llama.cpp:189,327):
 * with eos_id >= 0 a sequence that has produced eos_id is marked on the DEVICE, keeps emitting eos_id, and from the
 * next step on neither appends to nor reads its KV cache (the attention kernels skip it). -1 = off (default).
 * Flags are cleared by reset / fill_kv_random / every prompt pass. */
int b200_engine_set_eos(b200_engine* e, int eos_id);
int b200_engine_get_finished(b200_engine* e, unsigned char* flags_host /* [batch] */);
/* Continuous batching, data-path half: one sequence slot back to "nothing cached" (flag cleared, pages returned) while the
 * others keep their caches and positions; then b200_engine_prefill_ragged with length 0 for the slots that sit the pass out
 * (they keep the token they were about to decode). The scheduler that decides WHEN is the caller's. No reference
 * counterpart: llama.cpp:319-361 runs a fixed batch to the end. */
int b200_engine_reset_slot(b200_engine* e, int slot);
/* create_ex with the size of each layer's page pool (B200_ENGINE_KV_PAGED; 0 = batch * ceil(max_seq_len / 128)) */
b200_engine* b200_engine_create_paged(const b200_model_config* cfg, int batch, unsigned flags, int n_pages);
int b200_engine_kv_is_paged(const b200_engine* e);
/* paged cache: pool size / pages currently free / DEVICE page lists [batch, ceil(max_seq_len / 128)] */
int b200_engine_kv_pages_total(const b200_engine* e);
int b200_engine_kv_pages_free(const b200_engine* e);
const int* b200_engine_kv_page_table(const b200_engine* e);
/* paged cache: the pages of every sequence whose `finished` flag is set (b200_engine_set_eos) go back to the pool;
 * returns the number of pages released (>= 0). Synchronises the engine's stream. */
int b200_engine_release_finished(b200_engine* e);
/* paged cache: rows [0, cached rows of sequence b) of K and V gathered into caller DEVICE buffers laid out as the
 * reference's dense cache [layers, B, hk, max_seq, d] (tests; rows beyond a sequence's length are left untouched) */
int b200_engine_kv_gather(b200_engine* e, void* k_dst, void* v_dst);
int b200_engine_kv_is_fp8(const b200_engine* e);
/* fp8 cache only: K and V dequantised to fp16 into caller DEVICE buffers [layers, B, hk, max_seq, d] (tests) */
int b200_engine_kv_dequant(b200_engine* e, void* k_dst, void* v_dst);
void b200_engine_destroy(b200_engine* e);
const char* b200_engine_last_error(void);

size_t b200_engine_weight_bytes(const b200_engine* e);      /* this rank's weights         */
size_t b200_engine_kv_bytes(const b200_engine* e);          /* K + V cache                 */
size_t b200_engine_weight_elems(const b200_engine* e, int slot);
int b200_engine_set_mode(b200_engine* e, unsigned flags);
int b200_engine_set_allreduce(b200_engine* e, b200_allreduce_fn fn, void* user);
/* use the fused one-shot NVLink allreduce + residual add instead of the hook (NULL = back to the hook) */
int b200_engine_set_p2p(b200_engine* e, b200_p2p* comm);
void* b200_engine_stream(b200_engine* e);

/* weights: deterministic N(0, std^2) fill on the device (norm weights 1 + N(0,std^2)), or
 * explicit upload/download of one tensor (host pointers, element type = model dtype) */
int b200_engine_init_random(b200_engine* e, uint64_t seed, float std);
int b200_engine_set_weight(b200_engine* e, int slot, int layer, const void* host_src, size_t bytes);
int b200_engine_get_weight(b200_engine* e, int slot, int layer, void* host_dst, size_t bytes);

/* KV cache: fill rows [0, len) of every layer with U(-amp, amp) (bench configs that start from a
 * given context length) and set the step so the next decoded token sits at position `len`. */
int b200_engine_fill_kv_random(b200_engine* e, int len, uint64_t seed, float amp);
int b200_engine_reset(b200_engine* e);            /* step <- 0 (cache contents kept)          */
int b200_engine_get_step(const b200_engine* e);   /* tokens already in the cache              */

/* One decode step for the whole batch, all on the engine's stream, no host sync:
 *   ids_in (device int32 [batch], NULL = use the ids produced by the previous step)
 *   -> embedding -> layers -> final norm -> LM head -> argmax -> ids_out (device, inside engine)
 * The new token's K/V land at row `step`, then step <- step + 1. */
int b200_engine_decode_step(b200_engine* e, const int* ids_in_dev);
/* The call a user of the reference makes per token (llama.cpp:443 H2D id, :266 D2H id):
 * host ids in (pinned or pageable), one decode step, host ids out; synchronises. */
int b200_engine_decode_step_host(b200_engine* e, const int* ids_in_host, int* ids_out_host);

/* The prompt pass (Llama<T>::firstTokenGen, llama.cpp:274-318): `len` tokens per sequence
 * (ids_host int32 [batch, len], sequence-major), appended after the tokens already cached
 * (history_length = current step, as in the reference's multi-round chat): embedding -> context
 * decoder layers (tcgen05 GEMMs + flash context attention; K/V rows written to the cache) -> final
 * norm + LM head on each sequence's last token -> top-1. On return step += len, the engine's current
 * ids are the first generated tokens (decode_step(NULL) continues), and ids_out_host (may be NULL:
 * then no host sync) receives them. */
int b200_engine_prefill(b200_engine* e, const int* ids_host, int len, int* ids_out_host);

/* Ragged batch (per-sequence lengths: the `step` per request SURVEY 8f-2 asks for instead of the
 * reference's single host scalar, llama.h): sequence b gets lens_host[b] >= 1 prompt tokens,
 * ids_host holds them packed back to back (sum(lens) ids, the padding-free layout the context
 * decoder works on, llama.cpp:158-185), each appended after what THAT sequence has cached. When
 * the lengths end up different, the decode steps that follow run with per-sequence positions and
 * key counts (b200_rope_decode_ragged / b200_decode_attention_ragged, or the GEMV epilogue with
 * b200_gemv_args.step_stride = 1 in the fused order; CUDA-graphed, not the persistent kernel) and every sequence advances by one token per step; b200_engine_get_step() then
 * reports the longest sequence, b200_engine_get_seq_steps() all of them (tokens cached, host int
 * [batch]). b200_engine_reset / fill_kv_random return the engine to one shared step. */
int b200_engine_prefill_ragged(b200_engine* e, const int* ids_host, const int* lens_host,
                               int* ids_out_host);
int b200_engine_get_seq_steps(const b200_engine* e, int* steps_out_host);
int b200_engine_prefill_launches(const b200_engine* e);   /* kernel launches of the last prefill */

/* The generation loop of Llama<T>::Response (llama.cpp:363-448) without the tokenizer: prefill the
 * prompt ([batch, prompt_len] host ids), then decode until every sequence has produced eos_id or
 * max_new_tokens tokens are out. out_ids [batch, max_new_tokens] (host; padded with eos_id),
 * n_generated [batch] (host, may be NULL). Returns the number of decode steps run, < 0 on error. */
int b200_engine_generate(b200_engine* e, const int* prompt_ids_host, int prompt_len, int max_new_tokens,
                         int eos_id, int* out_ids, int* n_generated);
/* the same for prompts of different lengths: ids packed back to back + lens_host[batch]
 * (b200_engine_prefill_ragged), every sequence decoding from its own length */
int b200_engine_generate_ragged(b200_engine* e, const int* prompt_ids_host, const int* lens_host,
                                int max_new_tokens, int eos_id, int* out_ids, int* n_generated);

/* Load the reference's weight directory (LlamaWeight<T>::loadWeights llama_weights.cc:41-53,
 * LlamaLayerWeight<T>::loadWeights layer_weights.cc:48-66): raw fp32 files model.norm.weight.bin,
 * lm_head.weight.bin, model.embed_tokens.weight.bin, model.layers.<l>.{input_layernorm,
 * post_attention_layernorm,self_attn.qkv,self_attn.o_proj,mlp.gate_up_proj,mlp.down_proj}.weight.bin,
 * converted to the engine dtype; with tp_size > 1 each rank keeps its head / FFN-column shard. */
int b200_engine_load_weights(b200_engine* e, const char* dir);

/* device pointers into the engine (valid until destroy): logits [batch, vocab], ids [batch],
 * hidden [batch, H], caches; for parity tests and the bench. */
void* b200_engine_logits(b200_engine* e);
int* b200_engine_ids_out(b200_engine* e);
void* b200_engine_hidden(b200_engine* e);
void* b200_engine_k_cache(b200_engine* e);
void* b200_engine_v_cache(b200_engine* e);
/* device pointer of one weight tensor (bench: per-kernel roofline runs on the real weights) */
void* b200_engine_weight_ptr(b200_engine* e, int slot, int layer);
/* kernel launches one decode step issues (graph nodes or stream launches) */
int b200_engine_launches_per_step(const b200_engine* e);
/* debug (B200_STEP_TRACE=1 at create time): per-op %globaltimer stamps of CTA 0 of the persistent
 * kernel's most recent launch: [2*i] = op i finished, [2*i+1] = barrier after op i passed. Returns
 * the number of entries written (0 when tracing is off or the persistent path is not in use). */
int b200_engine_step_trace(b200_engine* e, unsigned long long* host_dst, int max_entries);

#ifdef __cplusplus
}
#endif
#endif /* B200_ENGINE_H */
