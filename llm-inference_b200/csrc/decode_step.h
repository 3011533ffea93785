// Program of the persistent whole-step decode kernel (decode_step.cu), built by the engine.
#pragma once
#include <cuda_runtime.h>
#include <stddef.h>
#include <vector>

#include "p2p_internal.h"

namespace b200 {

// kOpAttnO: split-KV attention items, then - without a grid barrier - the o projection split along K into (head, part)
// items: the CTA of item (head, part) waits on the (sequence, kv head) counters of the attention items it depends on,
// merges that head's split partials, multiplies the head's d outputs by ITS block of the packed o weights
// (W2: [head_num * parts][d][N / parts], rows = input columns) and adds the partial rows into acc_out (red.add).
// Fields: N = hidden size, K = head_num * head_size, units = head_num * parts, act_stride = parts, y = counters.
// kOpExchange (tensor parallel): the fp32 accumulator of a K-split projection holds THIS RANK's partial sums. CTA c
// rounds its slice to T and pushes it as 16-byte low-latency lines {data, call, data, call} into slot `rank` of every
// rank's receive buffer (p2p_allreduce.cu's wire format, peer stores over NVLink), clears the slice, then spins on the
// lines of all ranks for the same slice, sums them in rank order, adds the residual and stores `hidden` - the
// allreduce after o_proj / down_proj and launchAddResidual in one op. Fields: acc_out = accumulator, y = hidden
// (in place), N = hidden size, units = index of the exchange within the step (call tag = base + units + 1).
enum { kOpGemv = 0, kOpAttn = 1, kOpFinal = 2, kOpAttnO = 3, kOpExchange = 4 };
// kEpiSwigluDown: gate_up GEMV + SwiGLU, then - without a grid barrier - the CTA multiplies the slice of `act` it has
// just produced by ITS rows of the transposed down weights W2 [I, H] and adds the partial [H] into acc_out with
// red.global.add.v4.f32 (the down projection split along K over the CTAs; L2 does the cross-CTA sum)
enum { kEpiPlain = 0, kEpiResidual = 1, kEpiSwiglu = 2, kEpiRopeKV = 3, kEpiSwigluDown = 4 };
// kXResidAcc: x = T(x + T(acc_in)) - the residual add of a K-split projection, done where its result is first needed
enum { kXPlain = 0, kXEmbed = 1, kXCombine = 2, kXResidAcc = 3 };

struct StepOp {
    int type;        // kOp*
    int epi;         // kEpi* (gemv)
    int xsrc;        // kX*   (gemv)
    int norm;        // 0 none, 1 (x*inv)*gamma, 2 (gamma*x)*inv   (gemv prologue)
    int N, K, units;
    int n_zero;      // floats of `zero` cleared (split over the CTAs) when this op starts
    int act_stride;  // kEpiSwigluDown: halves between the batch rows of the CTA's act slice in shared memory; kOpAttnO: parts
    int no_barrier;  // 1: no grid barrier after this op (its consumers wait on finer-grained counters instead)
    int n_zero2;     // floats of `zero2` cleared when this op starts
    int pad2, pad3, pad4;
    const void* W;
    const void* x;
    void* y;
    const void* gamma;
    void* k_cache;   // layer slice [B, hk, max_seq, d] (RoPE/KV epilogue, attention)
    void* v_cache;
    const void* W2;        // kEpiSwigluDown: down weights transposed, [N = I rows, K = H]
    const float* acc_in;   // kXResidAcc: fp32 sums [nb, K] of the preceding K-split projection
    float* acc_out;        // kEpiSwigluDown: fp32 accumulator [nb, K]
    void* xout;            // kXEmbed / kXResidAcc: CTA 0 stores the new residual stream [nb, K] here
    float* zero;           // accumulator to clear for its next use (consumed one op earlier), or nullptr
    float* zero2;          // a second range to clear (the arrival counters), or nullptr
    unsigned* head_cnt;    // kEpiRopeKV: per qkv head, row pairs finished (released by the producing CTAs);
                           // kOpAttnO: the same array - an attention item waits for its q, k and v heads
};
static_assert(sizeof(StepOp) % 16 == 0, "StepOp travels in 16-byte cp.async pieces");

struct StepParams {
    const StepOp* ops;
    int n_ops;
    int nb;               // sequences
    int head_num, kv_head_num, head_size, max_seq_len, rotary_dim;
    float eps;
    const float* rope_den;
    int* step_dev;        // reference `step` of the token being decoded (tokens in cache + 1)
    int* ids_cur;
    int* ids_next;
    const void* embed;
    int vocab, hidden;
    void* qkv;            // [B, h+2hk, d]
    void* attn;           // [B, h*d]  (splits == 1 only)
    void* logits;         // [B, V]
    float* part_ml;       // [B, h, splits, 2]
    float* part_o;        // [B, h, splits, d]
    int splits;
    unsigned* sync;       // [0] arrival counter (monotonic), [1] base of the current launch, [2] error flag
    int stages;
    size_t xs_bytes;
    P2PWire wire;         // tensor parallel: the ranks' receive buffers (wire.world == 0: single GPU)
    int n_exchanges;      // kOpExchange ops per step (the communicator's call counter advances by this per launch)
    unsigned long long* trace;   // optional (B200_STEP_TRACE=1): [0,2n) CTA 0 (op done, barrier passed) %globaltimer stamps, [2n,3n) x staged,
                                 // [3n,4n) ring stages landed at GEMV start, [4n] L2 prefetches, [4n+1,5n+1) / [5n+1,6n+1) cycles warp 0 of the
                                 // first / last CTA waited for weights, [6n+1 + op*grid + cta] op-done stamp of every CTA
};

struct StepProgram {
    StepParams params{};
    StepOp* ops_dev = nullptr;
    unsigned* sync_dev = nullptr;
    void* attn_ws = nullptr;   // split-KV partials (sized by b200_decode_attention_workspace_bytes)
    size_t smem = 0;
    int nb_t = 0, g = 0, grid = 0;
};


// nullptr when the configuration is outside what the persistent kernel covers (fp16, batch <= 4,
// head_size 128, h/hk in {1, 2, 4, 8}, cooperative launch available)
StepProgram* step_program_create(const std::vector<StepOp>& ops, StepParams params, int batch, int head_num,
                                 int kv_head_num, int head_size, int k_max, void* attn_ws);
// `step` = reference step of the token being decoded (tokens already cached + 1)
int step_program_launch(const StepProgram* sp, int step, cudaStream_t s);
void step_program_destroy(StepProgram* sp);
// debug: per-op timestamps of CTA 0 from the most recent launch (enabled by B200_STEP_TRACE=1)
int step_program_trace(const StepProgram* sp, unsigned long long* host_dst, int max_entries);

}  // namespace b200
