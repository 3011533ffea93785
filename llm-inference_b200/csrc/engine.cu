// Decode-step driver (C ABI in include/b200_engine.h).
// Host-side order of operations follows the reference's per-token path —
// llama.cpp:319-361 (continueTokenGen), self_decoder.cpp:53-85 (layer loop),
// masked_self_attention.cpp:55-92, ffn.cpp:53-93, llama.cpp:215-271 (LM head + top-1) —
// with its per-call Malloc/Free + cudaDeviceSynchronize replaced by one arena, one stream and
// CUDA-graph replay (one graph per split-KV geometry; `step` lives in device memory so the same
// graph serves every position of its bucket).
#include <algorithm>
#include <cstdio>
#include <map>
#include <string>
#include <vector>

#include "b200_engine.h"
#include "common.cuh"
#include "decode_step.h"
#include "launch_count.h"

namespace b200 {

static thread_local std::string g_err;
static int fail(const std::string& m, int code = B200_EINVAL) {
    g_err = m;
    return code;
}
#define CU(call)                                                                            \
    do {                                                                                    \
        cudaError_t e_ = (call);                                                            \
        if (e_ != cudaSuccess)                                                              \
            return fail(std::string(#call) + ": " + cudaGetErrorString(e_), (int)e_);       \
    } while (0)
#define RC(call)                                                        \
    do {                                                                \
        int rc_ = (call);                                               \
        if (rc_ != 0) return fail(std::string(#call) + " failed", rc_); \
    } while (0)

int kv_fill_fp8(void* cache8, float* scale, size_t rows_total, int max_seq, int len, int d, uint64_t seed, float amp,
                cudaStream_t s);   // kv_fp8.cu

struct LayerW {
    void *attn_norm, *qkv, *o, *ffn_norm, *gate_up, *down;
};

}  // namespace b200

using namespace b200;

struct b200_engine {
    b200_model_config cfg;
    int batch;
    size_t esz;
    // per-rank shapes
    int h, hk, d, H, I, V, qkv_n;
    // memory
    char* arena = nullptr;
    size_t arena_bytes = 0;
    std::vector<LayerW> layers;
    void *embed = nullptr, *final_norm = nullptr, *lm_head = nullptr;
    char *k_cache = nullptr, *v_cache = nullptr;
    size_t kv_bytes_each = 0, layer_kv_elems = 0;
    // fp8 KV cache (B200_ENGINE_KV_FP8, kv_fp8.cu): one byte per cached element + one fp32 scale per cached row
    bool kv_fp8 = false;
    size_t kv_esz = 0;            // bytes per cached element (esz, or 1)
    float *k_scale = nullptr, *v_scale = nullptr;   // [layers, B, hk, max_seq]
    size_t layer_kv_rows = 0;
    // paged KV cache (B200_ENGINE_KV_PAGED, kv_paged.cu): k_cache / v_cache are [layers][n_pages, hk, 128, d] pools; one
    // page list per sequence, shared by all layers. The host owns the allocator (free stack + pinned mirror of the lists);
    // the device only ever reads the lists.
    bool kv_paged = false;
    int n_pages = 0, max_pages = 0;        // pool size; list length = ceil(max_seq / 128)
    int* page_table = nullptr;             // device [batch, max_pages]
    int* page_table_host = nullptr;        // pinned mirror (entries are only appended while a sequence lives)
    int* page_rows_dev = nullptr;          // device [batch]: rows per sequence, for gather (tests)
    std::vector<int> free_pages, seq_pages;
    std::vector<char> seq_released;        // retired sequences whose pages went back (b200_engine_release_finished)
    bool pages_dirty = false;              // the host lists hold entries the device copy does not have yet
    char* act_arena = nullptr;
    void *hidden, *residual, *qkv, *attn, *act, *gate_up, *logits, *partial;
    int *ids_cur, *ids_next, *step_dev;
    unsigned char* finished = nullptr;   // [batch] device flags (b200_engine_set_eos); part of the activation arena
    int eos_id = -1;                      // -1: no retirement
    float* rope_den;
    void* attn_ws;
    size_t attn_ws_bytes;
    void* lin_ws;          // split-K partials of the tcgen05 weight-streaming linears (batch > 8)
    size_t lin_ws_bytes;
    int* pinned_ids = nullptr;
    // prefill scratch (grown on demand by b200_engine_prefill)
    char* pf_arena = nullptr;
    size_t pf_bytes = 0;
    int* pf_pinned = nullptr;
    size_t pf_pinned_ints = 0;
    cudaEvent_t pf_copied = nullptr;   // recorded after the H2D copies out of pf_pinned
    int prefill_launches = 0;
    // control
    cudaStream_t stream = nullptr;
    unsigned mode = B200_ENGINE_FUSED | B200_ENGINE_GRAPH | B200_ENGINE_PERSISTENT;
    b200_allreduce_fn allreduce = nullptr;
    void* allreduce_user = nullptr;
    b200_p2p* p2p = nullptr;   // fused one-shot NVLink allreduce + residual add (p2p_allreduce.cu)
    int step = 0;  // host mirror of *step_dev: tokens already in the cache (ragged batch: of the longest sequence)
    // ragged batch (per-sequence lengths): seq_steps[b] = reference `step` of sequence b (device), seq_cached its host
    // mirror minus one. Uniform batches keep every entry equal to `step` and never read the device array.
    int* seq_steps = nullptr;
    std::vector<int> seq_cached;
    bool ragged = false;
    std::map<int, cudaGraphExec_t> graphs;  // key: split-KV count
    b200::StepProgram* program = nullptr;   // persistent whole-step kernel (fp16, batch <= 4, tp 1)
    // fused FFN of the persistent kernel: down weights stored TRANSPOSED [I, H] per layer (the K-split form streams
    // rows of it), refreshed lazily whenever a down weight changes; two fp32 accumulators [batch, H] (layer parity)
    char* down_t = nullptr;
    float* ffn_acc = nullptr;
    bool down_t_dirty = true;
    // attention -> o projection without a grid barrier (kOpAttnO): o weights packed per (head, part) item, one fp32
    // accumulator [batch, H] and the (sequence, kv head) arrival counters behind it
    char* o_packed = nullptr;
    float* o_acc = nullptr;     // [batch*H floats][counters: batch*hk unsigned, padded to 4]
    int o_parts = 0;
    int launches_per_step = 0;
};

namespace b200 {

static size_t align256(size_t x) { return (x + 255) & ~(size_t)255; }

static size_t slot_elems(const b200_engine* e, int slot) {
    switch (slot) {
        case B200_W_ATTN_NORM: case B200_W_FFN_NORM: case B200_W_FINAL_NORM: return (size_t)e->H;
        case B200_W_QKV: return (size_t)e->qkv_n * e->H;
        case B200_W_O: return (size_t)e->H * e->h * e->d;
        case B200_W_GATE_UP: return (size_t)2 * e->I * e->H;
        case B200_W_DOWN: return (size_t)e->H * e->I;
        case B200_W_EMBED: case B200_W_LM_HEAD: return (size_t)e->V * e->H;
    }
    return 0;
}

static void* slot_ptr(b200_engine* e, int slot, int layer) {
    if (slot <= B200_W_DOWN) {
        if (layer < 0 || layer >= (int)e->layers.size()) return nullptr;
        LayerW& w = e->layers[layer];
        switch (slot) {
            case B200_W_ATTN_NORM: return w.attn_norm;
            case B200_W_QKV: return w.qkv;
            case B200_W_O: return w.o;
            case B200_W_FFN_NORM: return w.ffn_norm;
            case B200_W_GATE_UP: return w.gate_up;
            case B200_W_DOWN: return w.down;
        }
    }
    if (slot == B200_W_EMBED) return e->embed;
    if (slot == B200_W_FINAL_NORM) return e->final_norm;
    if (slot == B200_W_LM_HEAD) return e->lm_head;
    return nullptr;
}

// counter-based normal / uniform fill (splitmix64 hash -> Box-Muller); value = mean + std*n
template <typename T>
__global__ void fill_normal_kernel(T* p, size_t n, uint64_t seed, float mean, float std) {
    for (size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x; i < n;
         i += (size_t)gridDim.x * blockDim.x) {
        uint64_t z = seed + 0x9E3779B97F4A7C15ull * (i + 1);
        z = (z ^ (z >> 30)) * 0xBF58476D1CE4E5B9ull;
        z = (z ^ (z >> 27)) * 0x94D049BB133111EBull;
        z ^= z >> 31;
        const float u1 = ((uint32_t)(z >> 40) + 1.0f) * (1.0f / 16777217.0f);
        const float u2 = (uint32_t)((z >> 8) & 0xffffff) * (1.0f / 16777216.0f);
        const float nrm = sqrtf(-2.f * logf(u1)) * cospif(2.f * u2);
        p[i] = DT<T>::from_f(mean + std * nrm);
    }
}
template <typename T>
__global__ void fill_uniform_rows_kernel(T* cache, size_t rows_total, int max_seq, int len, int d,
                                         uint64_t seed, float amp) {
    // cache viewed as [rows_total = layers*B*hk][max_seq][d]; fill rows t < len
    const size_t n = rows_total * (size_t)len * d;
    for (size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x; i < n;
         i += (size_t)gridDim.x * blockDim.x) {
        const size_t r = i / ((size_t)len * d);
        const size_t rem = i - r * (size_t)len * d;
        uint64_t z = seed + 0x9E3779B97F4A7C15ull * (i + 1);
        z = (z ^ (z >> 30)) * 0xBF58476D1CE4E5B9ull;
        z = (z ^ (z >> 27)) * 0x94D049BB133111EBull;
        z ^= z >> 31;
        const float u = (uint32_t)(z >> 40) * (1.0f / 16777216.0f);
        cache[r * (size_t)max_seq * d + rem] = DT<T>::from_f(amp * (2.f * u - 1.f));
    }
}

// dst [cols, rows] = src [rows, cols]^T, 32 x 32 tiles through shared memory (load-time weight packing)
template <typename T>
__global__ void transpose_kernel(const T* __restrict__ src, T* __restrict__ dst, int rows, int cols) {
    __shared__ T tile[32][33];
    const int c0 = blockIdx.x * 32, r0 = blockIdx.y * 32;
    for (int j = threadIdx.y; j < 32; j += blockDim.y) {
        const int r = r0 + j, c = c0 + threadIdx.x;
        if (r < rows && c < cols) tile[j][threadIdx.x] = src[(size_t)r * cols + c];
    }
    __syncthreads();
    for (int j = threadIdx.y; j < 32; j += blockDim.y) {
        const int c = c0 + j, r = r0 + threadIdx.x;
        if (r < rows && c < cols) dst[(size_t)c * rows + r] = tile[threadIdx.x][j];
    }
}

// o weights [H, h*d] -> blocks of item (head, part): [d input columns][H / parts output rows] (load-time packing)
__global__ void pack_o_kernel(const __half* __restrict__ W, __half* __restrict__ dst, int H, int heads, int d, int parts) {
    const int hp = H / parts;
    const size_t total = (size_t)H * heads * d;
    for (size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x; i < total; i += (size_t)gridDim.x * blockDim.x) {
        const int r = (int)(i % hp);
        size_t t = i / hp;
        const int j = (int)(t % d);
        t /= d;
        const int part = (int)(t % parts), head = (int)(t / parts);
        dst[i] = W[(size_t)(part * hp + r) * ((size_t)heads * d) + (size_t)head * d + j];
    }
}

// `finished` retirement (sampling.cu:46,81-83: finished[b] = (id == end_id)): once a sequence has produced eos_id it
// keeps emitting eos_id, and the attention / append kernels skip it (b200_decode_attention_set_finished)
__device__ __forceinline__ int retire(int* ids_next, unsigned char* finished, int eos_id, int b) {
    int id = ids_next[b];
    if (finished != nullptr && eos_id >= 0) {
        if (finished[b]) { id = eos_id; ids_next[b] = id; }
        else if (id == eos_id) finished[b] = 1;
    }
    return id;
}

__global__ void advance_kernel(int* ids_cur, int* ids_next, int* step, int batch, unsigned char* finished, int eos_id) {
    pdl_wait();
    for (int b = threadIdx.x; b < batch; b += blockDim.x) ids_cur[b] = retire(ids_next, finished, eos_id, b);
    if (threadIdx.x == 0) *step += 1;
}

// ragged batch: every sequence moves on by one token from its own length
__global__ void advance_ragged_kernel(int* ids_cur, int* ids_next, int* step, int* seq_steps, int batch,
                                      unsigned char* finished, int eos_id) {
    pdl_wait();
    for (int b = threadIdx.x; b < batch; b += blockDim.x) {
        const bool was = finished != nullptr && eos_id >= 0 && finished[b];
        ids_cur[b] = retire(ids_next, finished, eos_id, b);
        if (!was) seq_steps[b] += 1;   // a retired sequence stays where it ended
    }
    if (threadIdx.x == 0) *step += 1;
}

}  // namespace b200

// The program of the persistent whole-step kernel: the same op order as enqueue_step()'s fused
// branch, one entry per GEMV / attention phase.
static void build_program(b200_engine* e) {
    const b200_model_config& c = e->cfg;
    // tensor parallel: only with the peer-memory communicator attached (b200_engine_set_p2p rebuilds the program)
    // (a communicator attached to a tp_size = 1 engine gives the same exchange form with one rank: one-GPU test coverage)
    if (c.tp_size > 1 && !(e->p2p && e->p2p->connected)) return;
    const bool tp = e->p2p != nullptr && e->p2p->connected;
    if (c.dtype != B200_F16 || e->batch > 4) return;
    std::vector<StepOp> ops;
    const int H = e->H, L = c.num_layers;
    // Fused FFN (gate_up + SwiGLU + K-split down in one op): needs whole W2 rows per ring stage
    // (H a power of two <= 16384) and the per-thread accumulators in registers (batch tile x packets per thread <= 8)
    const int nb_t = e->batch <= 1 ? 1 : e->batch <= 2 ? 2 : 4;
    const int ppt = (H / 8 + 255) / 256;
    bool fuse_ffn = L >= 1 && H >= 256 && (H & (H - 1)) == 0 && ppt <= 4 && nb_t * ppt <= 8;   // H <= 8192; 8192: batch <= 2
    {
        const char* env = getenv("B200_STEP_NO_FFN_FUSION");
        if (env && env[0] == '1' && !tp) fuse_ffn = false;
    }
    if (tp && !fuse_ffn) return;   // the tensor-parallel program exists only in the K-split form (its accumulators ARE the partials)
    if (fuse_ffn) {
        const size_t wt = (size_t)L * e->I * H * e->esz, ab = (size_t)2 * e->batch * H * sizeof(float);
        if (cudaMalloc(&e->down_t, wt) != cudaSuccess || cudaMalloc(&e->ffn_acc, ab) != cudaSuccess) {
            cudaGetLastError();
            cudaFree(e->down_t);
            e->down_t = nullptr;
            fuse_ffn = false;
        } else {
            cudaMemset(e->ffn_acc, 0, ab);
            e->down_t_dirty = true;
        }
    }
    // attention -> o projection as one op (no grid barrier in between): parts = the largest power of two with
    // head_num * parts <= CTAs and >= 64 output rows per part; needs the fused FFN's residual plumbing
    int parts = 0;
    if (fuse_ffn && e->d == 128) {
        parts = 1;
        while (e->h * parts * 2 <= sm_count() && H / (parts * 2) >= 64) parts *= 2;
        const int row_len = H / parts;
        const char* env = getenv("B200_STEP_NO_ATTN_O_FUSION");
        if ((env && env[0] == '1') || H % parts || row_len % 8 || (32768 % (row_len * 2)) || (row_len / 8 + 255) / 256 > 2 ||
            nb_t * ((row_len / 8 + 255) / 256) > 8)
            parts = 0;
    }
    if (tp && !parts) {
        cudaFree(e->down_t); cudaFree(e->ffn_acc);
        e->down_t = nullptr; e->ffn_acc = nullptr;
        return;
    }
    if (parts) {
        const size_t wb = (size_t)L * H * e->h * e->d * e->esz;
        // arrival counters: [batch * hk attention items done][h + 2 hk qkv heads done], each block padded to 4
        const size_t cnt = (((size_t)e->batch * e->hk + 3) & ~(size_t)3) + (((size_t)e->h + 2 * e->hk + 3) & ~(size_t)3);
        const size_t ab = (size_t)e->batch * H * sizeof(float) + cnt * sizeof(unsigned);
        if (cudaMalloc(&e->o_packed, wb) != cudaSuccess || cudaMalloc(&e->o_acc, ab) != cudaSuccess) {
            cudaGetLastError();
            cudaFree(e->o_packed);
            e->o_packed = nullptr;
            parts = 0;
        } else {
            cudaMemset(e->o_acc, 0, ab);
            e->o_parts = parts;
        }
    }
    // Residual stream with K-split projections: the residual add of a projection is formed in the NEXT op's x staging
    // (T(x + T(sum))), which publishes the result to the OTHER buffer - CTAs of the same staging still read the old one.
    //   FFN only fused : layer l works in buf[l % 2] (o_proj adds in place), the decoder output lands in e->hidden
    //   attn-o fused too: A = e->hidden holds h_l, B = e->residual holds h_l + o_proj; A again after the FFN
    void* buf[2] = {e->hidden, e->hidden};
    if (fuse_ffn && !parts) {
        buf[L % 2] = e->hidden;
        buf[(L + 1) % 2] = e->residual;
    }
    void* const bufA = e->hidden;
    void* const bufB = e->residual;
    auto acc = [&](int l) { return e->ffn_acc + (size_t)(l % 2) * e->batch * H; };
    const int n_acc = e->batch * H;
    unsigned* counters = parts ? reinterpret_cast<unsigned*>(e->o_acc + n_acc) : nullptr;
    const int n_cnt_attn = (e->batch * e->hk + 3) & ~3;
    const int n_cnt = n_cnt_attn + ((e->h + 2 * e->hk + 3) & ~3);
    unsigned* head_cnt = parts ? counters + n_cnt_attn : nullptr;
    {
        const char* env = getenv("B200_STEP_QKV_BARRIER");   // 1: keep the grid barrier between qkv and attention
        if (env && env[0] == '1') head_cnt = nullptr;
    }
    for (int l = 0; l < L; ++l) {
        LayerW& w = e->layers[l];
        void* kc = e->k_cache + (size_t)l * e->layer_kv_elems * e->esz;
        void* vc = e->v_cache + (size_t)l * e->layer_kv_elems * e->esz;
        void* hid = parts ? bufA : buf[l % 2];   // (tensor parallel: always e->hidden = bufA)
        StepOp qkv{};
        qkv.type = kOpGemv; qkv.epi = kEpiRopeKV; qkv.norm = 1; qkv.N = e->qkv_n; qkv.K = H; qkv.units = e->qkv_n / 2;
        qkv.W = w.qkv; qkv.y = e->qkv; qkv.gamma = w.attn_norm; qkv.k_cache = kc; qkv.v_cache = vc;
        if (l == 0) { qkv.xsrc = kXEmbed; qkv.xout = hid; qkv.x = hid; }
        else if (tp) { qkv.xsrc = kXPlain; qkv.x = hid; }   // the exchange op has already formed the residual sum in `hidden`
        else if (fuse_ffn) { qkv.xsrc = kXResidAcc; qkv.x = parts ? bufB : buf[(l - 1) % 2]; qkv.acc_in = acc(l - 1); qkv.xout = hid; }
        else { qkv.xsrc = kXPlain; qkv.x = hid; }
        if (parts && l >= 1 && !tp) { qkv.zero = e->o_acc; qkv.n_zero = n_acc; }   // consumed by the previous layer's FFN staging
        StepOp att{};
        att.type = kOpAttn; att.k_cache = kc; att.v_cache = vc;
        // acc(l - 1) was consumed by this layer's qkv staging: cleared one barrier later - at the attention op, or at
        // the FFN op when no grid barrier separates qkv and attention
        if (fuse_ffn && l >= 1 && !head_cnt && !tp) { att.zero = acc(l - 1); att.n_zero = n_acc; }
        StepOp gu{};
        gu.type = kOpGemv; gu.xsrc = kXPlain; gu.norm = 2; gu.N = e->I; gu.K = H; gu.units = e->I;
        gu.W = w.gate_up; gu.x = hid; gu.gamma = w.ffn_norm;
        ops.push_back(qkv);
        if (parts) {
            att.type = kOpAttnO;
            att.N = H; att.K = e->h * e->d; att.units = e->h * parts; att.act_stride = parts;
            att.W2 = e->o_packed + (size_t)l * H * e->h * e->d * e->esz;
            att.acc_out = e->o_acc;
            att.y = counters;
            ops.push_back(att);
            gu.xsrc = kXResidAcc; gu.x = bufA; gu.acc_in = e->o_acc; gu.xout = bufB;
            gu.zero2 = reinterpret_cast<float*>(counters); gu.n_zero2 = n_cnt;   // all waits on them have been passed
            if (head_cnt) {
                ops[ops.size() - 2].head_cnt = head_cnt;   // the qkv op pushed above releases, ...
                ops[ops.size() - 2].no_barrier = 1;
                ops.back().head_cnt = head_cnt;            // ... the kOpAttnO op's attention items wait
                if (l >= 1 && !tp) { gu.zero = acc(l - 1); gu.n_zero = n_acc; }
            }
            if (tp) {
                // tensor parallel: o_acc holds this rank's partial; the exchange op sums it over the ranks, adds the
                // residual into `hidden` and clears it; the FFN then reads `hidden` like any other op
                StepOp ex{};
                ex.type = kOpExchange; ex.N = H; ex.units = 2 * l; ex.acc_out = e->o_acc; ex.y = e->hidden;
                ops.push_back(ex);
                gu.xsrc = kXPlain; gu.x = e->hidden; gu.acc_in = nullptr; gu.xout = nullptr;
            }
        } else {
            StepOp o{};
            o.type = kOpGemv; o.epi = kEpiResidual; o.xsrc = kXCombine; o.N = H; o.K = e->h * e->d; o.units = (H + 1) / 2;
            o.W = w.o; o.x = e->attn; o.y = hid;
            ops.push_back(att); ops.push_back(o);
        }
        if (fuse_ffn) {
            gu.epi = kEpiSwigluDown;
            gu.act_stride = (((e->I + sm_count() - 1) / sm_count()) + 8) & ~7;
            gu.W2 = e->down_t + (size_t)l * e->I * H * e->esz;
            gu.acc_out = tp ? acc(0) : acc(l);
            ops.push_back(gu);
            if (tp) {
                StepOp ex{};
                ex.type = kOpExchange; ex.N = H; ex.units = 2 * l + 1; ex.acc_out = acc(0); ex.y = e->hidden;
                ops.push_back(ex);
            }
        } else {
            gu.epi = kEpiSwiglu;
            gu.y = e->act;
            StepOp dn{};
            dn.type = kOpGemv; dn.epi = kEpiResidual; dn.xsrc = kXPlain; dn.N = H; dn.K = e->I; dn.units = (H + 1) / 2;
            dn.W = w.down; dn.x = e->act; dn.y = hid;
            ops.push_back(gu); ops.push_back(dn);
        }
    }
    StepOp head{};
    head.type = kOpGemv; head.epi = kEpiPlain; head.norm = 1; head.N = e->V; head.K = H; head.units = (e->V + 1) / 2;
    head.W = e->lm_head; head.y = e->logits; head.gamma = e->final_norm;
    if (L == 0) { head.xsrc = kXEmbed; head.x = e->hidden; head.xout = e->hidden; }
    else if (tp) { head.xsrc = kXPlain; head.x = e->hidden; }
    else if (fuse_ffn) {
        head.xsrc = kXResidAcc; head.x = parts ? bufB : buf[(L - 1) % 2]; head.acc_in = acc(L - 1);
        head.xout = parts ? bufA : buf[L % 2];
        if (parts) { head.zero = e->o_acc; head.n_zero = n_acc; }
    }
    else { head.xsrc = kXPlain; head.x = e->hidden; }
    StepOp fin{};
    fin.type = kOpFinal;
    if (fuse_ffn && !tp) { fin.zero = acc(L - 1); fin.n_zero = n_acc; }
    ops.push_back(head); ops.push_back(fin);
    StepParams p{};
    p.head_num = e->h; p.kv_head_num = e->hk; p.head_size = e->d; p.max_seq_len = c.max_seq_len;
    p.rotary_dim = c.rotary_dim; p.eps = c.rms_eps; p.rope_den = e->rope_den;
    p.step_dev = e->step_dev; p.ids_cur = e->ids_cur; p.ids_next = e->ids_next;
    p.embed = e->embed; p.vocab = e->V; p.hidden = H;
    p.qkv = e->qkv; p.attn = e->attn; p.logits = e->logits;
    if (tp) {
        p.wire = p2p_wire(e->p2p);
        p.n_exchanges = 2 * L;
    }
    const int k_max = fuse_ffn ? std::max(H, e->h * e->d) : std::max(std::max(H, e->I), e->h * e->d);
    e->program = step_program_create(ops, p, e->batch, e->h, e->hk, e->d, k_max, e->attn_ws);
    if (!e->program && e->down_t) {
        cudaFree(e->down_t);
        cudaFree(e->ffn_acc);
        cudaFree(e->o_packed);
        cudaFree(e->o_acc);
        e->down_t = nullptr;
        e->ffn_acc = nullptr;
        e->o_packed = nullptr;
        e->o_acc = nullptr;
    }
}

// the persistent kernel's transposed copy of the down weights follows every change of them
static int refresh_down_t(b200_engine* e) {
    if (!e->down_t || !e->down_t_dirty) return 0;
    const dim3 blk(32, 8), grd((e->I + 31) / 32, (e->H + 31) / 32);
    for (int l = 0; l < e->cfg.num_layers; ++l)
        transpose_kernel<__half><<<grd, blk, 0, e->stream>>>((const __half*)e->layers[l].down,
                                                             (__half*)e->down_t + (size_t)l * e->I * e->H, e->H, e->I);
    if (e->o_packed)
        for (int l = 0; l < e->cfg.num_layers; ++l)
            pack_o_kernel<<<148 * 8, 256, 0, e->stream>>>((const __half*)e->layers[l].o,
                                                          (__half*)e->o_packed + (size_t)l * e->H * e->h * e->d, e->H,
                                                          e->h, e->d, e->o_parts);
    CU(cudaGetLastError());
    e->down_t_dirty = false;
    return 0;
}

extern "C" const char* b200_engine_last_error(void) { return g_err.c_str(); }

extern "C" b200_engine* b200_engine_create(const b200_model_config* c, int batch) {
    return b200_engine_create_ex(c, batch, 0u);
}

extern "C" b200_engine* b200_engine_create_ex(const b200_model_config* c, int batch, unsigned flags) {
    return b200_engine_create_paged(c, batch, flags, 0);
}

/* n_pages: size of each layer's K (V) page pool with B200_ENGINE_KV_PAGED (0 = batch * ceil(max_seq_len / 128), the
 * capacity of the dense cache); ignored without the flag */
extern "C" b200_engine* b200_engine_create_paged(const b200_model_config* c, int batch, unsigned flags, int n_pages) {
    if (!c || batch < 1 || batch > c->max_batch) { fail("bad config / batch"); return nullptr; }
    if (flags & B200_ENGINE_KV_PAGED) {
        const int G = c->kv_head_num > 0 ? c->head_num / c->kv_head_num : 0;
        if (c->dtype != B200_F16 || c->head_size != 128 || (G != 1 && G != 2 && G != 4 && G != 8)) {
            fail("B200_ENGINE_KV_PAGED needs fp16, head size 128 and head_num / kv_head_num in {1, 2, 4, 8} ({1, 2, 4} with KV_FP8)");
            return nullptr;
        }
    }
    if (flags & B200_ENGINE_KV_FP8) {
        const int G = c->kv_head_num > 0 ? c->head_num / c->kv_head_num : 0;
        if (c->dtype != B200_F16 || c->head_size != 128 || (G != 1 && G != 2 && G != 4)) {
            fail("B200_ENGINE_KV_FP8 needs fp16, head size 128 and head_num / kv_head_num in {1, 2, 4}");
            return nullptr;
        }
    }
    if (c->tp_size < 1 || c->head_num % c->tp_size || c->kv_head_num % c->tp_size ||
        c->inter_size % c->tp_size || c->head_num % c->kv_head_num) {
        fail("heads / inter size not divisible by tp_size");
        return nullptr;
    }
    if (c->dtype != B200_F32 && c->dtype != B200_F16) { fail("dtype"); return nullptr; }
    if (batch > 1024) { fail("batch > 1024"); return nullptr; }
    b200_engine* e = new b200_engine();
    e->cfg = *c;
    e->batch = batch;
    e->esz = c->dtype == B200_F16 ? 2 : 4;
    e->h = c->head_num / c->tp_size;
    e->hk = c->kv_head_num / c->tp_size;
    e->d = c->head_size;
    e->H = c->head_num * c->head_size;
    e->I = c->inter_size / c->tp_size;
    e->V = c->vocab_size;
    e->qkv_n = (e->h + 2 * e->hk) * e->d;
    const size_t es = e->esz;
    auto fail_free = [&](const char* what, cudaError_t err) {
        fail(std::string(what) + ": " + cudaGetErrorString(err), (int)err);
        b200_engine_destroy(e);
        return (b200_engine*)nullptr;
    };
    // ---- weights arena
    size_t off = 0;
    std::vector<size_t> offs;
    auto take = [&](size_t elems) { size_t o = off; off += align256(elems * es); return o; };
    struct LO { size_t an, qkv, o, fn, gu, dn; };
    std::vector<LO> lo(c->num_layers);
    for (int l = 0; l < c->num_layers; ++l) {
        lo[l].an = take(slot_elems(e, B200_W_ATTN_NORM));
        lo[l].qkv = take(slot_elems(e, B200_W_QKV));
        lo[l].o = take(slot_elems(e, B200_W_O));
        lo[l].fn = take(slot_elems(e, B200_W_FFN_NORM));
        lo[l].gu = take(slot_elems(e, B200_W_GATE_UP));
        lo[l].dn = take(slot_elems(e, B200_W_DOWN));
    }
    const size_t o_embed = take(slot_elems(e, B200_W_EMBED));
    const size_t o_fnorm = take(slot_elems(e, B200_W_FINAL_NORM));
    const size_t o_head = take(slot_elems(e, B200_W_LM_HEAD));
    e->arena_bytes = off;
    cudaError_t err = cudaMalloc(&e->arena, off);
    if (err != cudaSuccess) return fail_free("cudaMalloc(weights)", err);
    e->layers.resize(c->num_layers);
    for (int l = 0; l < c->num_layers; ++l)
        e->layers[l] = {e->arena + lo[l].an, e->arena + lo[l].qkv, e->arena + lo[l].o,
                        e->arena + lo[l].fn, e->arena + lo[l].gu, e->arena + lo[l].dn};
    e->embed = e->arena + o_embed;
    e->final_norm = e->arena + o_fnorm;
    e->lm_head = e->arena + o_head;
    // ---- KV cache [layers, B, hk, max_seq, d] x2, 64-bit sizes (reference D21)
    e->kv_fp8 = (flags & B200_ENGINE_KV_FP8) != 0;
    e->kv_esz = e->kv_fp8 ? 1 : es;
    e->layer_kv_elems = (size_t)batch * e->hk * c->max_seq_len * e->d;
    e->layer_kv_rows = (size_t)batch * e->hk * c->max_seq_len;
    e->kv_paged = (flags & B200_ENGINE_KV_PAGED) != 0;
    if (e->kv_paged) {
        e->max_pages = (c->max_seq_len + B200_KV_PAGE_ROWS - 1) / B200_KV_PAGE_ROWS;
        if (n_pages < 0) { fail("n_pages < 0"); b200_engine_destroy(e); return nullptr; }
        e->n_pages = n_pages > 0 ? n_pages : batch * e->max_pages;
        e->layer_kv_rows = (size_t)e->n_pages * e->hk * B200_KV_PAGE_ROWS;
        e->layer_kv_elems = e->layer_kv_rows * e->d;
        const size_t tb = (size_t)batch * e->max_pages * sizeof(int);
        if ((err = cudaMalloc(&e->page_table, tb + (size_t)batch * sizeof(int))) != cudaSuccess) return fail_free("cudaMalloc(page lists)", err);
        cudaMemset(e->page_table, 0, tb + (size_t)batch * sizeof(int));
        e->page_rows_dev = e->page_table + (size_t)batch * e->max_pages;
        if ((err = cudaMallocHost(&e->page_table_host, tb + (size_t)batch * sizeof(int))) != cudaSuccess) return fail_free("cudaMallocHost(page lists)", err);
        memset(e->page_table_host, 0, tb + (size_t)batch * sizeof(int));
        e->seq_pages.assign(batch, 0);
        e->seq_released.assign(batch, 0);
        e->free_pages.resize(e->n_pages);
        for (int i = 0; i < e->n_pages; ++i) e->free_pages[i] = e->n_pages - 1 - i;   // page 0 is handed out first
    }
    e->kv_bytes_each = e->layer_kv_elems * c->num_layers * e->kv_esz;
    if (e->kv_fp8) {
        const size_t sb = e->layer_kv_rows * c->num_layers * sizeof(float);
        if ((err = cudaMalloc(&e->k_scale, sb)) != cudaSuccess) return fail_free("cudaMalloc(K scales)", err);
        if ((err = cudaMalloc(&e->v_scale, sb)) != cudaSuccess) return fail_free("cudaMalloc(V scales)", err);
        cudaMemset(e->k_scale, 0, sb);
        cudaMemset(e->v_scale, 0, sb);
    }
    if ((err = cudaMalloc(&e->k_cache, e->kv_bytes_each)) != cudaSuccess) return fail_free("cudaMalloc(K)", err);
    if ((err = cudaMalloc(&e->v_cache, e->kv_bytes_each)) != cudaSuccess) return fail_free("cudaMalloc(V)", err);
    cudaMemset(e->k_cache, 0, e->kv_bytes_each);
    cudaMemset(e->v_cache, 0, e->kv_bytes_each);
    // ---- activations
    e->attn_ws_bytes = b200_decode_attention_workspace_bytes(batch, e->h, e->d, c->max_seq_len);
    off = 0;
    const size_t B = batch;
    const size_t o_hidden = take(B * e->H), o_resid = take(B * e->H), o_qkv = take(B * e->qkv_n),
                 o_attn = take(B * e->h * e->d), o_act = take(B * e->I),
                 o_gu = take(B * 2 * e->I), o_logits = take(B * e->V), o_part = take(B * e->H);
    const size_t o_ids = off; off += align256(3 * B * sizeof(int) + sizeof(int) + 64);
    const size_t o_fin = off; off += align256(B);
    const size_t o_den = off; off += align256((size_t)c->rotary_dim / 2 * sizeof(float) + 64);
    const size_t o_ws = off; off += align256(e->attn_ws_bytes);
    e->lin_ws_bytes = 0;
    {
        const int shapes[5][2] = {{e->qkv_n, e->H}, {e->H, e->h * e->d}, {2 * e->I, e->H}, {e->H, e->I}, {e->V, e->H}};
        for (auto& sh : shapes) {
            const size_t w = b200_linear_workspace_bytes(batch, sh[0], sh[1], 1, c->dtype);
            if (w > e->lin_ws_bytes) e->lin_ws_bytes = w;
        }
    }
    const size_t o_lws = off; off += align256(e->lin_ws_bytes + 256);
    if ((err = cudaMalloc(&e->act_arena, off)) != cudaSuccess) return fail_free("cudaMalloc(act)", err);
    cudaMemset(e->act_arena, 0, off);
    char* a = e->act_arena;
    e->hidden = a + o_hidden; e->residual = a + o_resid; e->qkv = a + o_qkv; e->attn = a + o_attn;
    e->act = a + o_act; e->gate_up = a + o_gu; e->logits = a + o_logits; e->partial = a + o_part;
    e->ids_cur = (int*)(a + o_ids); e->ids_next = e->ids_cur + B; e->step_dev = e->ids_next + B;
    e->seq_steps = e->step_dev + 1;
    e->finished = (unsigned char*)(a + o_fin);
    e->seq_cached.assign(B, 0);
    e->attn_ws = a + o_ws;
    e->lin_ws = a + o_lws;
    e->rope_den = (float*)(a + o_den);
    b200_rope_table(e->rope_den, c->rotary_dim, c->rotary_base, nullptr);
    if ((err = cudaMallocHost(&e->pinned_ids, 2 * B * sizeof(int))) != cudaSuccess) return fail_free("cudaMallocHost", err);
    if ((err = cudaStreamCreateWithFlags(&e->stream, cudaStreamNonBlocking)) != cudaSuccess) return fail_free("stream", err);
    const int ref_step0 = 1;  // reference `step` of the first token (tokens in cache + 1)
    cudaMemcpy(e->step_dev, &ref_step0, sizeof(int), cudaMemcpyHostToDevice);
    cudaDeviceSynchronize();
    if (!e->kv_fp8 && !e->kv_paged) build_program(e);   // the persistent step kernel reads and appends a dense fp16 cache
    // default launch order by batch (measured, tools/order_sweep.sh + route_sweep.sh, 7B fp16, ms per step at
    // context 512 / 4096; GEMV = the tensor-core consumer, gemv.cu):
    //   batch 2: persistent 3.53 / 4.72, fused GEMV graph 3.30 / 4.24, reference launch order graph 3.49 / 4.42
    //   batch 3: persistent 4.58 / 6.91, fused GEMV graph 3.69 / 4.91, reference launch order graph 3.82 / 5.03
    //   batch 4: persistent 4.83 / 7.16, fused GEMV graph 3.90 / 5.53, reference launch order graph 3.93 / 5.53
    // (the persistent kernel still runs the scalar FHFMA loop, issue-bound from 2 token rows on). So batch 1 keeps the
    // persistent kernel, fp16 batches 2-4 default to the graphed fused-GEMV order; batches >= 5 run the reference
    // launch order on the tcgen05 weight-streaming GEMM whatever the FUSED flag says (enqueue_step).
    // Round 2 (2 grid barriers per layer, K-split o / down, 4 KB copies): batch 2 persistent 3.13 / 4.37 ms vs fused GEMV
    // graph 3.31 / 4.38; batch 3 persistent 4.01 / 6.51 vs 3.71 / 5.08 -> the persistent kernel is the default up to batch 2.
    if (c->dtype == B200_F16 && c->tp_size == 1 && batch >= 3) e->mode = B200_ENGINE_FUSED | B200_ENGINE_GRAPH;
    return e;
}

extern "C" void b200_engine_destroy(b200_engine* e) {
    if (!e) return;
    for (auto& kv : e->graphs) cudaGraphExecDestroy(kv.second);
    b200::step_program_destroy(e->program);
    if (e->stream) cudaStreamDestroy(e->stream);
    if (e->pinned_ids) cudaFreeHost(e->pinned_ids);
    if (e->pf_pinned) cudaFreeHost(e->pf_pinned);
    if (e->pf_copied) cudaEventDestroy(e->pf_copied);
    cudaFree(e->pf_arena);
    cudaFree(e->down_t);
    cudaFree(e->ffn_acc);
    cudaFree(e->o_packed);
    cudaFree(e->o_acc);
    cudaFree(e->arena);
    cudaFree(e->k_cache);
    cudaFree(e->v_cache);
    cudaFree(e->k_scale);
    cudaFree(e->v_scale);
    cudaFree(e->page_table);
    if (e->page_table_host) cudaFreeHost(e->page_table_host);
    cudaFree(e->act_arena);
    delete e;
}

// ------------------------------------------------------------------------------------------
// accessors / weights
// ------------------------------------------------------------------------------------------
extern "C" size_t b200_engine_weight_bytes(const b200_engine* e) { return e ? e->arena_bytes : 0; }
extern "C" size_t b200_engine_kv_bytes(const b200_engine* e) {
    return e ? 2 * e->kv_bytes_each + (e->kv_fp8 ? 2 * e->layer_kv_rows * e->cfg.num_layers * sizeof(float) : 0) : 0;
}
extern "C" size_t b200_engine_weight_elems(const b200_engine* e, int slot) { return e ? slot_elems(e, slot) : 0; }
extern "C" void* b200_engine_stream(b200_engine* e) { return e ? (void*)e->stream : nullptr; }
extern "C" void* b200_engine_logits(b200_engine* e) { return e->logits; }
extern "C" int* b200_engine_ids_out(b200_engine* e) { return e->ids_next; }
extern "C" void* b200_engine_hidden(b200_engine* e) { return e->hidden; }
extern "C" void* b200_engine_k_cache(b200_engine* e) { return e->k_cache; }
extern "C" void* b200_engine_v_cache(b200_engine* e) { return e->v_cache; }
extern "C" int b200_engine_get_step(const b200_engine* e) { return e->step; }
extern "C" int b200_engine_kv_is_fp8(const b200_engine* e) { return e && e->kv_fp8 ? 1 : 0; }
extern "C" int b200_engine_kv_is_paged(const b200_engine* e) { return e && e->kv_paged ? 1 : 0; }
extern "C" int b200_engine_kv_pages_total(const b200_engine* e) { return e && e->kv_paged ? e->n_pages : 0; }
extern "C" int b200_engine_kv_pages_free(const b200_engine* e) { return e && e->kv_paged ? (int)e->free_pages.size() : 0; }
extern "C" const int* b200_engine_kv_page_table(const b200_engine* e) { return e && e->kv_paged ? e->page_table : nullptr; }
/* paged cache only: K and V rows of every sequence gathered into dense [layers, B, hk, max_seq, d] DEVICE buffers */
extern "C" int b200_engine_kv_gather(b200_engine* e, void* k_dst, void* v_dst) {
    if (!e || !e->kv_paged || !k_dst || !v_dst) return B200_EINVAL;
    CU(cudaStreamSynchronize(e->stream));
    int* rows_host = e->page_table_host + (size_t)e->batch * e->max_pages;
    int max_rows = 0;
    for (int b = 0; b < e->batch; ++b) {
        rows_host[b] = std::min(e->seq_cached[b], e->seq_pages[b] * B200_KV_PAGE_ROWS);
        max_rows = std::max(max_rows, rows_host[b]);
    }
    CU(cudaMemcpyAsync(e->page_rows_dev, rows_host, (size_t)e->batch * sizeof(int), cudaMemcpyHostToDevice, e->stream));
    const size_t dense_layer = (size_t)e->batch * e->hk * e->cfg.max_seq_len * e->d * e->esz;
    for (int l = 0; l < e->cfg.num_layers; ++l) {
        if (e->kv_fp8) {   // fp8 pages: dequantised on the way out
            const size_t po8 = (size_t)l * e->layer_kv_elems, so = (size_t)l * e->layer_kv_rows;
            RC(b200_kv_dequant_fp8_paged(e->k_cache + po8, e->k_scale + so, (char*)k_dst + l * dense_layer, e->page_table, e->max_pages,
                                         e->page_rows_dev, e->batch, e->hk, e->cfg.max_seq_len, e->d, max_rows, e->stream));
            RC(b200_kv_dequant_fp8_paged(e->v_cache + po8, e->v_scale + so, (char*)v_dst + l * dense_layer, e->page_table, e->max_pages,
                                         e->page_rows_dev, e->batch, e->hk, e->cfg.max_seq_len, e->d, max_rows, e->stream));
            continue;
        }
        const size_t po = (size_t)l * e->layer_kv_elems * e->esz;
        RC(b200_kv_gather_pages(e->k_cache + po, (char*)k_dst + l * dense_layer, e->page_table, e->max_pages, e->page_rows_dev,
                                e->batch, e->hk, e->cfg.max_seq_len, e->d, (int)e->esz, max_rows, e->stream));
        RC(b200_kv_gather_pages(e->v_cache + po, (char*)v_dst + l * dense_layer, e->page_table, e->max_pages, e->page_rows_dev,
                                e->batch, e->hk, e->cfg.max_seq_len, e->d, (int)e->esz, max_rows, e->stream));
    }
    CU(cudaStreamSynchronize(e->stream));
    return B200_OK;
}
/* fp8 cache only: the whole K and V caches dequantised to fp16 into caller DEVICE buffers [layers, B, hk, max_seq, d] */
extern "C" int b200_engine_kv_dequant(b200_engine* e, void* k_dst, void* v_dst) {
    if (!e || !e->kv_fp8 || !k_dst || !v_dst) return B200_EINVAL;
    if (e->kv_paged) return b200_engine_kv_gather(e, k_dst, v_dst);
    const size_t rows = e->layer_kv_rows * e->cfg.num_layers;
    RC(b200_kv_dequant_fp8(e->k_cache, e->k_scale, k_dst, rows, e->d, e->stream));
    RC(b200_kv_dequant_fp8(e->v_cache, e->v_scale, v_dst, rows, e->d, e->stream));
    CU(cudaStreamSynchronize(e->stream));
    return B200_OK;
}
extern "C" void* b200_engine_weight_ptr(b200_engine* e, int slot, int layer) { return e ? slot_ptr(e, slot, layer) : nullptr; }
extern "C" int b200_engine_launches_per_step(const b200_engine* e) { return e->launches_per_step; }
extern "C" int b200_engine_step_trace(b200_engine* e, unsigned long long* host_dst, int max_entries) {
    if (!e || !e->program) return 0;
    cudaStreamSynchronize(e->stream);
    return step_program_trace(e->program, host_dst, max_entries);
}

static void drop_graphs(b200_engine* e) {
    for (auto& kv : e->graphs) cudaGraphExecDestroy(kv.second);
    e->graphs.clear();
}

extern "C" int b200_engine_set_mode(b200_engine* e, unsigned flags) {
    if (!e) return B200_EINVAL;
    cudaStreamSynchronize(e->stream);
    drop_graphs(e);
    e->down_t_dirty = true;   // also the hook for callers that wrote weights through b200_engine_weight_ptr
    e->mode = flags;
    return B200_OK;
}
extern "C" int b200_engine_set_allreduce(b200_engine* e, b200_allreduce_fn fn, void* user) {
    if (!e) return B200_EINVAL;
    drop_graphs(e);
    e->allreduce = fn;
    e->allreduce_user = user;
    return B200_OK;
}

extern "C" int b200_engine_set_p2p(b200_engine* e, b200_p2p* comm) {
    if (!e) return B200_EINVAL;
    cudaStreamSynchronize(e->stream);
    drop_graphs(e);
    e->p2p = comm;
    {
        // the persistent step program carries the communicator's peer pointers (exchange form): rebuild it
        b200::step_program_destroy(e->program);
        e->program = nullptr;
        cudaFree(e->down_t); cudaFree(e->ffn_acc); cudaFree(e->o_packed); cudaFree(e->o_acc);
        e->down_t = nullptr; e->ffn_acc = nullptr; e->o_packed = nullptr; e->o_acc = nullptr;
        if (!e->kv_fp8 && !e->kv_paged) build_program(e);
    }
    return B200_OK;
}

// ------------------------------------------------------------------------------------------
// page allocator of the paged KV cache (host side; the device only reads the lists)
// ------------------------------------------------------------------------------------------
// Entries of a living sequence's list are only ever APPENDED, so a queued list upload that runs late and picks up newer
// entries is harmless (kernels ahead of it never look past their own row count). Pages change owner only through
// release_pages, which callers run behind a stream synchronisation.
static void release_pages(b200_engine* e, int b) {
    int* list = e->page_table_host + (size_t)b * e->max_pages;
    for (int i = e->seq_pages[b] - 1; i >= 0; --i) e->free_pages.push_back(list[i]);
    e->seq_pages[b] = 0;
}
// make sure sequence b owns the pages of rows [0, rows); returns 1 if its list grew, 0 if not, < 0 on exhaustion
static int ensure_pages(b200_engine* e, int b, int rows) {
    const int need = (rows + B200_KV_PAGE_ROWS - 1) / B200_KV_PAGE_ROWS;
    if (need > e->max_pages) return fail("paged KV: sequence longer than max_seq_len");
    if (need <= e->seq_pages[b]) return 0;
    if ((size_t)(need - e->seq_pages[b]) > e->free_pages.size())
        return fail("paged KV: page pool exhausted (" + std::to_string(e->n_pages) + " pages of 128 rows)", B200_ENOMEM);
    int* list = e->page_table_host + (size_t)b * e->max_pages;
    while (e->seq_pages[b] < need) {
        list[e->seq_pages[b]++] = e->free_pages.back();
        e->free_pages.pop_back();
    }
    e->pages_dirty = true;
    return 1;
}
static int upload_pages(b200_engine* e) {
    CU(cudaMemcpyAsync(e->page_table, e->page_table_host, (size_t)e->batch * e->max_pages * sizeof(int),
                       cudaMemcpyHostToDevice, e->stream));
    e->pages_dirty = false;
    return 0;
}
// pages for `rows_b[b]` rows of every sequence (rows_b[b] < 0: leave sequence b alone)
static int ensure_all_pages(b200_engine* e, const int* rows_b) {
    for (int b = 0; b < e->batch; ++b) {
        if (rows_b[b] < 0) continue;
        const int r = ensure_pages(e, b, rows_b[b]);
        if (r < 0) return r;   // lists that grew before the failure stay marked dirty and go up with the next call
    }
    return e->pages_dirty ? upload_pages(e) : 0;
}

static int set_step(b200_engine* e, int n) {
    if (e->kv_paged) {   // a new position for everybody: hand every page back, then take what n rows need
        CU(cudaStreamSynchronize(e->stream));
        for (int b = 0; b < e->batch; ++b) release_pages(e, b);
        e->seq_released.assign(e->batch, 0);
        std::vector<int> rows(e->batch, n);
        const int rc = ensure_all_pages(e, rows.data());
        if (rc != 0) return rc;
    }
    e->step = n;
    e->seq_cached.assign(e->batch, n);
    e->ragged = false;
    const int ref_step = n + 1;  // reference `step`: tokens in the cache including the new one
    CU(cudaMemcpyAsync(e->step_dev, &ref_step, sizeof(int), cudaMemcpyHostToDevice, e->stream));
    CU(cudaMemsetAsync(e->finished, 0, e->batch, e->stream));   // a new position: nobody is retired
    CU(cudaStreamSynchronize(e->stream));
    return B200_OK;
}

extern "C" int b200_engine_reset(b200_engine* e) { return e ? set_step(e, 0) : B200_EINVAL; }

/* One sequence slot back to "nothing cached" while the others keep their caches and positions (the data-path half of
 * continuous batching: a retired sequence's slot takes the next request; the reference's loop, llama.cpp:319-361, only
 * ever runs a fixed batch to the end). Clears the slot's `finished` flag, returns its pages (paged cache) and puts the
 * engine into the per-sequence (ragged) step form. Follow with b200_engine_prefill_ragged, length 0 for the other slots. */
extern "C" int b200_engine_reset_slot(b200_engine* e, int b) {
    if (!e || b < 0 || b >= e->batch) return fail("reset_slot: bad slot");
    CU(cudaStreamSynchronize(e->stream));
    if (!e->ragged) {   // the per-sequence step array takes over from the shared step
        std::vector<int> steps(e->batch);
        for (int i = 0; i < e->batch; ++i) steps[i] = e->seq_cached[i] + 1;
        CU(cudaMemcpy(e->seq_steps, steps.data(), sizeof(int) * e->batch, cudaMemcpyHostToDevice));
        e->ragged = true;
    }
    const int one = 1;
    CU(cudaMemcpy(e->seq_steps + b, &one, sizeof(int), cudaMemcpyHostToDevice));
    CU(cudaMemset(e->finished + b, 0, 1));
    e->seq_cached[b] = 0;
    if (e->kv_paged) {
        release_pages(e, b);
        e->seq_released[b] = 0;
    }
    return B200_OK;
}

/* eos_id >= 0 switches sequence retirement on (sampling.cu:81-83: finished[b] = (id == end_id)); -1 off. The flags are
 * cleared by reset / fill_kv / every prompt pass. The persistent whole-step kernel does not look at them (its default use
 * is batch 1, where the generation loop simply stops): engines with retirement on run the multi-kernel orders. */
extern "C" int b200_engine_set_eos(b200_engine* e, int eos_id) {
    if (!e) return B200_EINVAL;
    CU(cudaStreamSynchronize(e->stream));
    drop_graphs(e);   // the advance kernel's arguments are baked into the graphs
    e->eos_id = eos_id < 0 ? -1 : eos_id;
    CU(cudaMemsetAsync(e->finished, 0, e->batch, e->stream));
    return B200_OK;
}
extern "C" int b200_engine_get_finished(b200_engine* e, unsigned char* flags_host) {
    if (!e || !flags_host) return B200_EINVAL;
    CU(cudaMemcpyAsync(flags_host, e->finished, e->batch, cudaMemcpyDeviceToHost, e->stream));
    CU(cudaStreamSynchronize(e->stream));
    return B200_OK;
}

template <typename T>
static void fill_normal(void* p, size_t n, uint64_t seed, float mean, float std, cudaStream_t s) {
    size_t blocks = (n + 255) / 256;
    if (blocks > 148 * 16) blocks = 148 * 16;
    fill_normal_kernel<T><<<(unsigned)blocks, 256, 0, s>>>((T*)p, n, seed, mean, std);
}

extern "C" int b200_engine_init_random(b200_engine* e, uint64_t seed, float std) {
    if (!e) return B200_EINVAL;
    uint64_t k = seed * 1000003ull;
    auto fill = [&](void* p, int slot, float mean) {
        const size_t n = slot_elems(e, slot);
        if (e->cfg.dtype == B200_F16) fill_normal<__half>(p, n, ++k, mean, std, e->stream);
        else fill_normal<float>(p, n, ++k, mean, std, e->stream);
    };
    for (auto& w : e->layers) {
        fill(w.attn_norm, B200_W_ATTN_NORM, 1.f);
        fill(w.qkv, B200_W_QKV, 0.f);
        fill(w.o, B200_W_O, 0.f);
        fill(w.ffn_norm, B200_W_FFN_NORM, 1.f);
        fill(w.gate_up, B200_W_GATE_UP, 0.f);
        fill(w.down, B200_W_DOWN, 0.f);
    }
    fill(e->embed, B200_W_EMBED, 0.f);
    fill(e->final_norm, B200_W_FINAL_NORM, 1.f);
    fill(e->lm_head, B200_W_LM_HEAD, 0.f);
    e->down_t_dirty = true;
    CU(cudaStreamSynchronize(e->stream));
    CU(cudaGetLastError());
    return B200_OK;
}

extern "C" int b200_engine_set_weight(b200_engine* e, int slot, int layer, const void* src, size_t bytes) {
    if (!e || !src) return B200_EINVAL;
    void* p = slot_ptr(e, slot, layer);
    if (!p || bytes != slot_elems(e, slot) * e->esz) return fail("set_weight: bad slot / size");
    CU(cudaMemcpy(p, src, bytes, cudaMemcpyHostToDevice));
    if (slot == B200_W_DOWN || slot == B200_W_O) e->down_t_dirty = true;
    return B200_OK;
}
extern "C" int b200_engine_get_weight(b200_engine* e, int slot, int layer, void* dst, size_t bytes) {
    if (!e || !dst) return B200_EINVAL;
    void* p = slot_ptr(e, slot, layer);
    if (!p || bytes != slot_elems(e, slot) * e->esz) return fail("get_weight: bad slot / size");
    CU(cudaMemcpy(dst, p, bytes, cudaMemcpyDeviceToHost));
    return B200_OK;
}

/* paged cache: pages of the sequences the device has retired go back to the pool (they are never read again: the
 * attention kernels skip a finished sequence). Returns the number of pages released. */
extern "C" int b200_engine_release_finished(b200_engine* e) {
    if (!e) return B200_EINVAL;
    if (!e->kv_paged || e->eos_id < 0) return 0;
    std::vector<unsigned char> flags(e->batch);
    const int rc = b200_engine_get_finished(e, flags.data());   // synchronises the stream
    if (rc != 0) return rc > 0 ? -1000 - rc : rc;
    int released = 0;
    for (int b = 0; b < e->batch; ++b)
        if (flags[b] && e->seq_pages[b] > 0) {
            released += e->seq_pages[b];
            release_pages(e, b);
            e->seq_released[b] = 1;
        }
    return released;
}

extern "C" int b200_engine_fill_kv_random(b200_engine* e, int len, uint64_t seed, float amp) {
    if (!e || len < 0 || len >= e->cfg.max_seq_len) return fail("fill_kv: len out of range");
    if (e->kv_paged) {   // every page of every pool gets synthetic rows; set_step hands out the pages of `len` rows
        const size_t prow = (size_t)e->cfg.num_layers * e->n_pages * e->hk;
        if (e->kv_fp8) {
            RC(kv_fill_fp8(e->k_cache, e->k_scale, prow, B200_KV_PAGE_ROWS, B200_KV_PAGE_ROWS, e->d, seed * 2 + 1, amp, e->stream));
            RC(kv_fill_fp8(e->v_cache, e->v_scale, prow, B200_KV_PAGE_ROWS, B200_KV_PAGE_ROWS, e->d, seed * 2 + 2, amp, e->stream));
            return set_step(e, len);
        }
        const size_t n = prow * B200_KV_PAGE_ROWS * e->d;
        size_t blocks = (n + 255) / 256;
        if (blocks > 148 * 32) blocks = 148 * 32;
        fill_uniform_rows_kernel<__half><<<(unsigned)blocks, 256, 0, e->stream>>>(
            (__half*)e->k_cache, prow, B200_KV_PAGE_ROWS, B200_KV_PAGE_ROWS, e->d, seed * 2 + 1, amp);
        fill_uniform_rows_kernel<__half><<<(unsigned)blocks, 256, 0, e->stream>>>(
            (__half*)e->v_cache, prow, B200_KV_PAGE_ROWS, B200_KV_PAGE_ROWS, e->d, seed * 2 + 2, amp);
        CU(cudaGetLastError());
        return set_step(e, len);
    }
    const size_t rows = (size_t)e->cfg.num_layers * e->batch * e->hk;
    if (len > 0 && e->kv_fp8) {
        RC(kv_fill_fp8(e->k_cache, e->k_scale, rows, e->cfg.max_seq_len, len, e->d, seed * 2 + 1, amp, e->stream));
        RC(kv_fill_fp8(e->v_cache, e->v_scale, rows, e->cfg.max_seq_len, len, e->d, seed * 2 + 2, amp, e->stream));
    } else if (len > 0) {
        const size_t n = rows * (size_t)len * e->d;
        size_t blocks = (n + 255) / 256;
        if (blocks > 148 * 32) blocks = 148 * 32;
        if (e->cfg.dtype == B200_F16) {
            fill_uniform_rows_kernel<__half><<<(unsigned)blocks, 256, 0, e->stream>>>(
                (__half*)e->k_cache, rows, e->cfg.max_seq_len, len, e->d, seed * 2 + 1, amp);
            fill_uniform_rows_kernel<__half><<<(unsigned)blocks, 256, 0, e->stream>>>(
                (__half*)e->v_cache, rows, e->cfg.max_seq_len, len, e->d, seed * 2 + 2, amp);
        } else {
            fill_uniform_rows_kernel<float><<<(unsigned)blocks, 256, 0, e->stream>>>(
                (float*)e->k_cache, rows, e->cfg.max_seq_len, len, e->d, seed * 2 + 1, amp);
            fill_uniform_rows_kernel<float><<<(unsigned)blocks, 256, 0, e->stream>>>(
                (float*)e->v_cache, rows, e->cfg.max_seq_len, len, e->d, seed * 2 + 2, amp);
        }
        CU(cudaGetLastError());
    }
    return set_step(e, len);
}

// ------------------------------------------------------------------------------------------
// one decode step
// ------------------------------------------------------------------------------------------
namespace b200 {
int attn_pick_splits_public(int batch, int kv_head_num, int step, int group);  // decode_attention.cu
}

static int gemv(b200_engine* e, const void* x, const void* W, void* y, int N, int K, unsigned flags,
                const void* gamma = nullptr, int layer = -1, bool push_p2p = false) {
    b200_gemv_args a{};
    if (push_p2p) a.p2p = e->p2p;
    a.x = x; a.W = W; a.y = y; a.gamma = gamma; a.eps = e->cfg.rms_eps;
    a.n_tokens = e->batch; a.N = N; a.K = K; a.flags = flags;
    if (flags & B200_GEMV_ROPE_KV) {
        a.k_cache_layer = e->k_cache + (size_t)layer * e->layer_kv_elems * e->esz;
        a.v_cache_layer = e->v_cache + (size_t)layer * e->layer_kv_elems * e->esz;
        a.head_num = e->h; a.kv_head_num = e->hk; a.head_size = e->d;
        a.max_seq_len = e->cfg.max_seq_len; a.step = e->step + 1;
        a.rotary_dim = e->cfg.rotary_dim; a.rotary_base = e->cfg.rotary_base;
        a.step_dev = e->ragged ? e->seq_steps : e->step_dev;
        a.step_stride = e->ragged ? 1 : 0;
        a.rope_den = e->rope_den;
    }
    // batches above the GEMV tile are walked in tiles of B200_GEMV_MAX_TOKENS
    const int B = e->batch;
    for (int t0 = 0; t0 < B; t0 += B200_GEMV_MAX_TOKENS) {
        b200_gemv_args t = a;
        const int nt = (B - t0 < B200_GEMV_MAX_TOKENS) ? B - t0 : B200_GEMV_MAX_TOKENS;
        t.n_tokens = nt;
        t.x = (const char*)x + (size_t)t0 * K * e->esz;
        t.y = (char*)y + (size_t)t0 * N * e->esz;
        if (flags & B200_GEMV_ROPE_KV) {
            const size_t per_seq = (size_t)e->hk * e->cfg.max_seq_len * e->d * e->esz;
            t.k_cache_layer = (char*)a.k_cache_layer + (size_t)t0 * per_seq;
            t.v_cache_layer = (char*)a.v_cache_layer + (size_t)t0 * per_seq;
            if (a.step_stride) t.step_dev = a.step_dev + t0;
        }
        int rc = b200_gemv_fused(&t, e->cfg.dtype, e->stream);
        if (rc != 0) return rc;
    }
    return 0;
}

static int tp_allreduce_add(b200_engine* e) {
    // partial [B,H] -> sum over ranks -> hidden += partial
    if (e->p2p) {
        RC(b200_p2p_allreduce_add(e->p2p, e->partial, e->hidden, (size_t)e->batch * e->H, e->cfg.dtype, e->stream));
        return 0;
    }
    if (e->allreduce) RC(e->allreduce(e->partial, (size_t)e->batch * e->H, e->cfg.dtype, e->stream, e->allreduce_user));
    RC(b200_add_residual(e->partial, e->hidden, e->batch, e->H, e->cfg.dtype, e->stream));
    return 0;
}

// in-place sum of buf over the tensor-parallel group: one-shot NVLink kernel when it fits, else the hook
static int tp_sum(b200_engine* e, void* buf, size_t count) {
    if (e->p2p && count * e->esz <= b200_p2p_max_message_bytes(e->p2p))
        RC(b200_p2p_allreduce(e->p2p, buf, count, e->cfg.dtype, e->stream));
    else if (e->allreduce)
        RC(e->allreduce(buf, count, e->cfg.dtype, e->stream, e->allreduce_user));
    else if (e->p2p)   // never skip the exchange silently: partial sums would pass for results
        return fail("tensor-parallel sum: message larger than the b200_p2p communicator and no allreduce hook installed");
    return 0;
}

static int enqueue_step_body(b200_engine* e);
static int enqueue_step(b200_engine* e) {
    b200_decode_attention_set_finished(e->eos_id >= 0 ? e->finished : nullptr);
    const int rc = enqueue_step_body(e);
    b200_decode_attention_set_finished(nullptr);
    return rc;
}
static int enqueue_step_body(b200_engine* e) {
    const int dt = e->cfg.dtype, B = e->batch, H = e->H;
    const int ref_step = e->step + 1;
    cudaStream_t s = e->stream;
    const bool tp = e->cfg.tp_size > 1;
    // fused GEMV path: batches that fit the GEMV tile. Larger batches run the reference's launch
    // sequence with the linears on the tcgen05 weight-streaming GEMM (weights are read once per
    // step instead of once per 8-token tile).
    // fp16 batches of 5..8 also take the tensor-core path: eight x-rows of K = 11008 crowd the GEMV's weight ring out
    // of shared memory (measured at batch 8 x KV 4096: 12.0 ms per step fused vs the figure in DESIGN §4c)
    // (the fused qkv GEMV appends fp16 rows in its epilogue: an fp8 cache takes the reference launch order)
    const bool fused = (e->mode & B200_ENGINE_FUSED) && B <= (dt == B200_F16 && !tp ? 4 : B200_GEMV_MAX_TOKENS) && !e->kv_fp8 &&
                       !e->kv_paged;
    auto linear = [&](const void* x, const void* W, void* y, int N, int K) {
        return b200_linear(x, W, nullptr, y, B, N, K, 1, dt, e->lin_ws, e->lin_ws_bytes, s);
    };
    RC(b200_embedding(e->ids_cur, e->embed, e->hidden, B, H, e->V, dt, s));
    for (int l = 0; l < e->cfg.num_layers; ++l) {
        LayerW& w = e->layers[l];
        void* kc = e->k_cache + (size_t)l * e->layer_kv_elems * e->kv_esz;
        void* vc = e->v_cache + (size_t)l * e->layer_kv_elems * e->kv_esz;
        if (fused) {
            RC(gemv(e, e->hidden, w.qkv, e->qkv, e->qkv_n, H, B200_GEMV_NORM | B200_GEMV_ROPE_KV, w.attn_norm, l));
            if (e->ragged)
                RC(b200_decode_attention_ragged(e->qkv, kc, vc, e->attn, e->attn_ws, e->attn_ws_bytes, B, e->h, e->hk,
                                                e->d, e->cfg.max_seq_len, ref_step, e->seq_steps, 0, dt, s));
            else
                RC(b200_decode_attention_ex(e->qkv, kc, vc, e->attn, e->attn_ws, e->attn_ws_bytes, B, e->h, e->hk,
                                            e->d, e->cfg.max_seq_len, ref_step, e->step_dev, 0, dt, s));
            const bool fused_ar = tp && e->p2p && dt == B200_F16 && (H & 3) == 0 && B <= B200_GEMV_MAX_TOKENS;
            if (!tp) {
                RC(gemv(e, e->attn, w.o, e->hidden, H, e->h * e->d, B200_GEMV_RESIDUAL));
            } else if (fused_ar) {   // GEMV epilogue pushes its rows to the peers; the reduce kernel sums + adds
                RC(gemv(e, e->attn, w.o, e->partial, H, e->h * e->d, 0, nullptr, -1, true));
                RC(b200_p2p_reduce_add(e->p2p, e->hidden, (size_t)B * H, dt, s));
            } else {
                RC(gemv(e, e->attn, w.o, e->partial, H, e->h * e->d, 0));
                RC(tp_allreduce_add(e));
            }
            RC(gemv(e, e->hidden, w.gate_up, e->act, e->I, H,
                    B200_GEMV_NORM | B200_GEMV_NORM_GFIRST | B200_GEMV_SWIGLU, w.ffn_norm));
            if (!tp) {
                RC(gemv(e, e->act, w.down, e->hidden, H, e->I, B200_GEMV_RESIDUAL));
            } else if (fused_ar) {
                RC(gemv(e, e->act, w.down, e->partial, H, e->I, 0, nullptr, -1, true));
                RC(b200_p2p_reduce_add(e->p2p, e->hidden, (size_t)B * H, dt, s));
            } else {
                RC(gemv(e, e->act, w.down, e->partial, H, e->I, 0));
                RC(tp_allreduce_add(e));
            }
        } else {
            // the reference's launch sequence, one kernel per launch* (self_decoder.cpp:53-85)
            RC(b200_rmsnorm(e->hidden, e->residual, w.attn_norm, B, H, e->cfg.rms_eps, dt, s));
            RC(linear(e->hidden, w.qkv, e->qkv, e->qkv_n, H));
            if (e->kv_fp8) {   // RoPE, then the new rows enter the fp8 cache (quantised once), then attention over the cache
                float* ksc = e->k_scale + (size_t)l * e->layer_kv_rows;
                float* vsc = e->v_scale + (size_t)l * e->layer_kv_rows;
                const int* sd = e->ragged ? e->seq_steps : e->step_dev;
                const int stride = e->ragged ? 1 : 0;
                if (e->ragged)
                    RC(b200_rope_decode_ragged(e->qkv, B, e->h, e->hk, e->d, e->seq_steps, e->cfg.rotary_dim,
                                               e->cfg.rotary_base, e->rope_den, dt, s));
                else
                    RC(b200_rope_decode_tab(e->qkv, B, e->h, e->hk, e->d, ref_step, e->step_dev, e->cfg.rotary_dim,
                                            e->cfg.rotary_base, e->rope_den, dt, s));
                if (e->kv_paged) {
                    RC(b200_kv_append_fp8_paged(e->qkv, kc, vc, ksc, vsc, e->page_table, e->max_pages, B, e->h, e->hk, e->d, ref_step,
                                                sd, stride, s));
                    RC(b200_decode_attention_fp8_paged(e->qkv, kc, vc, ksc, vsc, e->page_table, e->max_pages, e->attn, e->attn_ws,
                                                       e->attn_ws_bytes, B, e->h, e->hk, e->d, ref_step, sd, stride, s));
                } else {
                    RC(b200_kv_append_fp8(e->qkv, kc, vc, ksc, vsc, B, e->h, e->hk, e->d, e->cfg.max_seq_len, ref_step, sd, stride, s));
                    RC(b200_decode_attention_fp8(e->qkv, kc, vc, ksc, vsc, e->attn, e->attn_ws, e->attn_ws_bytes, B, e->h, e->hk,
                                                 e->d, e->cfg.max_seq_len, ref_step, sd, stride, s));
                }
            } else if (e->kv_paged) {   // RoPE, then the attention appends the new row into / reads the rows out of the pages
                if (e->ragged)
                    RC(b200_rope_decode_ragged(e->qkv, B, e->h, e->hk, e->d, e->seq_steps, e->cfg.rotary_dim,
                                               e->cfg.rotary_base, e->rope_den, dt, s));
                else
                    RC(b200_rope_decode_tab(e->qkv, B, e->h, e->hk, e->d, ref_step, e->step_dev, e->cfg.rotary_dim,
                                            e->cfg.rotary_base, e->rope_den, dt, s));
                RC(b200_decode_attention_paged(e->qkv, kc, vc, e->page_table, e->max_pages, e->attn, e->attn_ws, e->attn_ws_bytes,
                                               B, e->h, e->hk, e->d, ref_step, e->ragged ? e->seq_steps : e->step_dev,
                                               e->ragged ? 1 : 0, 1, s));
            } else if (e->ragged) {   // per-sequence positions and key counts; ref_step (the longest) only sizes the launch
                RC(b200_rope_decode_ragged(e->qkv, B, e->h, e->hk, e->d, e->seq_steps, e->cfg.rotary_dim,
                                           e->cfg.rotary_base, e->rope_den, dt, s));
                RC(b200_decode_attention_ragged(e->qkv, kc, vc, e->attn, e->attn_ws, e->attn_ws_bytes, B, e->h, e->hk,
                                                e->d, e->cfg.max_seq_len, ref_step, e->seq_steps, 1, dt, s));
            } else {
                RC(b200_rope_decode_tab(e->qkv, B, e->h, e->hk, e->d, ref_step, e->step_dev, e->cfg.rotary_dim,
                                        e->cfg.rotary_base, e->rope_den, dt, s));
                RC(b200_decode_attention_ex(e->qkv, kc, vc, e->attn, e->attn_ws, e->attn_ws_bytes, B, e->h, e->hk, e->d,
                                            e->cfg.max_seq_len, ref_step, e->step_dev, 1, dt, s));
            }
            RC(linear(e->attn, w.o, e->hidden, H, e->h * e->d));
            if (tp) RC(tp_sum(e, e->hidden, (size_t)B * H));
            RC(b200_fused_add_bias_residual_rmsnorm(e->residual, e->hidden, nullptr, w.ffn_norm, B, H, e->cfg.rms_eps, dt, s));
            RC(linear(e->hidden, w.gate_up, e->gate_up, 2 * e->I, H));
            RC(b200_swiglu(e->gate_up, e->act, B, e->I, dt, s));
            RC(linear(e->act, w.down, e->hidden, H, e->I));
            if (tp) RC(tp_sum(e, e->hidden, (size_t)B * H));
            RC(b200_add_residual(e->residual, e->hidden, B, H, dt, s));
        }
    }
    if (fused) {
        RC(gemv(e, e->hidden, e->lm_head, e->logits, e->V, H, B200_GEMV_NORM, e->final_norm));
    } else {
        // llama.cpp:341 normalises decoder_output in place; we do it on a copy so that
        // b200_engine_hidden() shows the decoder output in both modes
        CU(cudaMemcpyAsync(e->partial, e->hidden, (size_t)B * H * e->esz, cudaMemcpyDeviceToDevice, s));
        RC(b200_rmsnorm(e->partial, nullptr, e->final_norm, B, H, e->cfg.rms_eps, dt, s));
        RC(linear(e->partial, e->lm_head, e->logits, e->V, H));
    }
    RC(b200_argmax(e->logits, e->ids_next, B, e->V, dt, s));
    if (e->ragged)
        CU(launch_k(advance_ragged_kernel, dim3(1), dim3(1024), 0, s, true, e->ids_cur, e->ids_next,
                    e->step_dev, e->seq_steps, B, e->eos_id >= 0 ? e->finished : nullptr, e->eos_id));
    else
        CU(launch_k(advance_kernel, dim3(1), dim3(1024), 0, s, true, e->ids_cur, e->ids_next, e->step_dev, B,
                    e->eos_id >= 0 ? e->finished : nullptr, e->eos_id));
    count_launch();
    return 0;
}

static void stepped(b200_engine* e) {
    e->step += 1;
    for (int& c : e->seq_cached) c += 1;
}

static int run_step(b200_engine* e) {
    if (e->step + 1 > e->cfg.max_seq_len) return fail("KV cache full (step == max_seq_len)");
    if (e->kv_paged) {   // the row this step appends may open a new page (a retired sequence with no pages stays without)
        std::vector<int> rows(e->batch);
        for (int b = 0; b < e->batch; ++b)
            rows[b] = e->seq_released[b] ? -1 : e->seq_cached[b] + 1;
        const int rc = ensure_all_pages(e, rows.data());
        if (rc != 0) return rc;
    }
    if (e->program && !e->ragged && (e->mode & B200_ENGINE_FUSED) && (e->mode & B200_ENGINE_PERSISTENT) &&
        !(e->eos_id >= 0 && e->batch > 1)) {
        RC(refresh_down_t(e));
        RC(step_program_launch(e->program, e->step + 1, e->stream));
        e->launches_per_step = 1;
        stepped(e);
        return 0;
    }
    const bool use_graph = (e->mode & B200_ENGINE_GRAPH) != 0;
    if (!use_graph) {
        const uint64_t before = b200_launch_count();
        RC(enqueue_step(e));
        e->launches_per_step = (int)(b200_launch_count() - before);
        stepped(e);
        return 0;
    }
    // one graph per split count of the decode attention; ragged batches record different kernels
    const int key = attn_pick_splits_public(e->batch, e->hk, e->step + 1, e->h / e->hk) | (e->ragged ? 1 << 20 : 0);
    auto it = e->graphs.find(key);
    if (it == e->graphs.end()) {
        const uint64_t before = b200_launch_count();
        cudaGraph_t g = nullptr;
        CU(cudaStreamBeginCapture(e->stream, cudaStreamCaptureModeThreadLocal));
        int rc = enqueue_step(e);
        cudaError_t ce = cudaStreamEndCapture(e->stream, &g);
        if (rc != 0) { if (g) cudaGraphDestroy(g); return rc; }
        if (ce != cudaSuccess) return fail(std::string("graph capture: ") + cudaGetErrorString(ce), (int)ce);
        cudaGraphExec_t ge = nullptr;
        ce = cudaGraphInstantiate(&ge, g, 0);
        cudaGraphDestroy(g);
        if (ce != cudaSuccess) return fail(std::string("graph instantiate: ") + cudaGetErrorString(ce), (int)ce);
        e->launches_per_step = (int)(b200_launch_count() - before);
        // the capture itself launched nothing: take its count back, replays are counted below
        it = e->graphs.emplace(key, ge).first;
    } else {
        count_launch((uint64_t)e->launches_per_step);
    }
    CU(cudaGraphLaunch(it->second, e->stream));
    stepped(e);
    return 0;
}

extern "C" int b200_engine_decode_step(b200_engine* e, const int* ids_in_dev) {
    if (!e) return B200_EINVAL;
    if (ids_in_dev)
        CU(cudaMemcpyAsync(e->ids_cur, ids_in_dev, sizeof(int) * e->batch, cudaMemcpyDeviceToDevice, e->stream));
    return run_step(e);
}

extern "C" int b200_engine_decode_step_host(b200_engine* e, const int* ids_in_host, int* ids_out_host) {
    if (!e || !ids_out_host) return B200_EINVAL;
    const size_t nb = sizeof(int) * e->batch;
    if (ids_in_host) {
        memcpy(e->pinned_ids, ids_in_host, nb);
        CU(cudaMemcpyAsync(e->ids_cur, e->pinned_ids, nb, cudaMemcpyHostToDevice, e->stream));
    }
    int rc = run_step(e);
    if (rc != 0) return rc;
    CU(cudaMemcpyAsync(e->pinned_ids + e->batch, e->ids_next, nb, cudaMemcpyDeviceToHost, e->stream));
    CU(cudaStreamSynchronize(e->stream));
    memcpy(ids_out_host, e->pinned_ids + e->batch, nb);
    return 0;
}


// ------------------------------------------------------------------------------------------
// prefill: the prompt pass of Llama<T>::firstTokenGen (llama.cpp:274-318) =
// launchInputEmbedding -> LlamaContextDecoder<T>::forward (context_decoder.cpp:48-140; per layer
// RMSNorm -> qkv GEMM -> bias/RoPE/transpose -> ConcatKV -> attention -> o GEMM -> fused add+norm ->
// gate_up GEMM -> SwiGLU -> down GEMM -> AddResidual) -> final RMSNorm -> LM head on the last token of
// each sequence (llama.cpp:221-232) -> top-1. All sequences of the engine share one `step`
// (llama.h keeps a single host `step` as well), so every sequence gets `len` prompt tokens, appended
// after the `step` tokens already cached (history_length = step: the reference's multi-round chat).
// The attention is b200_context_attention (tcgen05 flash kernel); for shapes it does not cover
// (fp32, head size != 128) the reference-order GPU sequence runs instead.
// ------------------------------------------------------------------------------------------
namespace b200 {
// after a prompt pass: the token each sequence decodes next. A sequence that took no tokens in this pass (lens[b] == 0: the
// others were being (re)filled around it) keeps the token it was about to decode, mirrored into ids_next for the readers.
__global__ void set_ids_step_kernel(int* ids_cur, int* ids_next, int* step, int batch, int ref_step, const int* lens) {
    pdl_wait();
    if (threadIdx.x < batch) {
        if (lens != nullptr && lens[threadIdx.x] == 0) ids_next[threadIdx.x] = ids_cur[threadIdx.x];
        else ids_cur[threadIdx.x] = ids_next[threadIdx.x];
    }
    if (threadIdx.x == 0) *step = ref_step;
}
}  // namespace b200

// lens_host == nullptr: every sequence gets `uniform_len` tokens. Otherwise sequence b gets lens_host[b] tokens and
// ids_host holds them packed back to back (the reference's padding-free [num_tokens] layout, llama.cpp:158-185);
// each sequence's history is what IT has cached so far.
static int prefill_impl(b200_engine* e, const int* ids_host, const int* lens_host, int uniform_len, int* ids_out_host) {
    const int dt = e->cfg.dtype, B = e->batch, H = e->H;
    long long total = 0;
    int len = 0, ctx = 0;   // longest prompt / longest context of this call: the padded extents
    std::vector<int> len_b(B), hist_b(B), cum_b(B + 1, 0);
    bool any_zero = false;
    for (int b = 0; b < B; ++b) {
        len_b[b] = lens_host ? lens_host[b] : uniform_len;
        hist_b[b] = e->seq_cached[b];
        // ragged form only: a sequence may sit a pass out (0 tokens) while others are (re)filled - it keeps its cache, its
        // position, its `finished` flag and the token it decodes next (continuous batching: b200_engine_reset_slot + this)
        if (len_b[b] < (lens_host ? 0 : 1)) return fail("prefill: a prompt length is negative (or 0 in the uniform form)");
        any_zero |= len_b[b] == 0;
        if (hist_b[b] + len_b[b] > e->cfg.max_seq_len - 1) return fail("prefill: prompt does not fit the KV cache");
        total += len_b[b];
        cum_b[b + 1] = (int)total;
        len = std::max(len, len_b[b]);
        ctx = std::max(ctx, hist_b[b] + len_b[b]);
    }
    if (total < 1) return fail("prefill: no prompt tokens at all");
    if ((long long)B * len >= (1ll << 24)) return fail("prefill: too many tokens");
    const int Nt = (int)total;
    const size_t Np = (size_t)B * len;   // rows of the re-padded [B, h, len, d] buffers
    bool ragged = false;
    for (int b = 1; b < B; ++b) ragged |= hist_b[b] + len_b[b] != hist_b[0] + len_b[0];
    const size_t es = e->esz;
    cudaStream_t s = e->stream;
    const uint64_t launches0 = b200_launch_count();
    // ---- scratch layout
    size_t off = 0;
    auto take = [&](size_t bytes) { size_t o = off; off += align256(bytes); return o; };
    const size_t o_x = take((size_t)Nt * H * es), o_res = take((size_t)Nt * H * es);
    const size_t o_qkv = take((size_t)Nt * e->qkv_n * es);
    const size_t o_q = take(Np * e->h * e->d * es), o_k = take(Np * e->hk * e->d * es),
                 o_v = take(Np * e->hk * e->d * es);
    const size_t o_att = take((size_t)Nt * e->h * e->d * es);
    const size_t o_gu = take((size_t)Nt * 2 * e->I * es), o_act = take((size_t)Nt * e->I * es);
    const size_t o_ids = take((size_t)Nt * 4), o_pad = take((size_t)Nt * 4), o_cum = take((size_t)(B + 1) * 4),
                 o_len = take((size_t)B * 4), o_hist = take((size_t)B * 4), o_ctx = take((size_t)B * 4);
    size_t lin_ws = 0;
    {
        const int shapes[4][2] = {{e->qkv_n, H}, {H, e->h * e->d}, {2 * e->I, H}, {H, e->I}};
        for (auto& sh : shapes) lin_ws = std::max(lin_ws, b200_linear_workspace_bytes(Nt, sh[0], sh[1], 1, dt));
    }
    const size_t o_lws = take(lin_ws + 256);
    const bool flash = dt == B200_F16 && e->d == 128 && !(getenv("B200_NO_TCGEN05") && getenv("B200_NO_TCGEN05")[0] == '1');
    if (any_zero && !flash)
        return fail("prefill: sequences that sit a pass out need the flash context attention (fp16, head size 128)", B200_EUNSUPPORTED);
    if (e->kv_fp8) {
        if (!flash) return fail("fp8 KV cache: the prompt pass needs the flash context attention", B200_EUNSUPPORTED);
        for (int b = 0; b < B; ++b)
            if (hist_b[b] != 0 && len_b[b] > 0)
                return fail("fp8 KV cache: a prompt pass over an existing history is not supported (first pass only)",
                            B200_EUNSUPPORTED);
    }
    if (e->kv_paged) {
        if (!flash) return fail("paged KV cache: the prompt pass needs the flash context attention", B200_EUNSUPPORTED);
        for (int b = 0; b < B; ++b)
            if (e->seq_released[b] && len_b[b] > 0)
                return fail("paged KV cache: a retired sequence gave its pages back - b200_engine_reset_slot before its next prompt");
        // the pages of history + prompt (fp16 pages: the context attention reads the history through the page lists; fp8
        // pages are first-pass only, checked above)
        for (int b = 0; b < B; ++b) {
            const int r = ensure_pages(e, b, hist_b[b] + len_b[b]);
            if (r < 0) return r;   // the pool cannot hold these prompts: nothing launched, cached rows and lengths untouched (lists
                                   // that grew before the failure keep their pages: the sequences grow into them, reset returns them)
        }
        RC(upload_pages(e));
    }
    size_t o_mask = 0, o_kr = 0, o_vr = 0, o_sc = 0, o_pv = 0;
    if (!flash) {   // reference-order attention: dense mask, repeated KV, materialised scores
        o_mask = take((size_t)B * len * ctx * es);
        o_kr = take((size_t)B * e->h * ctx * e->d * es);
        o_vr = take((size_t)B * e->h * ctx * e->d * es);
        o_sc = take((size_t)B * e->h * len * ctx * es);
        o_pv = take(Np * e->h * e->d * es);
    }
    if (off > e->pf_bytes) {
        CU(cudaStreamSynchronize(s));
        cudaFree(e->pf_arena);
        e->pf_arena = nullptr;
        e->pf_bytes = 0;
        CU(cudaMalloc(&e->pf_arena, off));
        e->pf_bytes = off;
    }
    const size_t need_ints = (size_t)Nt + 4 * B;
    // the staging buffer is reused by every prompt pass: the copies a previous (unsynchronised) pass queued
    // from it must have run before it is overwritten or freed
    if (e->pf_copied) CU(cudaEventSynchronize(e->pf_copied));
    else CU(cudaEventCreateWithFlags(&e->pf_copied, cudaEventDisableTiming));
    if (need_ints > e->pf_pinned_ints) {
        if (e->pf_pinned) cudaFreeHost(e->pf_pinned);
        e->pf_pinned = nullptr;
        e->pf_pinned_ints = 0;
        CU(cudaMallocHost(&e->pf_pinned, need_ints * sizeof(int)));
        e->pf_pinned_ints = need_ints;
    }
    char* a = e->pf_arena;
    void *x = a + o_x, *res = a + o_res, *qkv = a + o_qkv, *q = a + o_q, *k = a + o_k, *v = a + o_v, *att = a + o_att,
         *gu = a + o_gu, *act = a + o_act, *lws = a + o_lws;
    int *ids = (int*)(a + o_ids), *pad = (int*)(a + o_pad), *cum = (int*)(a + o_cum), *lens = (int*)(a + o_len),
        *hists = (int*)(a + o_hist), *ctxs = (int*)(a + o_ctx);
    // ---- host -> device: token ids and the three length vectors (contiguous in the arena)
    memcpy(e->pf_pinned, ids_host, (size_t)Nt * sizeof(int));
    for (int b = 0; b < B; ++b) {
        e->pf_pinned[Nt + b] = len_b[b];
        e->pf_pinned[Nt + B + b] = hist_b[b];
        e->pf_pinned[Nt + 2 * B + b] = hist_b[b] + len_b[b];
        e->pf_pinned[Nt + 3 * B + b] = hist_b[b] + len_b[b] + 1;   // the reference `step` of its first decode
    }
    CU(cudaMemcpyAsync(ids, e->pf_pinned, (size_t)Nt * sizeof(int), cudaMemcpyHostToDevice, s));
    CU(cudaMemcpyAsync(lens, e->pf_pinned + Nt, (size_t)B * sizeof(int), cudaMemcpyHostToDevice, s));
    CU(cudaMemcpyAsync(hists, e->pf_pinned + Nt + B, (size_t)B * sizeof(int), cudaMemcpyHostToDevice, s));
    CU(cudaMemcpyAsync(ctxs, e->pf_pinned + Nt + 2 * B, (size_t)B * sizeof(int), cudaMemcpyHostToDevice, s));
    CU(cudaMemcpyAsync(e->seq_steps, e->pf_pinned + Nt + 3 * B, (size_t)B * sizeof(int), cudaMemcpyHostToDevice, s));
    CU(cudaEventRecord(e->pf_copied, s));
    if (!any_zero) {
        CU(cudaMemsetAsync(e->finished, 0, B, s));   // a prompt pass starts every sequence afresh
    } else {
        for (int b = 0; b < B; ++b)
            if (len_b[b] > 0) CU(cudaMemsetAsync(e->finished + b, 0, 1, s));   // ... every sequence that takes part in it
    }
    auto linear = [&](const void* in, const void* W, void* out, int N, int K) {
        return b200_linear(in, W, nullptr, out, Nt, N, K, 1, dt, lws, lin_ws + 256, s);
    };
    const bool tp = e->cfg.tp_size > 1 && (e->allreduce || e->p2p);
    RC(b200_cal_padding_offset(pad, cum, lens, B, len, s));
    if (!flash && ragged) {   // shorter sequences leave rows of the repeated K/V unwritten: 0 * stale NaN must not reach PV
        CU(cudaMemsetAsync(a + o_kr, 0, (size_t)B * e->h * ctx * e->d * es, s));
        CU(cudaMemsetAsync(a + o_vr, 0, (size_t)B * e->h * ctx * e->d * es, s));
        CU(cudaMemsetAsync(q, 0, Np * e->h * e->d * es, s));
    }
    if (!flash) RC(b200_build_causal_mask(a + o_mask, lens, ctxs, B, len, ctx, 1, dt, s));
    RC(b200_embedding(ids, e->embed, x, Nt, H, e->V, dt, s));
    for (int l = 0; l < e->cfg.num_layers; ++l) {
        LayerW& w = e->layers[l];
        const size_t layer_off = (size_t)l * e->layer_kv_elems;
        RC(b200_rmsnorm(x, res, w.attn_norm, Nt, H, e->cfg.rms_eps, dt, s));
        RC(linear(x, w.qkv, qkv, e->qkv_n, H));
        if (e->kv_fp8) {
            // fp8 cache: this pass's k, v go to an fp16 scratch slice [B, hk, len, d] (history is empty, checked above), the
            // flash kernel attends over it, and the rows enter the fp8 cache quantised once
            RC(b200_qkv_bias_rope_to_cache(q, k, v, qkv, nullptr, pad, hists, B, len, Nt, e->h, e->hk, e->d, len,
                                           e->cfg.rotary_dim, e->cfg.rotary_base, dt, s));
            RC(b200_context_attention(q, k, v, att, pad, lens, hists, B, len, Nt, e->h, e->hk, e->d, len, dt, s));
            if (e->kv_paged) {
                RC(b200_kv_quantize_fp8_paged(k, e->k_cache + layer_off, e->k_scale + (size_t)l * e->layer_kv_rows, e->page_table,
                                              e->max_pages, lens, hists, B, e->hk, len, e->d, len, s));
                RC(b200_kv_quantize_fp8_paged(v, e->v_cache + layer_off, e->v_scale + (size_t)l * e->layer_kv_rows, e->page_table,
                                              e->max_pages, lens, hists, B, e->hk, len, e->d, len, s));
            } else {
            RC(b200_kv_quantize_fp8(k, e->k_cache + layer_off, e->k_scale + (size_t)l * e->layer_kv_rows, lens, hists, B, e->hk,
                                    len, e->cfg.max_seq_len, e->d, len, s));
            RC(b200_kv_quantize_fp8(v, e->v_cache + layer_off, e->v_scale + (size_t)l * e->layer_kv_rows, lens, hists, B, e->hk,
                                    len, e->cfg.max_seq_len, e->d, len, s));
            }
        } else if (e->kv_paged) {
            // paged cache: the RoPE kernel writes the rotated k and the v rows straight into the sequences' pages behind their
            // history (launchAddFusedQKVBiasTransposeAndRoPE + launchConcatKVCache in one pass, as for the dense cache), and the
            // flash kernel attends over the pages: a page is one of its 128-key tiles
            RC(b200_qkv_bias_rope_to_pages(q, e->k_cache + layer_off * es, e->v_cache + layer_off * es, e->page_table, e->max_pages,
                                           qkv, nullptr, pad, hists, B, len, Nt, e->h, e->hk, e->d, e->cfg.rotary_dim,
                                           e->cfg.rotary_base, dt, s));
            RC(b200_context_attention_paged(q, e->k_cache + layer_off * es, e->v_cache + layer_off * es, e->page_table, e->max_pages,
                                            e->n_pages, att, pad, lens, hists, B, len, Nt, e->h, e->hk, e->d, dt, s));
        } else {
        // bias/RoPE/transpose with ConcatKV folded in: k, v land in the cache rows directly
        RC(b200_qkv_bias_rope_to_cache(q, e->k_cache + layer_off * es, e->v_cache + layer_off * es, qkv, nullptr, pad, hists,
                                       B, len, Nt, e->h, e->hk, e->d, e->cfg.max_seq_len, e->cfg.rotary_dim,
                                       e->cfg.rotary_base, dt, s));
        if (flash) {
            RC(b200_context_attention(q, e->k_cache + layer_off * es, e->v_cache + layer_off * es, att, pad, lens, hists, B,
                                      len, Nt, e->h, e->hk, e->d, e->cfg.max_seq_len, dt, s));
        } else {
            RC(b200_repeat_kv_cache(e->k_cache, e->v_cache, a + o_kr, a + o_vr, layer_off, ctxs, B, e->h, e->hk,
                                    e->cfg.max_seq_len, ctx, e->d, dt, s));
            RC(b200_batched_gemm(q, a + o_kr, a + o_sc, B * e->h, len, ctx, e->d, 1, dt, s));
            RC(b200_scale_mask_softmax(a + o_sc, a + o_mask, a + o_sc, B, e->h, len, ctx, 1.0f / sqrtf((float)e->d), dt, s));
            RC(b200_batched_gemm(a + o_sc, a + o_vr, a + o_pv, B * e->h, len, e->d, ctx, 0, dt, s));
            RC(b200_transpose_remove_padding(a + o_pv, att, pad, Nt, B, len, e->h, e->d, dt, s));
        }
        }
        RC(linear(att, w.o, x, H, e->h * e->d));
        if (tp) RC(tp_sum(e, x, (size_t)Nt * H));
        RC(b200_fused_add_bias_residual_rmsnorm(res, x, nullptr, w.ffn_norm, Nt, H, e->cfg.rms_eps, dt, s));
        int rc_glu = b200_linear_swiglu(x, w.gate_up, act, Nt, e->I, H, dt, s);   // GEMM with SwiGLU epilogue
        if (rc_glu == B200_EUNSUPPORTED) {
            RC(linear(x, w.gate_up, gu, 2 * e->I, H));
            RC(b200_swiglu(gu, act, Nt, e->I, dt, s));
        } else if (rc_glu != 0) {
            return fail("b200_linear_swiglu failed", rc_glu);
        }
        int rc_res = tp ? B200_EUNSUPPORTED : b200_linear_residual(act, w.down, res, x, Nt, H, e->I, dt, s);   // down + AddResidual
        if (rc_res == B200_EUNSUPPORTED) {
            RC(linear(act, w.down, x, H, e->I));
            if (tp) RC(tp_sum(e, x, (size_t)Nt * H));
            RC(b200_add_residual(res, x, Nt, H, dt, s));
        } else if (rc_res != 0) {
            return fail("b200_linear_residual failed", rc_res);
        }
    }
    // ---- last token of each sequence -> final norm -> LM head -> top-1 (llama.cpp:221-232, :305)
    if (!lens_host) {
        CU(cudaMemcpy2DAsync(e->hidden, (size_t)H * es, (const char*)x + (size_t)(len - 1) * H * es, (size_t)len * H * es,
                             (size_t)H * es, (size_t)B, cudaMemcpyDeviceToDevice, s));
    } else {
        for (int b = 0; b < B; ++b)
            if (len_b[b] > 0)   // (a sequence that sat the pass out keeps its stale row: its logits are not used)
                CU(cudaMemcpyAsync((char*)e->hidden + (size_t)b * H * es, (const char*)x + (size_t)(cum_b[b + 1] - 1) * H * es,
                                   (size_t)H * es, cudaMemcpyDeviceToDevice, s));
    }
    CU(cudaMemcpyAsync(e->partial, e->hidden, (size_t)B * H * es, cudaMemcpyDeviceToDevice, s));
    RC(b200_rmsnorm(e->partial, nullptr, e->final_norm, B, H, e->cfg.rms_eps, dt, s));
    RC(b200_linear(e->partial, e->lm_head, nullptr, e->logits, B, e->V, H, 1, dt, e->lin_ws, e->lin_ws_bytes, s));
    RC(b200_argmax(e->logits, e->ids_next, B, e->V, dt, s));
    CU(launch_k(set_ids_step_kernel, dim3(1), dim3(1024), 0, s, false, e->ids_cur, e->ids_next, e->step_dev, B, ctx + 1,
                any_zero ? (const int*)lens : (const int*)nullptr));
    count_launch();
    e->step = ctx;
    for (int b = 0; b < B; ++b) e->seq_cached[b] = hist_b[b] + len_b[b];
    e->ragged = ragged;
    e->prefill_launches = (int)(b200_launch_count() - launches0);
    if (ids_out_host) {
        CU(cudaMemcpyAsync(e->pinned_ids, e->ids_next, sizeof(int) * B, cudaMemcpyDeviceToHost, s));
        CU(cudaStreamSynchronize(s));
        memcpy(ids_out_host, e->pinned_ids, sizeof(int) * B);
    }
    return 0;
}
extern "C" int b200_engine_prefill_launches(const b200_engine* e) { return e ? e->prefill_launches : 0; }

extern "C" int b200_engine_prefill(b200_engine* e, const int* ids_host, int len, int* ids_out_host) {
    if (!e || !ids_host || len < 1) return fail("prefill: bad arguments");
    return prefill_impl(e, ids_host, nullptr, len, ids_out_host);
}

extern "C" int b200_engine_prefill_ragged(b200_engine* e, const int* ids_host, const int* lens_host, int* ids_out_host) {
    if (!e || !ids_host || !lens_host) return fail("prefill_ragged: bad arguments");
    return prefill_impl(e, ids_host, lens_host, 0, ids_out_host);
}

extern "C" int b200_engine_get_seq_steps(const b200_engine* e, int* steps_out_host) {
    if (!e || !steps_out_host) return B200_EINVAL;
    for (int b = 0; b < e->batch; ++b) steps_out_host[b] = e->seq_cached[b];
    return B200_OK;
}


// ------------------------------------------------------------------------------------------
// weight files: the reference's on-disk format (LlamaWeight<T>::loadWeights, llama_weights.cc:41-53;
// LlamaLayerWeight<T>::loadWeights, layer_weights.cc:48-66): one raw little-endian **fp32** file per
// tensor (loadWeightFromBin<T, float>, weight_utils.cu:91-181 converts to T), names
//   <dir>/model.norm.weight.bin  <dir>/lm_head.weight.bin  <dir>/model.embed_tokens.weight.bin
//   <dir>/model.layers.<l>.{input_layernorm,post_attention_layernorm}.weight.bin
//   <dir>/model.layers.<l>.self_attn.{qkv,o_proj}.weight.bin   <dir>/model.layers.<l>.mlp.{gate_up_proj,down_proj}.weight.bin
// qkv rows are q heads, k heads, v heads; gate_up rows gate then up (act_kernel.cu:28-29). With
// tp_size > 1 each rank keeps its head / FFN-column shard (Megatron split): rows of qkv and
// gate_up, columns of o_proj and down_proj.
// ------------------------------------------------------------------------------------------
namespace b200 {
static int read_f32(const std::string& path, std::vector<float>& buf, size_t elems) {
    FILE* f = fopen(path.c_str(), "rb");
    if (!f) return fail("load_weights: cannot open " + path);
    buf.resize(elems);
    const size_t got = fread(buf.data(), sizeof(float), elems, f);
    int extra = fgetc(f);
    fclose(f);
    if (got != elems || extra != EOF) return fail("load_weights: " + path + " does not hold the expected number of fp32 values");
    return 0;
}
// rows [r0, r0+nr) x cols [c0, c0+nc) of a row-major fp32 [rows, cols] matrix -> dst (engine dtype), appended at dst_row
static void pack_block(const std::vector<float>& src, size_t cols, size_t r0, size_t nr, size_t c0, size_t nc, int dtype,
                       std::vector<unsigned char>& dst, size_t dst_row) {
    const size_t es = dtype == B200_F16 ? 2 : 4;
    if (dst.size() < (dst_row + nr) * nc * es) dst.resize((dst_row + nr) * nc * es);
    for (size_t r = 0; r < nr; ++r) {
        const float* s = src.data() + (r0 + r) * cols + c0;
        if (dtype == B200_F16) {
            __half* d = reinterpret_cast<__half*>(dst.data()) + (dst_row + r) * nc;
            for (size_t c = 0; c < nc; ++c) d[c] = __float2half_rn(s[c]);
        } else {
            memcpy(dst.data() + (dst_row + r) * nc * 4, s, nc * 4);
        }
    }
}
}  // namespace b200

extern "C" int b200_engine_load_weights(b200_engine* e, const char* dir) {
    if (!e || !dir) return B200_EINVAL;
    const b200_model_config& c = e->cfg;
    std::string base(dir);
    if (!base.empty() && base.back() != '/') base += '/';
    const int dt = c.dtype, tp = c.tp_size, rk = c.tp_rank;
    const size_t H = e->H, d = e->d, I_all = c.inter_size, I = e->I, V = e->V;
    const size_t h_all = c.head_num, hk_all = c.kv_head_num, h = e->h, hk = e->hk;
    std::vector<float> buf;
    std::vector<unsigned char> out;
    auto upload = [&](int slot, int layer) {
        const size_t bytes = slot_elems(e, slot) * e->esz;
        if (out.size() < bytes) return fail("load_weights: internal size mismatch");
        return b200_engine_set_weight(e, slot, layer, out.data(), bytes);
    };
    auto whole = [&](const std::string& name, size_t rows, size_t cols, int slot, int layer) {
        RC(read_f32(base + name, buf, rows * cols));
        out.clear();
        pack_block(buf, cols, 0, rows, 0, cols, dt, out, 0);
        return upload(slot, layer);
    };
    RC(whole("model.norm.weight.bin", 1, H, B200_W_FINAL_NORM, 0));
    RC(whole("lm_head.weight.bin", V, H, B200_W_LM_HEAD, 0));
    RC(whole("model.embed_tokens.weight.bin", V, H, B200_W_EMBED, 0));
    for (int l = 0; l < c.num_layers; ++l) {
        const std::string lp = "model.layers." + std::to_string(l);
        RC(whole(lp + ".input_layernorm.weight.bin", 1, H, B200_W_ATTN_NORM, l));
        RC(whole(lp + ".post_attention_layernorm.weight.bin", 1, H, B200_W_FFN_NORM, l));
        // qkv [(h + 2 hk) d, H]: this rank's q heads, then its k heads, then its v heads
        RC(read_f32(base + lp + ".self_attn.qkv.weight.bin", buf, (h_all + 2 * hk_all) * d * H));
        out.clear();
        pack_block(buf, H, (size_t)rk * h * d, h * d, 0, H, dt, out, 0);
        pack_block(buf, H, (h_all + (size_t)rk * hk) * d, hk * d, 0, H, dt, out, h * d);
        pack_block(buf, H, (h_all + hk_all + (size_t)rk * hk) * d, hk * d, 0, H, dt, out, (h + hk) * d);
        RC(upload(B200_W_QKV, l));
        // o_proj [H, h_all d]: this rank's input columns
        RC(read_f32(base + lp + ".self_attn.o_proj.weight.bin", buf, H * h_all * d));
        out.clear();
        pack_block(buf, h_all * d, 0, H, (size_t)rk * h * d, h * d, dt, out, 0);
        RC(upload(B200_W_O, l));
        // gate_up [2 I_all, H]: gate rows of this rank, then its up rows
        RC(read_f32(base + lp + ".mlp.gate_up_proj.weight.bin", buf, 2 * I_all * H));
        out.clear();
        pack_block(buf, H, (size_t)rk * I, I, 0, H, dt, out, 0);
        pack_block(buf, H, I_all + (size_t)rk * I, I, 0, H, dt, out, I);
        RC(upload(B200_W_GATE_UP, l));
        // down [H, I_all]: this rank's input columns
        RC(read_f32(base + lp + ".mlp.down_proj.weight.bin", buf, H * I_all));
        out.clear();
        pack_block(buf, I_all, 0, H, (size_t)rk * I, I, dt, out, 0);
        RC(upload(B200_W_DOWN, l));
    }
    (void)tp;
    return B200_OK;
}

// ------------------------------------------------------------------------------------------
// generation loop of Llama<T>::Response (llama.cpp:363-448) without the tokenizer: prompt ids in,
// firstTokenGen (prefill) once, continueTokenGen until every sequence has produced eos_id or
// max_new_tokens are out. out_ids [batch, max_new_tokens] (sequence-major; positions after a
// sequence's EOS are filled with eos_id), n_generated [batch] (may be NULL). As in the reference
// the EOS token itself is not counted / emitted for a sequence (llama.cpp:415-419 breaks before
// printing it). Returns the number of decode steps run (>= 0) or a negative / CUDA error code.
// ------------------------------------------------------------------------------------------
static int generate_impl(b200_engine* e, const int* prompt_ids_host, const int* lens_host, int prompt_len,
                         int max_new_tokens, int eos_id, int* out_ids, int* n_generated) {
    if (!e || !prompt_ids_host || !out_ids || max_new_tokens < 1) return B200_EINVAL;
    const int B = e->batch;
    std::vector<int> cur(B), done(B, 0), count(B, 0);
    // batches: a sequence that has produced eos_id is retired on the device (no more KV traffic for it) while the others
    // go on - the reference's `finished` flags; restored to the caller's setting on the way out
    const int eos_before = e->eos_id;
    struct Restore {
        b200_engine* e; int eos;
        ~Restore() { if (e->eos_id != eos) b200_engine_set_eos(e, eos); }
    } restore{e, eos_before};
    if (B > 1 && eos_id >= 0 && e->eos_id != eos_id) b200_engine_set_eos(e, eos_id);
    int rc = lens_host ? b200_engine_prefill_ragged(e, prompt_ids_host, lens_host, cur.data())
                       : b200_engine_prefill(e, prompt_ids_host, prompt_len, cur.data());
    if (rc != 0) return rc > 0 ? -1000 - rc : rc;
    for (int i = 0; i < B * max_new_tokens; ++i) out_ids[i] = eos_id;
    int steps = 0;
    for (int t = 0; t < max_new_tokens; ++t) {
        int alive = 0;
        bool newly_done = false;
        for (int b = 0; b < B; ++b) {
            // the first token (from the prompt pass) is emitted unconditionally, as llama.cpp:410-413 does
            if (!done[b] && t > 0 && cur[b] == eos_id) { done[b] = 1; newly_done = true; }
            if (!done[b]) { out_ids[(size_t)b * max_new_tokens + t] = cur[b]; count[b] = t + 1; ++alive; }
        }
        if (!alive || t == max_new_tokens - 1) break;
        if (e->step + 1 >= e->cfg.max_seq_len) break;   // cache full
        if (newly_done && e->kv_paged) b200_engine_release_finished(e);   // retired sequences give their pages back
        rc = b200_engine_decode_step_host(e, nullptr, cur.data());   // continues from the ids the engine holds
        if (rc != 0) return rc > 0 ? -1000 - rc : rc;
        ++steps;
    }
    if (n_generated) for (int b = 0; b < B; ++b) n_generated[b] = count[b];
    return steps;
}

extern "C" int b200_engine_generate(b200_engine* e, const int* prompt_ids_host, int prompt_len, int max_new_tokens,
                                    int eos_id, int* out_ids, int* n_generated) {
    if (prompt_len < 1) return B200_EINVAL;
    return generate_impl(e, prompt_ids_host, nullptr, prompt_len, max_new_tokens, eos_id, out_ids, n_generated);
}

// prompts of different lengths: ids packed back to back, lens_host[batch] (b200_engine_prefill_ragged)
extern "C" int b200_engine_generate_ragged(b200_engine* e, const int* prompt_ids_host, const int* lens_host,
                                           int max_new_tokens, int eos_id, int* out_ids, int* n_generated) {
    if (!lens_host) return B200_EINVAL;
    return generate_impl(e, prompt_ids_host, lens_host, 0, max_new_tokens, eos_id, out_ids, n_generated);
}
