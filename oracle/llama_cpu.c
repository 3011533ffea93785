/*
 * oracle/llama_cpu.c — CPU restatement of the reference's LLaMA decoder kernels.
 *
 * TEST INFRASTRUCTURE ONLY. Nothing under llm-inference_b200/ may import, link or call this;
 * only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference legs do.
 *
 * Each function follows the reference CUDA kernel (file:line cited, paths relative to the
 * reference repo root) with the SURVEY.md §2.3 defects D1-D22 resolved as listed there.
 * Data are fp32 arrays (fp16 inputs are widened exactly by the caller); sums are accumulated
 * in double so the oracle is at least as accurate as either implementation it arbitrates.
 * Pinned against the reference's own unit-test CPU functions and golden vectors by
 * tests/test_oracle_*.py (see oracle/ref_wrappers.cu for the compiled reference side).
 *
 * Build: gcc -O2 -shared -fPIC -o liboracle.so llama_cpu.c -lm -lpthread   (oracle/Makefile)
 * (no OpenMP runtime in this image: the one heavy loop, oracle_linear, uses pthreads)
 */
#include <math.h>
#include <stddef.h>
#include <stdint.h>
#include <string.h>

#define IDX size_t

#include <pthread.h>
#include <stdlib.h>
#include <unistd.h>

/* minimal parallel-for over [0, n): fn(ctx, begin, end) on ORACLE_THREADS (default: all cores) */
typedef void (*range_fn)(void* ctx, int begin, int end);
typedef struct { range_fn fn; void* ctx; int begin, end; } range_job;
static void* range_tramp(void* p) {
    range_job* j = (range_job*)p;
    j->fn(j->ctx, j->begin, j->end);
    return NULL;
}
int oracle_num_threads(void) {
    const char* e = getenv("ORACLE_THREADS");
    long n = e ? atol(e) : sysconf(_SC_NPROCESSORS_ONLN);
    if (n < 1) n = 1;
    if (n > 256) n = 256;
    return (int)n;
}
static void parallel_for(int n, range_fn fn, void* ctx) {
    int nt = oracle_num_threads();
    if (nt > n) nt = n;
    if (nt <= 1) { fn(ctx, 0, n); return; }
    pthread_t th[256];
    range_job jobs[256];
    for (int i = 0; i < nt; ++i) {
        jobs[i].fn = fn; jobs[i].ctx = ctx;
        jobs[i].begin = (int)((long long)n * i / nt);
        jobs[i].end = (int)((long long)n * (i + 1) / nt);
        pthread_create(&th[i], NULL, range_tramp, &jobs[i]);
    }
    for (int i = 0; i < nt; ++i) pthread_join(th[i], NULL);
}

/* ---- RMSNorm: src/kernels/rmsnorm_kernel.cu:63-115; tests/unittests/test_rmsnorm.cu:33-51.
 * x <- (x * rsqrt(mean(x^2)+eps)) * gamma; residual_out <- pre-norm x (D7: really saved). */
void oracle_rmsnorm(float* x, float* residual_out, const float* gamma, int n_tokens, int hidden,
                    float eps) {
    for (int t = 0; t < n_tokens; ++t) {
        float* row = x + (IDX)t * hidden;
        double ss = 0.0;
        for (int i = 0; i < hidden; ++i) ss += (double)row[i] * row[i];
        const float inv = (float)(1.0 / sqrt(ss / hidden + (double)eps));
        for (int i = 0; i < hidden; ++i) {
            if (residual_out) residual_out[(IDX)t * hidden + i] = row[i];
            row[i] = (row[i] * inv) * gamma[i];
        }
    }
}

/* ---- fused add-bias-residual + RMSNorm: src/kernels/fused_addresidual_norm.cu:62-128.
 * residual <- residual + out (+bias); out <- (gamma * residual) * inv_rms.  (D8: true sum) */
void oracle_fused_add_bias_residual_rmsnorm(float* residual, float* out, const float* bias,
                                            const float* gamma, int n_tokens, int hidden,
                                            float eps) {
    for (int t = 0; t < n_tokens; ++t) {
        float* r = residual + (IDX)t * hidden;
        float* o = out + (IDX)t * hidden;
        double ss = 0.0;
        for (int i = 0; i < hidden; ++i) {
            float v = o[i] + (bias ? bias[i] : 0.f);
            v = v + r[i];
            r[i] = v;
            ss += (double)v * v;
        }
        const float inv = (float)(1.0 / sqrt(ss / hidden + (double)eps));
        for (int i = 0; i < hidden; ++i) o[i] = (gamma[i] * r[i]) * inv;
    }
}

/* ---- AddResidual: src/kernels/add_residual.cu:9-31 (D10).  out <- out + residual */
void oracle_add_residual(const float* residual, float* out, int n_tokens, int hidden) {
    const IDX n = (IDX)n_tokens * hidden;
    for (IDX i = 0; i < n; ++i) out[i] += residual[i];
}

/* ---- SwiGLU: src/kernels/act_kernel.cu:7-31; tests/unittests/test_act.cu:14-25.
 * gate_up [n, 2, inter] -> out [n, inter] = silu(gate) * up, silu(x) = x / (1 + exp(-x)) */
void oracle_swiglu(const float* gate_up, float* out, int n_tokens, int inter) {
    for (int t = 0; t < n_tokens; ++t)
        for (int i = 0; i < inter; ++i) {
            const float g = gate_up[(IDX)t * 2 * inter + i];
            const float u = gate_up[(IDX)t * 2 * inter + inter + i];
            out[(IDX)t * inter + i] = (g / (1.0f + expf(-g))) * u;
        }
}

/* ---- RoPE helpers: src/kernels/qkv_bias_and_RoPE.cu:29-46 */
static void rope_cs(int pair, int rot_dim, float base, float pos, float* c, float* s) {
    /* power evaluated in double, rounded once (= correctly rounded powf); see common.cuh */
    const float ex = (float)(2 * pair) / (float)rot_dim;
    const float den = (float)pow((double)base, (double)ex);
    const float ang = pos / den;
    *c = cosf(ang);
    *s = sinf(ang);
}

/* ---- decode RoPE: src/kernels/qkv_bias_and_RoPE.cu:323-380, launcher :410-447.
 * qkv token-major [batch, h+2hk, d] (D5); q and k heads rotated at position step-1. */
void oracle_rope_decode(float* qkv, int batch, int head_num, int kv_head_num, int head_size,
                        int step, int rot_dim, float base) {
    const int qkv_heads = head_num + 2 * kv_head_num, hd = head_size / 2;
    for (int b = 0; b < batch; ++b)
        for (int head = 0; head < head_num + kv_head_num; ++head) {
            float* p = qkv + ((IDX)b * qkv_heads + head) * head_size;
            for (int i = 0; i < rot_dim / 2; ++i) {
                float c, s;
                rope_cs(i, rot_dim, base, (float)(step - 1), &c, &s);
                const float a = p[i], bb = p[i + hd];
                p[i] = c * a - s * bb;
                p[i + hd] = c * bb + s * a;
            }
        }
}

/* ---- prefill bias + RoPE + transposition: src/kernels/qkv_bias_and_RoPE.cu:50-145, :248-290
 * (D12: strides h+2hk, v copied, timestep = history_len[b] + local token id).
 * qkv [token_num, h+2hk, d] -> q [B,h,S,d], k,v [B,hk,S,d]; padded rows untouched. */
void oracle_qkv_bias_rope_transpose(float* q, float* k, float* v, const float* qkv,
                                    const float* bias, const int* padding_offset,
                                    const int* history_len, int batch, int seq_len, int token_num,
                                    int head_num, int kv_head_num, int head_size, int rot_dim,
                                    float base) {
    (void)batch;
    const int qkv_heads = head_num + 2 * kv_head_num, hd = head_size / 2;
    for (int t = 0; t < token_num; ++t) {
        const int dst = t + padding_offset[t];
        const int b = dst / seq_len, local = dst % seq_len;
        const float pos = (float)(history_len[b] + local);
        for (int head = 0; head < qkv_heads; ++head) {
            const float* src = qkv + ((IDX)t * qkv_heads + head) * head_size;
            const float* bs = bias ? bias + (IDX)head * head_size : NULL;
            float* d;
            if (head < head_num)
                d = q + (((IDX)b * head_num + head) * seq_len + local) * head_size;
            else if (head < head_num + kv_head_num)
                d = k + (((IDX)b * kv_head_num + (head - head_num)) * seq_len + local) * head_size;
            else
                d = v + (((IDX)b * kv_head_num + (head - head_num - kv_head_num)) * seq_len + local) * head_size;
            for (int i = 0; i < head_size; ++i) d[i] = src[i] + (bs ? bs[i] : 0.f);
            if (head < head_num + kv_head_num) {
                for (int i = 0; i < rot_dim / 2; ++i) {
                    float c, s;
                    rope_cs(i, rot_dim, base, pos, &c, &s);
                    const float a = d[i], bb = d[i + hd];
                    d[i] = c * a - s * bb;
                    d[i + hd] = c * bb + s * a;
                }
            }
        }
    }
}

/* ---- ConcatKVCache: src/kernels/concat_past_kv.cu:17-56, :103-144. Pure copy: works on raw
 * bytes (elem_size = 2 or 4) so fp16 parity is bit-exact too.
 * src [B, hk, max_q_len, d] -> cache + layer_offset: [B, hk, max_seq_len, d] */
void oracle_concat_kv(const void* src, void* cache, size_t layer_offset, const int* cur_len,
                      const int* history_len, int batch, int kv_heads, int max_q_len,
                      int max_seq_len, int head_size, int elem_size) {
    const IDX row = (IDX)head_size * elem_size;
    const char* s = (const char*)src;
    char* d = (char*)cache + layer_offset * elem_size;
    for (int b = 0; b < batch; ++b)
        for (int h = 0; h < kv_heads; ++h)
            for (int t = 0; t < max_q_len; ++t)
                if (t < cur_len[b])
                    memcpy(d + (((IDX)b * kv_heads + h) * max_seq_len + history_len[b] + t) * row,
                           s + (((IDX)b * kv_heads + h) * max_q_len + t) * row, row);
}

/* ---- RepeatKVCache: src/kernels/repeat_kv.cu:27-70, :76-120.
 * cache + layer_offset [B, hk, max_seq_len, d] -> dst [B, h, max_k_len, d], rows < context_len */
void oracle_repeat_kv(const void* cache, void* dst, size_t layer_offset, const int* context_len,
                      int batch, int head_num, int kv_heads, int max_seq_len, int max_k_len,
                      int head_size, int elem_size) {
    const IDX row = (IDX)head_size * elem_size;
    const char* s = (const char*)cache + layer_offset * elem_size;
    char* d = (char*)dst;
    const int rep = head_num / kv_heads;
    for (int b = 0; b < batch; ++b)
        for (int h = 0; h < head_num; ++h)
            for (int t = 0; t < max_k_len; ++t)
                if (t < context_len[b])
                    memcpy(d + (((IDX)b * head_num + h) * max_k_len + t) * row,
                           s + (((IDX)b * kv_heads + h / rep) * max_seq_len + t) * row, row);
}

/* ---- TransposeOutRemovePadding: src/kernels/fused_transpose_and_remv_pad.cu:18-47.
 * src [B, h, S, d] -> dst [token_num, h*d] */
void oracle_transpose_remove_padding(const void* src, void* dst, const int* padding_offset,
                                     int token_num, int seq_len, int head_num, int head_size,
                                     int elem_size) {
    const IDX row = (IDX)head_size * elem_size;
    for (int t = 0; t < token_num; ++t) {
        const int p = t + padding_offset[t];
        const int b = p / seq_len, local = p % seq_len;
        for (int h = 0; h < head_num; ++h)
            memcpy((char*)dst + ((IDX)t * head_num + h) * row,
                   (const char*)src + (((IDX)b * head_num + h) * seq_len + local) * row, row);
    }
}

/* ---- CalPaddingoffset: src/kernels/cal_paddingoffset.cu:51-72 (same serial loop).
 * Golden vector: tests/unittests/test_cal_paddingoffset.cu:54-56. */
void oracle_cal_padding_offset(int* padding_offset, int* cum_seqlens, const int* input_lens,
                               int batch, int max_q_len) {
    int ind = 0, cum_offset = 0, total = 0;
    for (int b = 0; b < batch; ++b) {
        const int len = input_lens[b];
        cum_seqlens[b] = total;
        for (int i = 0; i < len; ++i) padding_offset[ind++] = cum_offset;
        cum_offset += max_q_len - len;
        total += len;
    }
    cum_seqlens[batch] = total;
}

/* ---- BuildCausalMasks: src/kernels/build_causal_mask.cu:23-44 (history_visible = 0) and
 * tests/unittests/test_casual_mask.cu:12-32 (history_visible = 1); identical when klen == qlen (D17) */
void oracle_build_causal_mask(float* mask, const int* q_lens, const int* k_lens, int batch,
                              int max_q_len, int max_k_len, int history_visible) {
    for (int b = 0; b < batch; ++b)
        for (int q = 0; q < max_q_len; ++q)
            for (int k = 0; k < max_k_len; ++k) {
                const int ql = q_lens[b], kl = k_lens[b];
                int one = q < ql && k < kl && k <= q + (kl - ql);
                if (!history_visible) one = one && (k >= kl - ql);
                mask[((IDX)b * max_q_len + q) * max_k_len + k] = one ? 1.f : 0.f;
            }
}

/* ---- ScaleMaskAndSoftmax: src/kernels/attn_softmax_kernel.cu:109-173 (D16: max from -inf).
 * score = exp(s - max) / (sum + 1e-6), s = scale*qk + (1-mask)*(-10000) */
void oracle_scale_mask_softmax(const float* qk, const float* mask, float* score, int batch,
                               int head_num, int q_len, int k_len, float scale) {
    for (int b = 0; b < batch; ++b)
        for (int h = 0; h < head_num; ++h)
            for (int q = 0; q < q_len; ++q) {
                const float* in = qk + (((IDX)b * head_num + h) * q_len + q) * k_len;
                const float* m = mask + ((IDX)b * q_len + q) * k_len;
                float* o = score + (((IDX)b * head_num + h) * q_len + q) * k_len;
                float mx = -INFINITY;
                for (int k = 0; k < k_len; ++k) {
                    const float s = scale * in[k] + (1.f - m[k]) * (-10000.0f);
                    if (s > mx) mx = s;
                }
                double sum = 0.0;
                for (int k = 0; k < k_len; ++k)
                    sum += exp((double)(scale * in[k] + (1.f - m[k]) * (-10000.0f)) - (double)mx);
                const double inv = 1.0 / (sum + 1e-6);
                for (int k = 0; k < k_len; ++k)
                    o[k] = (float)(exp((double)(scale * in[k] + (1.f - m[k]) * (-10000.0f)) - (double)mx) * inv);
            }
}

/* ---- Linear: src/kernels/linear.cu:39-104 -> cublasGemmEx (cublas_utils.cc:32-71);
 * CPU pin tests/unittests/test_linear.cu:13-22 (trans_b = 0, W [K,N]).
 * y[M,N] = x[M,K] * (trans_b ? W[N,K]^T : W[K,N]) (+ bias) */
typedef struct { const float* x; const float* W; const float* bias; float* y; int M, N, K, trans_b; } linear_ctx;
static void linear_range(void* vp, int n0, int n1) {
    linear_ctx* c = (linear_ctx*)vp;
    const int M = c->M, N = c->N, K = c->K;
    for (int n = n0; n < n1; ++n)
        for (int m = 0; m < M; ++m) {
            double acc = 0.0;
            if (c->trans_b)
                for (int k = 0; k < K; ++k) acc += (double)c->x[(IDX)m * K + k] * c->W[(IDX)n * K + k];
            else
                for (int k = 0; k < K; ++k) acc += (double)c->x[(IDX)m * K + k] * c->W[(IDX)k * N + n];
            c->y[(IDX)m * N + n] = (float)acc + (c->bias ? c->bias[n] : 0.f);
        }
}
void oracle_linear(const float* x, const float* W, const float* bias, float* y, int M, int N,
                   int K, int trans_b) {
    linear_ctx c = {x, W, bias, y, M, N, K, trans_b};
    if ((double)M * N * K < 1e6) linear_range(&c, 0, N);
    else parallel_for(N, linear_range, &c);
}

/* ---- strided batched GEMM: src/kernels/linear.cu:127-229 (D14: correct C stride) */
void oracle_batched_gemm(const float* A, const float* B, float* C, int batch, int m, int n, int k,
                         int trans_b) {
    for (int b = 0; b < batch; ++b)
        oracle_linear(A + (IDX)b * m * k, B + (IDX)b * n * k, NULL, C + (IDX)b * m * n, m, n, k,
                      trans_b);
}

/* ---- decode attention: src/kernels/fused_decoder_self_attention.cu:80-291 with D1-D5:
 * qkv token-major [B, h+2hk, d]; cache layer slice [B, hk, max_seq, d]; kv head = head/(h/hk);
 * append_kv: write k,v at row step-1 first (:170-172, :274-276); scale = rsqrt(d) (:122);
 * softmax max-subtracted, denominator + 1e-6 (:203-221); out [B, h*d] (:272-287). */
void oracle_decode_attention(const float* qkv, float* k_cache, float* v_cache, float* out,
                             int batch, int head_num, int kv_head_num, int head_size,
                             int max_seq_len, int step, int append_kv) {
    const int qkv_heads = head_num + 2 * kv_head_num, rep = head_num / kv_head_num;
    const int d = head_size;
    if (append_kv)
        for (int b = 0; b < batch; ++b)
            for (int kh = 0; kh < kv_head_num; ++kh) {
                const float* tok = qkv + (IDX)b * qkv_heads * d;
                memcpy(k_cache + (((IDX)b * kv_head_num + kh) * max_seq_len + step - 1) * d,
                       tok + (IDX)(head_num + kh) * d, sizeof(float) * d);
                memcpy(v_cache + (((IDX)b * kv_head_num + kh) * max_seq_len + step - 1) * d,
                       tok + (IDX)(head_num + kv_head_num + kh) * d, sizeof(float) * d);
            }
    const double scale = 1.0 / sqrt((double)d);
    for (int b = 0; b < batch; ++b)
        for (int h = 0; h < head_num; ++h) {
            const float* q = qkv + ((IDX)b * qkv_heads + h) * d;
            const float* kc = k_cache + ((IDX)b * kv_head_num + h / rep) * max_seq_len * d;
            const float* vc = v_cache + ((IDX)b * kv_head_num + h / rep) * max_seq_len * d;
            double mx = -INFINITY;
            double sc[step > 0 ? step : 1];
            for (int j = 0; j < step; ++j) {
                double acc = 0.0;
                for (int e = 0; e < d; ++e) acc += (double)q[e] * kc[(IDX)j * d + e];
                sc[j] = acc * scale;
                if (sc[j] > mx) mx = sc[j];
            }
            double sum = 0.0;
            for (int j = 0; j < step; ++j) { sc[j] = exp(sc[j] - mx); sum += sc[j]; }
            const double inv = 1.0 / (sum + 1e-6);
            for (int e = 0; e < d; ++e) {
                double acc = 0.0;
                for (int j = 0; j < step; ++j) acc += sc[j] * vc[(IDX)j * d + e];
                out[((IDX)b * head_num + h) * d + e] = (float)(acc * inv);
            }
        }
}

/* ---- embedding: src/kernels/input_embedding.cu:5-28; tests/unittests/test_input_embedding.cu:35-41 */
void oracle_embedding(const int* ids, const float* table, float* out, int n_tokens, int hidden) {
    for (int t = 0; t < n_tokens; ++t)
        memcpy(out + (IDX)t * hidden, table + (IDX)ids[t] * hidden, sizeof(float) * hidden);
}

/* ---- top-K: src/kernels/topK.cu:25-190 (values descending; equal values -> lower index) */
void oracle_topk(const float* logits, int* ids, float* vals, int batch, int vocab, int k) {
    for (int b = 0; b < batch; ++b) {
        const float* row = logits + (IDX)b * vocab;
        for (int r = 0; r < k; ++r) {
            int best = -1;
            for (int j = 0; j < vocab; ++j) {
                int taken = 0;
                for (int q = 0; q < r; ++q) taken |= (ids[(IDX)b * k + q] == j);
                if (taken) continue;
                if (best < 0 || row[j] > row[best]) best = j;
            }
            ids[(IDX)b * k + r] = best;
            if (vals) vals[(IDX)b * k + r] = row[best];
        }
    }
}
