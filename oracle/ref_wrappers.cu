// oracle/ref_wrappers.cu — C-ABI shims around the REFERENCE's own code, compiled from the
// sources where they lie under /root/reference (never copied into this repo) into
// oracle/_ref/libref.so by `make -C oracle ref`.
//
// TEST INFRASTRUCTURE ONLY (same rule as oracle/llama_cpu.c).
//
// Two families are exported:
//   ref_cpu_*  the CPU reference functions that live inside tests/unittests/*.cu, reached by
//              including each test file inside its own namespace with `main` renamed;
//   ref_gpu_*  the reference's CUDA launchers (src/kernels/*.cu), a live GPU oracle for the
//              kernels that are not defective at the tested shape.
// System / reference headers are included first, outside any namespace, so the includes inside
// the test files are no-ops by the time they are seen in a namespace.
#include <algorithm>
#include <fstream>
#include <iostream>
#include <math.h>
#include <stdio.h>
#include <stdlib.h>
#include <string>
#include <thread>
#include <vector>
#include <cuda_fp16.h>
#include <cuda_runtime.h>

#include "src/kernels/act_kernel.h"
#include "src/kernels/add_residual.h"
#include "src/kernels/attn_softmax_kernel.h"
#include "src/kernels/build_causal_mask.h"
#include "src/kernels/cal_paddingoffset.h"
#include "src/kernels/concat_past_kv.h"
#include "src/kernels/fused_addresidual_norm.h"
#include "src/kernels/fused_decoder_self_attention.h"
#include "src/kernels/fused_transpose_and_remv_pad.h"
#include "src/kernels/linear.h"
#include "src/kernels/qkv_bias_and_RoPE.h"
#include "src/kernels/repeat_kv.h"
#include "src/kernels/rmsnorm_kernel.h"
#include "src/kernels/topK.h"
#include "src/utils/macro.h"
#include "src/weights/base_weights.h"
#include "src/weights/llama/attention_weights.h"

#define main ref_test_main
namespace t_rmsnorm {
#include "tests/unittests/test_rmsnorm.cu"
}
namespace t_addnorm {
#include "tests/unittests/test_fused_addresidual_norm.cu"
}
namespace t_act {
#include "tests/unittests/test_act.cu"
}
namespace t_rope {
#include "tests/unittests/test_bias_and_RoPE.cu"
}
namespace t_attn {
#include "tests/unittests/test_fused_decoder_attention.cu"
}
namespace t_linear {
#include "tests/unittests/test_linear.cu"
}
namespace t_mask {
#include "tests/unittests/test_casual_mask.cu"
}
#undef main

#define API extern "C"

// ------------------------------------------------------------------ CPU reference functions
API void ref_cpu_rmsnorm(float* out, float* scale, float eps, int hidden, int n_tokens) {
    t_rmsnorm::CPUfusedresidandRMSNorm(out, scale, eps, hidden, n_tokens);
}
API void ref_cpu_fused_add_norm(float* residual, float* out, float* bias, float* scale, float eps,
                            int hidden, int n_tokens) {
    t_addnorm::CPUfusedresidandRMSNorm(residual, out, bias, scale, eps, hidden, n_tokens);
}
API void ref_cpu_swiglu(float* in, float* out, int batch, int inter) {
    t_act::CPUSwiGLU<float>(in, out, batch, inter);
}
API void ref_cpu_rope(float* q, float* k, float* v, float* QKV, const float* bias,
                  const int* padding_offset, const int* history, const int* input_len, int batch,
                  int seq_len, int token_num, int head_num, int kv_head_num, int head_size,
                  int rot_dim, float base) {
    t_rope::CPUfunc(q, k, v, QKV, bias, padding_offset, history, input_len, batch, seq_len,
                    token_num, head_num, kv_head_num, head_size, rot_dim, base);
}
API void ref_cpu_masked_attn(float* q, float* k, float* v, float* k_cache, float* v_cache, float* out,
                         int batch, int heads, int head_size, int step) {
    t_attn::CPUMaskedAttn<float>(q, k, v, k_cache, v_cache, out, batch, heads, head_size, step);
}
API void ref_cpu_linear(float* in, float* w, float* out, int m, int k, int n) {
    t_linear::CPUlinear(in, w, out, m, k, n);
}
// The same CPUlinear over W stored as n/block consecutive column blocks [k, block] (its inner loop walks W
// with a stride of n floats: narrow blocks keep that stride inside a few cache lines), the blocks dealt
// round-robin to `threads` host threads. out [m, n] is overwritten. Used by bench.py --impl reference.
API void ref_cpu_linear_blocked(float* in, float* w_blocks, float* out, int m, int k, int n, int block, int threads) {
    const int nblk = n / block;
    if (threads < 1) threads = 1;
    auto work = [&](int t) {
        std::vector<float> tmp((size_t)m * block);
        for (int b = t; b < nblk; b += threads) {
            std::fill(tmp.begin(), tmp.end(), 0.f);   // CPUlinear accumulates into its output
            t_linear::CPUlinear(in, w_blocks + (size_t)b * k * block, tmp.data(), m, k, block);
            for (int i = 0; i < m; ++i)
                for (int j = 0; j < block; ++j) out[(size_t)i * n + (size_t)b * block + j] = tmp[(size_t)i * block + j];
        }
    };
    std::vector<std::thread> th;
    for (int t = 1; t < threads; ++t) th.emplace_back(work, t);
    work(0);
    for (auto& x : th) x.join();
}
API void ref_cpu_causal_mask(float* mask, const int* q_lens, const int* k_lens, int max_q, int max_k,
                         int batch) {
    t_mask::CPUbuildCasualMask(mask, q_lens, k_lens, max_q, max_k, batch);
}

// ------------------------------------------------------------------ GPU reference launchers
// all pointers are device pointers; every call synchronises and returns the CUDA error code.
static int done() { return (int)cudaDeviceSynchronize() | (int)cudaGetLastError(); }

API int ref_gpu_rmsnorm_f32(float* x, float* resid, float* gamma, int nt, int hidden, float eps) {
    TensorWrapper<float> t(Device::GPU, FP32, {nt, hidden}, x), r(Device::GPU, FP32, {nt, hidden}, resid);
    LayerNormWeight<float> w; w.gamma = gamma;
    launchRMSNorm(&t, &r, w, eps);
    return done();
}
API int ref_gpu_rmsnorm_f16(half* x, half* resid, half* gamma, int nt, int hidden, float eps) {
    TensorWrapper<half> t(Device::GPU, FP16, {nt, hidden}, x), r(Device::GPU, FP16, {nt, hidden}, resid);
    LayerNormWeight<half> w; w.gamma = gamma;
    launchRMSNorm(&t, &r, w, eps);
    return done();
}
API int ref_gpu_fused_add_norm_f32(float* resid, float* out, float* bias, float* gamma, int nt, int hidden, float eps) {
    TensorWrapper<float> r(Device::GPU, FP32, {nt, hidden}, resid), o(Device::GPU, FP32, {nt, hidden}, out);
    BaseWeight<float> n; n.bias = bias;
    launchFusedAddBiasResidualRMSNorm(&r, &o, n, gamma, eps);
    return done();
}
API int ref_gpu_act_f32(float* in, float* out, int nt, int inter) {
    TensorWrapper<float> i(Device::GPU, FP32, {nt, 2, inter}, in), o(Device::GPU, FP32, {nt, inter}, out);
    launchAct(&i, &o);
    return done();
}
API int ref_gpu_add_residual_f32(float* resid, float* out, int nt, int hidden) {
    TensorWrapper<float> r(Device::GPU, FP32, {nt, hidden}, resid), o(Device::GPU, FP32, {nt, hidden}, out);
    launchAddResidual(&r, &o, false);
    return done();
}
template <typename T>
static int concat_t(T* ks, T* vs, T* kd, T* vd, int* cur, int* hist, int layers, int layer, int b,
                    int hk, int s, int max_seq, int d) {
    DataType dt = getTensorType<T>();
    TensorWrapper<T> k1(Device::GPU, dt, {b, hk, s, d}, ks), v1(Device::GPU, dt, {b, hk, s, d}, vs);
    TensorWrapper<T> k2(Device::GPU, dt, {layers, b, hk, max_seq, d}, kd), v2(Device::GPU, dt, {layers, b, hk, max_seq, d}, vd);
    TensorWrapper<int> lid(Device::CPU, INT32, {1}, &layer);
    TensorWrapper<int> c(Device::GPU, INT32, {b}, cur), h(Device::GPU, INT32, {b}, hist);
    launchConcatKVCache(&k1, &v1, &lid, &c, &h, &k2, &v2);
    return done();
}
API int ref_gpu_concat_f32(float* ks, float* vs, float* kd, float* vd, int* cur, int* hist, int layers, int layer, int b, int hk, int s, int max_seq, int d) {
    return concat_t<float>(ks, vs, kd, vd, cur, hist, layers, layer, b, hk, s, max_seq, d);
}
API int ref_gpu_concat_f16(half* ks, half* vs, half* kd, half* vd, int* cur, int* hist, int layers, int layer, int b, int hk, int s, int max_seq, int d) {
    return concat_t<half>(ks, vs, kd, vd, cur, hist, layers, layer, b, hk, s, max_seq, d);
}
template <typename T>
static int repeat_t(T* ks, T* vs, T* kd, T* vd, int* ctx, int layers, int layer, int b, int h,
                    int hk, int max_seq, int max_k, int d) {
    DataType dt = getTensorType<T>();
    TensorWrapper<T> k1(Device::GPU, dt, {layers, b, hk, max_seq, d}, ks), v1(Device::GPU, dt, {layers, b, hk, max_seq, d}, vs);
    TensorWrapper<T> k2(Device::GPU, dt, {b, h, max_k, d}, kd), v2(Device::GPU, dt, {b, h, max_k, d}, vd);
    TensorWrapper<int> lid(Device::CPU, INT32, {1}, &layer), c(Device::GPU, INT32, {b}, ctx);
    launchRepeatKVCache(&k1, &v1, &c, &lid, &k2, &v2);
    return done();
}
API int ref_gpu_repeat_f32(float* ks, float* vs, float* kd, float* vd, int* ctx, int layers, int layer, int b, int h, int hk, int max_seq, int max_k, int d) {
    return repeat_t<float>(ks, vs, kd, vd, ctx, layers, layer, b, h, hk, max_seq, max_k, d);
}
API int ref_gpu_repeat_f16(half* ks, half* vs, half* kd, half* vd, int* ctx, int layers, int layer, int b, int h, int hk, int max_seq, int max_k, int d) {
    return repeat_t<half>(ks, vs, kd, vd, ctx, layers, layer, b, h, hk, max_seq, max_k, d);
}
API int ref_gpu_transpose_f32(float* src, int* pad, float* dst, int nt, int b, int h, int s, int d) {
    TensorWrapper<float> i(Device::GPU, FP32, {b, h, s, d}, src), o(Device::GPU, FP32, {nt, h, d}, dst);
    TensorWrapper<int> p(Device::GPU, INT32, {nt}, pad);
    launchTransposeOutRemovePadding(&i, &p, &o);
    return done();
}
API int ref_gpu_padding_offset(int* pad, int* cum, int* lens, int b, int max_q) {
    TensorWrapper<int> p(Device::GPU, INT32, {b, max_q}, pad), c(Device::GPU, INT32, {b + 1}, cum), l(Device::GPU, INT32, {b}, lens);
    launchCalPaddingoffset(&p, &c, &l);
    return done();
}
API int ref_gpu_causal_mask_f32(float* mask, int* ql, int* kl, int b, int mq, int mk) {
    TensorWrapper<float> m(Device::GPU, FP32, {b, mq, mk}, mask);
    TensorWrapper<int> q(Device::GPU, INT32, {b}, ql), k(Device::GPU, INT32, {b}, kl);
    launchBuildCausalMasks<float>(&m, &q, &k);
    return done();
}
API int ref_gpu_linear_f32(float* x, float* w, float* y, int m, int n, int k, int trans_b) {
    static cublasHandle_t h = nullptr;
    static cublasLtHandle_t lt = nullptr;
    static cublasWrapper* cw = nullptr;
    if (!cw) {
        cublasCreate(&h);
        cublasLtCreate(&lt);
        cw = new cublasWrapper(h, lt);
        cw->setFP32GemmConfig();
    }
    TensorWrapper<float> in(Device::GPU, FP32, {m, k}, x), out(Device::GPU, FP32, {m, n}, y);
    BaseWeight<float> W;
    W.data = w;
    W.shape = trans_b ? std::vector<int>{n, k} : std::vector<int>{k, n};
    launchLinearGemm(&in, W, &out, cw, false, trans_b != 0);
    return done();
}
API int ref_gpu_rope_decode_f32(float* qkv, int b, int qkv_heads, int d, int step, int rot, float base) {
    TensorWrapper<float> t(Device::GPU, FP32, {b, qkv_heads, d}, qkv);
    TensorWrapper<int> s(Device::CPU, INT32, {1}, &step);
    LLaMAAttentionStaticParams p{rot, base, 2048, false};
    launchRoPE(&t, &s, p);
    return done();
}
API int ref_gpu_softmax_f32(float* qk, float* mask, float* score, int b, int h, int q, int k, float scale) {
    TensorWrapper<float> a(Device::GPU, FP32, {b, h, q, k}, qk), m(Device::GPU, FP32, {b, q, k}, mask), o(Device::GPU, FP32, {b, h, q, k}, score);
    launchScaleMaskAndSoftmax(&a, &m, &o, scale);
    return done();
}
// launchTopKforBeamSearch (src/kernels/topK.cu:147): K = 5, 8 blocks per row, probs [rows, vocab].
// Its heap starts at val = 1e-20 / id = -1, so it is only meaningful on positive data.
API int ref_gpu_topk_f32(float* probs, int* tmp_ids, float* tmp_vals, int* final_ids, float* final_vals,
                         int rows, int vocab) {
    TensorWrapper<float> p(Device::GPU, FP32, {rows, vocab}, probs);
    TensorWrapper<int> ti(Device::GPU, INT32, {rows, 1, 8, 5}, tmp_ids), fi(Device::GPU, INT32, {rows, 1, 5}, final_ids);
    TensorWrapper<float> tv(Device::GPU, FP32, {rows, 1, 8, 5}, tmp_vals), fv(Device::GPU, FP32, {rows, 1, 5}, final_vals);
    launchTopKforBeamSearch(&p, &ti, &tv, &fi, &fv);
    return done();
}
