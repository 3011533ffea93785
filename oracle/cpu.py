"""numpy front end over oracle/liboracle.so (the C restatement in oracle/llama_cpu.c).

TEST INFRASTRUCTURE ONLY — imported by tests/, __graft_entry__.smoke() and bench.py's CPU
baseline legs; never by the product package. Functions mirror the names / argument order of
llm-inference_b200/ops.py so parity tests read symmetrically. All float data are float32
numpy arrays (fp16 test inputs are widened first: exact), copies take raw arrays of any dtype.
"""
import ctypes as C
import os
import subprocess

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
_lib = None


def build():
    subprocess.check_call(["make", "-s", "-C", _HERE, "liboracle.so"])


def lib():
    global _lib
    if _lib is None:
        path = os.path.join(_HERE, "liboracle.so")
        src = os.path.join(_HERE, "llama_cpu.c")
        if not os.path.exists(path) or os.path.getmtime(path) < os.path.getmtime(src):
            build()
        _lib = C.CDLL(path)
        _lib.oracle_num_threads.restype = C.c_int
    return _lib


def num_threads():
    return lib().oracle_num_threads()


def _f(a):
    assert a.dtype == np.float32 and a.flags["C_CONTIGUOUS"], (a.dtype, a.flags)
    return a.ctypes.data_as(C.c_void_p)


def _i(a):
    assert a.dtype == np.int32 and a.flags["C_CONTIGUOUS"]
    return a.ctypes.data_as(C.c_void_p)


def _opt(a):
    return None if a is None else _f(a)


def rmsnorm(x, residual_out, gamma, eps):
    nt, h = x.shape
    lib().oracle_rmsnorm(_f(x), _opt(residual_out), _f(gamma), nt, h, C.c_float(eps))


def fused_add_bias_residual_rmsnorm(residual, out, bias, gamma, eps):
    nt, h = out.shape
    lib().oracle_fused_add_bias_residual_rmsnorm(_f(residual), _f(out), _opt(bias), _f(gamma), nt,
                                                 h, C.c_float(eps))


def add_residual(residual, out):
    nt, h = out.shape
    lib().oracle_add_residual(_f(residual), _f(out), nt, h)


def swiglu(gate_up, out):
    lib().oracle_swiglu(_f(gate_up), _f(out), gate_up.shape[0], gate_up.shape[2])


def rope_decode(qkv, head_num, kv_head_num, step, rot_dim, base):
    b, _, d = qkv.shape
    lib().oracle_rope_decode(_f(qkv), b, head_num, kv_head_num, d, int(step), rot_dim,
                             C.c_float(base))


def qkv_bias_rope_transpose(q, k, v, qkv, bias, padding_offset, history_len, rot_dim, base):
    token_num, _, d = qkv.shape
    b, h, s, _ = q.shape
    hk = k.shape[1]
    lib().oracle_qkv_bias_rope_transpose(_f(q), _f(k), _f(v), _f(qkv), _opt(bias),
                                         _i(padding_offset), _i(history_len), b, s, token_num, h,
                                         hk, d, rot_dim, C.c_float(base))


def _raw(a):
    assert a.flags["C_CONTIGUOUS"]
    return a.ctypes.data_as(C.c_void_p)


def concat_kv(src, cache, layer_id, cur_len, history_len):
    b, hk, s, d = src.shape
    max_seq = cache.shape[3]
    off = int(layer_id) * b * hk * max_seq * d
    lib().oracle_concat_kv(_raw(src), _raw(cache), C.c_size_t(off), _i(cur_len), _i(history_len),
                           b, hk, s, max_seq, d, src.itemsize)


def repeat_kv(cache, dst, layer_id, context_len):
    b = context_len.shape[0]
    hk, max_seq = cache.shape[2], cache.shape[3]
    h, max_k, d = dst.shape[1], dst.shape[2], dst.shape[3]
    off = int(layer_id) * b * hk * max_seq * d
    lib().oracle_repeat_kv(_raw(cache), _raw(dst), C.c_size_t(off), _i(context_len), b, h, hk,
                           max_seq, max_k, d, cache.itemsize)


def transpose_remove_padding(src, dst, padding_offset):
    b, h, s, d = src.shape
    lib().oracle_transpose_remove_padding(_raw(src), _raw(dst), _i(padding_offset), dst.shape[0], s,
                                          h, d, src.itemsize)


def cal_padding_offset(padding_offset, cum_seqlens, input_lens):
    b, max_q = padding_offset.shape
    lib().oracle_cal_padding_offset(_i(padding_offset), _i(cum_seqlens), _i(input_lens), b, max_q)


def build_causal_mask(mask, q_lens, k_lens, history_visible=True):
    b, mq, mk = mask.shape
    lib().oracle_build_causal_mask(_f(mask), _i(q_lens), _i(k_lens), b, mq, mk,
                                   1 if history_visible else 0)


def scale_mask_softmax(qk, mask, score, scale):
    b, h, ql, kl = qk.shape
    lib().oracle_scale_mask_softmax(_f(qk), _f(mask), _f(score), b, h, ql, kl, C.c_float(scale))


def linear(x, W, y, trans_b=True, bias=None):
    m = x.shape[0]
    k = x.size // m
    n = y.size // y.shape[0]
    lib().oracle_linear(_f(x), _f(W), _opt(bias), _f(y), m, n, k, 1 if trans_b else 0)


def batched_gemm(A, B, Cm, trans_b):
    bc = A.shape[0] * A.shape[1]
    m, k = A.shape[2], A.shape[3]
    n = B.shape[2] if trans_b else B.shape[3]
    lib().oracle_batched_gemm(_f(A), _f(B), _f(Cm), bc, m, n, k, 1 if trans_b else 0)


def decode_attention(qkv, k_cache_layer, v_cache_layer, out, head_num, kv_head_num, step,
                     append_kv=True):
    b, _, d = qkv.shape
    max_seq = k_cache_layer.shape[2]
    lib().oracle_decode_attention(_f(qkv), _f(k_cache_layer), _f(v_cache_layer), _f(out), b,
                                  head_num, kv_head_num, d, max_seq, int(step),
                                  1 if append_kv else 0)


def embedding(ids, table, out):
    lib().oracle_embedding(_i(ids), _f(table), _f(out), out.shape[0], out.shape[1])


def topk(logits, ids, vals, k):
    b, v = logits.shape
    lib().oracle_topk(_f(logits), _i(ids), _opt(vals), b, v, k)
