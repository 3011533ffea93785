"""Torch-tensor front end over the C ABI, named after the reference's ``launch*`` entry points
(src/kernels/*.h). Shapes are read from the tensors exactly as the reference launchers read
them from ``TensorWrapper::shape``; ``step`` / ``layer_id`` are host ints (the reference keeps
them in CPU TensorWrappers). Every function launches on torch's current stream and raises
``RuntimeError`` (reference: ``std::runtime_error`` from LLM_CHECK, src/utils/macro.h:113-133) on
a bad shape or a CUDA error. Tensors must be CUDA + contiguous: there is no CPU path.
"""
import ctypes as C

import torch

from . import capi
from .capi import F16, F32


def _dt(t):
    if t.dtype == torch.float32:
        return F32
    if t.dtype == torch.float16:
        return F16
    raise RuntimeError(f"[b200][ERROR] unsupported dtype {t.dtype} (float32 / float16 only)")


def _p(t):
    if t is None:
        return None
    if not t.is_cuda:
        raise RuntimeError("[b200][ERROR] tensor is not on a CUDA device (no CPU fallback)")
    if not t.is_contiguous():
        raise RuntimeError("[b200][ERROR] tensor must be contiguous")
    return C.c_void_p(t.data_ptr())


def _stream():
    return C.c_void_p(torch.cuda.current_stream().cuda_stream)


def _check(rc, what):
    if rc != 0:
        kind = {-1: "invalid argument", -2: "unsupported shape", -3: "workspace too small"}.get(
            rc, f"CUDA error {rc}")
        raise RuntimeError(f"[b200][ERROR] {what}: {kind} Assertion fail")


def launchRMSNorm(decoder_out, decoder_residual, gamma, eps, is_last=False):
    """rmsnorm_kernel.h:12-17. decoder_out [Nt,H] in place; decoder_residual [Nt,H] or None."""
    nt, h = decoder_out.shape
    _check(capi.lib().b200_rmsnorm(_p(decoder_out), _p(decoder_residual), _p(gamma), nt, h, eps,
                                   _dt(decoder_out), _stream()), "launchRMSNorm")


def launchFusedAddBiasResidualRMSNorm(residual, decoder_out, bias, scale, eps):
    """fused_addresidual_norm.h. bias = norm.bias (may be None), scale = gamma."""
    nt, h = decoder_out.shape
    _check(capi.lib().b200_fused_add_bias_residual_rmsnorm(
        _p(residual), _p(decoder_out), _p(bias), _p(scale), nt, h, eps, _dt(decoder_out),
        _stream()), "launchFusedAddBiasResidualRMSNorm")


def launchAddResidual(residual, decoder_out, is_print=False):
    nt, h = decoder_out.shape
    _check(capi.lib().b200_add_residual(_p(residual), _p(decoder_out), nt, h, _dt(decoder_out),
                                        _stream()), "launchAddResidual")


def launchAct(inp, out):
    """act_kernel.h: inp [Nt,2,I] -> out [Nt,I]."""
    if inp.dim() != 3 or inp.shape[1] != 2:
        raise RuntimeError("[b200][ERROR] launchAct: input->shape[1] == 2 Assertion fail")
    _check(capi.lib().b200_swiglu(_p(inp), _p(out), inp.shape[0], inp.shape[2], _dt(inp),
                                  _stream()), "launchAct")


def launchRoPE(qkv_buf, step, rotary_embedding_dim, rotary_embedding_base, kv_head_num=None):
    """qkv_bias_and_RoPE.h launchRoPE: qkv_buf [B, h+2hk, d]. The reference infers
    h = shape[1]/3 (MHA only, qkv_bias_and_RoPE.cu:416); pass kv_head_num for GQA."""
    b, qh, d = qkv_buf.shape
    if kv_head_num is None:
        kv_head_num = qh // 3
    _check(capi.lib().b200_rope_decode(_p(qkv_buf), b, qh - 2 * kv_head_num, kv_head_num, d,
                                       int(step), rotary_embedding_dim, rotary_embedding_base,
                                       _dt(qkv_buf), _stream()), "launchRoPE")


def launchAddFusedQKVBiasTransposeAndRoPE(q_buf, k_buf, v_buf, QKV, qkv_bias, padding_offset,
                                          history_length, input_length, rotary_embedding_dim,
                                          rotary_embedding_base):
    token_num, _, d = QKV.shape
    b, h, s, _ = q_buf.shape
    hk = k_buf.shape[1]
    _check(capi.lib().b200_qkv_bias_rope_transpose(
        _p(q_buf), _p(k_buf), _p(v_buf), _p(QKV), _p(qkv_bias), _p(padding_offset),
        _p(history_length), _p(input_length), b, s, token_num, h, hk, d, rotary_embedding_dim,
        rotary_embedding_base, _dt(QKV), _stream()), "launchAddFusedQKVBiasTransposeAndRoPE")


def launchConcatKVCache(k_src, v_src, layer_id, cur_query_length, history_length, k_dst, v_dst):
    """k_src [B,hk,S,d]; k_dst [layers,B,hk,max_seq,d]."""
    b, hk, s, d = k_src.shape
    max_seq = k_dst.shape[3]
    off = int(layer_id) * b * hk * max_seq * d
    _check(capi.lib().b200_concat_kv_cache(
        _p(k_src), _p(v_src), _p(k_dst), _p(v_dst), off, _p(cur_query_length), _p(history_length),
        b, hk, s, max_seq, d, _dt(k_src), _stream()), "launchConcatKVCache")


def launchRepeatKVCache(k_cache_src, v_cache_src, context_length, layer_id, k_cache_dst,
                        v_cache_dst):
    b = context_length.shape[0]
    hk, max_seq = k_cache_src.shape[2], k_cache_src.shape[3]
    h, max_k, d = k_cache_dst.shape[1], k_cache_dst.shape[2], k_cache_dst.shape[3]
    off = int(layer_id) * b * hk * max_seq * d
    _check(capi.lib().b200_repeat_kv_cache(
        _p(k_cache_src), _p(v_cache_src), _p(k_cache_dst), _p(v_cache_dst), off,
        _p(context_length), b, h, hk, max_seq, max_k, d, _dt(k_cache_src), _stream()),
        "launchRepeatKVCache")


def launchTransposeOutRemovePadding(qkv_buf_w_pad, padding_offset, qkv_buf_wo_pad_1):
    b, h, s, d = qkv_buf_w_pad.shape
    nt = qkv_buf_wo_pad_1.shape[0]
    _check(capi.lib().b200_transpose_remove_padding(
        _p(qkv_buf_w_pad), _p(qkv_buf_wo_pad_1), _p(padding_offset), nt, b, s, h, d,
        _dt(qkv_buf_w_pad), _stream()), "launchTransposeOutRemovePadding")


def launchCalPaddingoffset(padding_offset, cum_seqlens, input_lengths):
    b, max_q = padding_offset.shape
    if input_lengths.shape[0] != b:
        raise RuntimeError("[b200][ERROR] input lengths numbers should equal to padding offset bs dim Assertion fail")
    if cum_seqlens.shape[0] - 1 != b:
        raise RuntimeError("[b200][ERROR] cum seqlen numbers should equal to padding offset bs dim + 1! Assertion fail")
    _check(capi.lib().b200_cal_padding_offset(_p(padding_offset), _p(cum_seqlens),
                                              _p(input_lengths), b, max_q, _stream()),
           "launchCalPaddingoffset")


def launchBuildCausalMasks(mask, q_lens, k_lens, history_visible=True):
    b, mq, mk = mask.shape
    _check(capi.lib().b200_build_causal_mask(_p(mask), _p(q_lens), _p(k_lens), b, mq, mk,
                                             1 if history_visible else 0, _dt(mask), _stream()),
           "launchBuildCausalMasks")


def launchScaleMaskAndSoftmax(qk, mask, attn_score, scale):
    b, h, ql, kl = qk.shape
    _check(capi.lib().b200_scale_mask_softmax(_p(qk), _p(mask), _p(attn_score), b, h, ql, kl,
                                              scale, _dt(qk), _stream()),
           "launchScaleMaskAndSoftmax")


def launchLinearGemm(inp, weight, output, trans_a=False, trans_b=False, bias=None):
    """linear.h launchLinearGemm (the cublasWrapper argument is dropped). weight [N,K] with
    trans_b, [K,N] without. 3-D input/output are flattened as the reference does (linear.cu:57-58)."""
    if trans_a:
        raise RuntimeError("[b200][ERROR] launchLinearGemm: trans_a is not used on the decoder path")
    m = inp.shape[0]
    k = inp.numel() // m
    n = output.numel() // output.shape[0]
    wk = weight.shape[1] if trans_b else weight.shape[0]
    wn = weight.shape[0] if trans_b else weight.shape[1]
    if wk != k or wn != n:
        raise RuntimeError("[b200][ERROR] 2nd dim of input MUST = 1st dim of weight Assertion fail")
    lib = capi.lib()
    nbytes = lib.b200_linear_workspace_bytes(m, n, k, 1 if trans_b else 0, _dt(inp))
    ws = _workspace(inp.device, nbytes, "linear") if nbytes else None
    _check(lib.b200_linear(_p(inp), _p(weight), _p(bias), _p(output), m, n, k, 1 if trans_b else 0,
                           _dt(inp), _p(ws), nbytes, _stream()), "launchLinearGemm")


def launchLinearStridedBatchGemm(input1, input2, output, trans_a=False, trans_b=False):
    """linear.h: input1 [B,h,m,k]; input2 [B,h,n,k] (trans_b) or [B,h,k,n]; output [B,h,m,n]."""
    if trans_a:
        raise RuntimeError("[b200][ERROR] launchLinearStridedBatchGemm: trans_a unsupported")
    bc = input1.shape[0] * input1.shape[1]
    m, k = input1.shape[2], input1.shape[3]
    n = input2.shape[2] if trans_b else input2.shape[3]
    k2 = input2.shape[3] if trans_b else input2.shape[2]
    if k2 != k:
        raise RuntimeError("[b200][ERROR] 2nd dim of input MUST = 1st dim of weight Assertion fail")
    _check(capi.lib().b200_batched_gemm(_p(input1), _p(input2), _p(output), bc, m, n, k,
                                        1 if trans_b else 0, _dt(input1), _stream()),
           "launchLinearStridedBatchGemm")


_ws = {}


def _workspace(device, nbytes, what):
    ws = _ws.get((device, what))
    if ws is None or ws.numel() < nbytes:
        ws = torch.empty(nbytes, dtype=torch.uint8, device=device)
        _ws[(device, what)] = ws
    return ws


def _attn_workspace(device, nbytes):
    return _workspace(device, nbytes, "attn")


def launchDecoderMaskedMHA(qkv_buf, layer_id, k_cache, v_cache, step, mha_output, append_kv=True,
                           step_dev=None):
    """fused_decoder_self_attention.h. qkv_buf [B,h+2hk,d]; caches [layers,B,hk,max_seq,d];
    mha_output [B,h*d]. (qkv bias / finished / static params of the reference signature are
    unused by the reference kernel, fused_decoder_self_attention.cu:341-344.)"""
    b, qh, d = qkv_buf.shape
    hk, max_seq = k_cache.shape[2], k_cache.shape[3]
    h = qh - 2 * hk
    lib = capi.lib()
    nbytes = lib.b200_decode_attention_workspace_bytes(b, h, d, max_seq)
    ws = _attn_workspace(qkv_buf.device, nbytes)
    kl, vl = k_cache[int(layer_id)], v_cache[int(layer_id)]
    _check(lib.b200_decode_attention_ex(
        _p(qkv_buf), _p(kl), _p(vl), _p(mha_output), _p(ws), ws.numel(), b, h, hk, d, max_seq,
        int(step), _p(step_dev), 1 if append_kv else 0, _dt(qkv_buf), _stream()),
        "launchDecoderMaskedMHA")


def gemv_fused(x, W, y, flags=0, bias=None, gamma=None, eps=0.0, N=None, k_cache_layer=None,
               v_cache_layer=None, head_num=0, kv_head_num=0, head_size=0, step=0, rotary_dim=0,
               rotary_base=10000.0, step_dev=None, rope_den=None, p2p=None):
    """b200_gemv_fused: decode-time linear with fused prologue / epilogue (engine path).
    p2p: a b200_p2p* handle (int / c_void_p): the outputs are pushed to the ranks' receive buffers instead of y."""
    a = capi.GemvArgs()
    a.x, a.W, a.y = x.data_ptr(), W.data_ptr(), y.data_ptr()
    a.bias = bias.data_ptr() if bias is not None else None
    a.gamma = gamma.data_ptr() if gamma is not None else None
    a.eps = eps
    a.n_tokens = x.shape[0]
    a.K = W.shape[1]
    a.N = N if N is not None else (W.shape[0] // 2 if flags & capi.GEMV_SWIGLU else W.shape[0])
    a.flags = flags
    if k_cache_layer is not None:
        a.k_cache_layer, a.v_cache_layer = k_cache_layer.data_ptr(), v_cache_layer.data_ptr()
        a.max_seq_len = k_cache_layer.shape[2]
    a.head_num, a.kv_head_num, a.head_size = head_num, kv_head_num, head_size
    a.step, a.rotary_dim, a.rotary_base = int(step), rotary_dim, rotary_base
    a.step_dev = step_dev.data_ptr() if step_dev is not None else None
    a.rope_den = rope_den.data_ptr() if rope_den is not None else None
    a.p2p = p2p
    _check(capi.lib().b200_gemv_fused(C.byref(a), _dt(x), _stream()), "b200_gemv_fused")


def launchInputEmbedding(input_ids, output, embed_table):
    nt, h = output.shape
    _check(capi.lib().b200_embedding(_p(input_ids), _p(embed_table), _p(output), nt, h,
                                     embed_table.shape[0], _dt(output), _stream()),
           "launchInputEmbedding")


def launchTopK(probs, topk_ids, topk_vals, k):
    b, v = probs.shape
    _check(capi.lib().b200_topk(_p(probs), _p(topk_ids), _p(topk_vals), b, v, k, _dt(probs),
                                _stream()), "launchTopKforBeamSearch")


def argmax(logits, out_ids):
    b, v = logits.shape
    _check(capi.lib().b200_argmax(_p(logits), _p(out_ids), b, v, _dt(logits), _stream()), "argmax")


def context_attention(q, k_cache_layer, v_cache_layer, out, padding_offset, input_len, history_len):
    """b200_context_attention: the tcgen05 flash kernel behind LLaMAContextAttentionLayer's fused path
    (replaces context_attention.cpp:136-161). q [B, h, S, d]; cache layer slices [B, hk, max_seq, d];
    out [num_tokens, h*d]; int32 device vectors as in the reference. Raises on unsupported shapes
    (fp32 / head size != 128): the caller then runs the reference-order sequence on the GPU."""
    b, h, s, d = q.shape
    hk, max_seq = k_cache_layer.shape[1], k_cache_layer.shape[2]
    _check(capi.lib().b200_context_attention(
        _p(q), _p(k_cache_layer), _p(v_cache_layer), _p(out), _p(padding_offset), _p(input_len),
        _p(history_len), b, s, out.shape[0], h, hk, d, max_seq, _dt(q), _stream()),
        "b200_context_attention")


def linear_swiglu(inp, weight, out):
    """b200_linear_swiglu: out [Nt, I] = silu(inp Wg^T) * (inp Wu^T), weight [2I, K] gate rows first
    (ffn.cpp:72-80 in one tcgen05 launch). Raises "unsupported shape" outside fp16 prefill shapes."""
    nt, k = inp.shape
    inter = out.shape[1]
    if weight.shape[0] != 2 * inter or weight.shape[1] != k:
        raise RuntimeError("[b200][ERROR] linear_swiglu: weight must be [2*inter, K] Assertion fail")
    _check(capi.lib().b200_linear_swiglu(_p(inp), _p(weight), _p(out), nt, inter, k, _dt(inp), _stream()),
           "b200_linear_swiglu")


def linear_residual(inp, weight, residual, out):
    """b200_linear_residual: out = residual + inp W^T (launchLinearGemm + launchAddResidual in one tcgen05 launch)"""
    nt, k = inp.shape
    n = out.shape[1]
    _check(capi.lib().b200_linear_residual(_p(inp), _p(weight), _p(residual), _p(out), nt, n, k, _dt(inp), _stream()),
           "b200_linear_residual")


def rope_decode_ragged(qkv_buf, seq_steps, rotary_embedding_dim, rotary_embedding_base, kv_head_num):
    """launchRoPE with one `step` per sequence (b200_rope_decode_ragged): seq_steps int32 DEVICE [B]"""
    b, qh, d = qkv_buf.shape
    _check(capi.lib().b200_rope_decode_ragged(_p(qkv_buf), b, qh - 2 * kv_head_num, kv_head_num, d, _p(seq_steps),
                                              rotary_embedding_dim, rotary_embedding_base, None, _dt(qkv_buf),
                                              _stream()), "rope_decode_ragged")


def decode_attention_ragged(qkv_buf, layer_id, k_cache, v_cache, seq_steps, max_step, mha_output, append_kv=True):
    """launchDecoderMaskedMHA with one `step` per sequence (b200_decode_attention_ragged)"""
    b, qh, d = qkv_buf.shape
    hk, max_seq = k_cache.shape[2], k_cache.shape[3]
    h = qh - 2 * hk
    lib = capi.lib()
    ws = _attn_workspace(qkv_buf.device, lib.b200_decode_attention_workspace_bytes(b, h, d, max_seq))
    kl, vl = k_cache[int(layer_id)], v_cache[int(layer_id)]
    _check(lib.b200_decode_attention_ragged(
        _p(qkv_buf), _p(kl), _p(vl), _p(mha_output), _p(ws), ws.numel(), b, h, hk, d, max_seq, int(max_step),
        _p(seq_steps), 1 if append_kv else 0, _dt(qkv_buf), _stream()), "decode_attention_ragged")
