"""ctypes loader + prototypes for ``include/b200_llama.h`` (the drop-in C ABI)."""
import ctypes as C
import os

F32, F16 = 0, 1
_HERE = os.path.dirname(os.path.abspath(__file__))


def lib_path():
    return os.path.join(_HERE, "lib", "libb200llama.so")


class GemvArgs(C.Structure):
    """mirror of ``b200_gemv_args``"""
    _fields_ = [
        ("x", C.c_void_p), ("W", C.c_void_p), ("bias", C.c_void_p), ("y", C.c_void_p),
        ("gamma", C.c_void_p), ("eps", C.c_float),
        ("n_tokens", C.c_int), ("N", C.c_int), ("K", C.c_int), ("flags", C.c_uint),
        ("k_cache_layer", C.c_void_p), ("v_cache_layer", C.c_void_p),
        ("head_num", C.c_int), ("kv_head_num", C.c_int), ("head_size", C.c_int),
        ("max_seq_len", C.c_int), ("step", C.c_int), ("rotary_dim", C.c_int),
        ("rotary_base", C.c_float), ("step_dev", C.c_void_p), ("rope_den", C.c_void_p), ("p2p", C.c_void_p),
        ("step_stride", C.c_int),
    ]


GEMV_NORM, GEMV_RESIDUAL, GEMV_SWIGLU, GEMV_ROPE_KV, GEMV_NORM_GFIRST = 1, 2, 4, 8, 16

_vp, _i, _f, _sz = C.c_void_p, C.c_int, C.c_float, C.c_size_t

# name -> (restype, argtypes); every symbol include/b200_llama.h declares is listed here and
# tests/test_capi_symbols.py checks the two stay in sync.
PROTOTYPES = {
    "b200_abi_version": (_i, []),
    "b200_build_info": (C.c_char_p, []),
    "b200_launch_count": (C.c_uint64, []),
    "b200_launch_count_reset": (None, []),
    "b200_rmsnorm": (_i, [_vp, _vp, _vp, _i, _i, _f, _i, _vp]),
    "b200_fused_add_bias_residual_rmsnorm": (_i, [_vp, _vp, _vp, _vp, _i, _i, _f, _i, _vp]),
    "b200_add_residual": (_i, [_vp, _vp, _i, _i, _i, _vp]),
    "b200_swiglu": (_i, [_vp, _vp, _i, _i, _i, _vp]),
    "b200_rope_decode": (_i, [_vp, _i, _i, _i, _i, _i, _i, _f, _i, _vp]),
    "b200_rope_decode_ex": (_i, [_vp, _i, _i, _i, _i, _i, _vp, _i, _f, _i, _vp]),
    "b200_qkv_bias_rope_transpose": (_i, [_vp] * 8 + [_i] * 7 + [_f, _i, _vp]),
    "b200_concat_kv_cache": (_i, [_vp, _vp, _vp, _vp, _sz, _vp, _vp, _i, _i, _i, _i, _i, _i, _vp]),
    "b200_repeat_kv_cache": (_i, [_vp, _vp, _vp, _vp, _sz, _vp, _i, _i, _i, _i, _i, _i, _i, _vp]),
    "b200_transpose_remove_padding": (_i, [_vp, _vp, _vp, _i, _i, _i, _i, _i, _i, _vp]),
    "b200_cal_padding_offset": (_i, [_vp, _vp, _vp, _i, _i, _vp]),
    "b200_build_causal_mask": (_i, [_vp, _vp, _vp, _i, _i, _i, _i, _i, _vp]),
    "b200_scale_mask_softmax": (_i, [_vp, _vp, _vp, _i, _i, _i, _i, _f, _i, _vp]),
    "b200_linear_workspace_bytes": (_sz, [_i, _i, _i, _i, _i]),
    "b200_linear": (_i, [_vp, _vp, _vp, _vp, _i, _i, _i, _i, _i, _vp, _sz, _vp]),
    "b200_gemv_fused": (_i, [C.POINTER(GemvArgs), _i, _vp]),
    "b200_rope_table": (_i, [_vp, _i, _f, _vp]),
    "b200_batched_gemm": (_i, [_vp, _vp, _vp, _i, _i, _i, _i, _i, _i, _vp]),
    "b200_decode_attention_workspace_bytes": (_sz, [_i, _i, _i, _i]),
    "b200_decode_attention": (_i, [_vp, _vp, _vp, _vp, _vp, _sz] + [_i] * 8 + [_vp]),
    "b200_decode_attention_ex": (_i, [_vp, _vp, _vp, _vp, _vp, _sz] + [_i] * 6 + [_vp, _i, _i, _vp]),
    "b200_decode_attention_ragged": (_i, [_vp, _vp, _vp, _vp, _vp, _sz] + [_i] * 6 + [_vp, _i, _i, _vp]),
    "b200_decode_attention_static": (_i, [_vp, _vp, _vp, _vp] + [_i] * 8 + [_vp]),
    "b200_decode_attention_set_finished": (None, [_vp]),
    "b200_kv_append_fp8": (_i, [_vp] * 5 + [_i] * 6 + [_vp, _i, _vp]),
    "b200_decode_attention_fp8": (_i, [_vp] * 7 + [_sz] + [_i] * 6 + [_vp, _i, _vp]),
    "b200_kv_quantize_fp8": (_i, [_vp] * 5 + [_i] * 6 + [_vp]),
    "b200_kv_dequant_fp8": (_i, [_vp, _vp, _vp, _sz, _i, _vp]),
    "b200_decode_attention_paged": (_i, [_vp] * 4 + [_i] + [_vp, _vp, _sz] + [_i] * 5 + [_vp, _i, _i, _vp]),
    "b200_qkv_bias_rope_to_pages": (_i, [_vp] * 4 + [_i] + [_vp] * 4 + [_i] * 7 + [C.c_float, _i, _vp]),
    "b200_context_attention_paged": (_i, [_vp] * 4 + [_i, _i] + [_vp] * 4 + [_i] * 7 + [_vp]),
    "b200_kv_scatter_pages": (_i, [_vp] * 3 + [_i] + [_vp, _vp] + [_i] * 6 + [_vp]),
    "b200_kv_gather_pages": (_i, [_vp] * 3 + [_i] + [_vp] + [_i] * 6 + [_vp]),
    "b200_kv_append_fp8_paged": (_i, [_vp] * 6 + [_i] * 6 + [_vp, _i, _vp]),
    "b200_decode_attention_fp8_paged": (_i, [_vp] * 6 + [_i] + [_vp, _vp, _sz] + [_i] * 5 + [_vp, _i, _vp]),
    "b200_kv_quantize_fp8_paged": (_i, [_vp] * 4 + [_i] + [_vp, _vp] + [_i] * 5 + [_vp]),
    "b200_kv_dequant_fp8_paged": (_i, [_vp] * 4 + [_i] + [_vp] + [_i] * 5 + [_vp]),
    "b200_rope_decode_ragged": (_i, [_vp, _i, _i, _i, _i, _vp, _i, C.c_float, _vp, _i, _vp]),
    "b200_context_attention": (_i, [_vp] * 7 + [_i] * 8 + [_vp]),
    "b200_embedding": (_i, [_vp, _vp, _vp, _i, _i, _i, _i, _vp]),
    "b200_argmax": (_i, [_vp, _vp, _i, _i, _i, _vp]),
    "b200_context_attention_set_trace": (None, [_vp]),
    "b200_rope_decode_tab": (_i, [_vp, _i, _i, _i, _i, _i, _vp, _i, C.c_float, _vp, _i, _vp]),
    "b200_qkv_bias_rope_to_cache": (_i, [_vp] * 7 + [_i] * 8 + [C.c_float, _i, _vp]),
    "b200_linear_residual": (_i, [_vp, _vp, _vp, _vp, _i, _i, _i, _i, _vp]),
    "b200_linear_swiglu": (_i, [_vp, _vp, _vp, _i, _i, _i, _i, _vp]),
    "b200_topk": (_i, [_vp, _vp, _vp, _i, _i, _i, _i, _vp]),
}

_lib = None


def lib():
    """Load (once) and return the CDLL. Raises if the CUDA library has not been built."""
    global _lib
    if _lib is None:
        path = lib_path()
        if not os.path.exists(path):
            raise RuntimeError(
                f"[b200][ERROR] {path} is missing: build it with `python -c 'import "
                "__graft_entry__ as g; g.build()'` (there is no CPU fallback)")
        handle = C.CDLL(path)
        for name, (res, args) in PROTOTYPES.items():
            fn = getattr(handle, name)
            fn.restype = res
            fn.argtypes = args
        _lib = handle
    return _lib
