"""The C-ABI library loads on a CPU-only host and exports every symbol include/b200_llama.h
declares; the ctypes prototype table covers exactly the same set. No compute calls here."""
import ctypes
import os
import re

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def declared_symbols():
    names = set()
    for hdr in ("b200_llama.h", "b200_engine.h"):
        text = open(os.path.join(ROOT, "include", hdr)).read()
        text = re.sub(r"/\*.*?\*/", "", text, flags=re.S)
        names |= set(re.findall(r"\b(b200_\w+)\s*\(", text))
    names.discard("b200_allreduce_fn")
    return sorted(names)


def test_header_declares_the_hot_path():
    names = declared_symbols()
    for must in ["b200_rmsnorm", "b200_fused_add_bias_residual_rmsnorm", "b200_add_residual",
                 "b200_swiglu", "b200_rope_decode", "b200_qkv_bias_rope_transpose",
                 "b200_concat_kv_cache", "b200_repeat_kv_cache", "b200_transpose_remove_padding",
                 "b200_cal_padding_offset", "b200_build_causal_mask", "b200_scale_mask_softmax",
                 "b200_linear", "b200_batched_gemm", "b200_decode_attention", "b200_gemv_fused"]:
        assert must in names


def test_library_exports_every_declared_symbol(pkg):
    path = pkg.lib_path()
    if not os.path.exists(path):
        pytest.fail(f"{path} not built: run __graft_entry__.build()")
    handle = ctypes.CDLL(path)
    missing = [n for n in declared_symbols() if not hasattr(handle, n)]
    assert not missing, f"declared in include/b200_llama.h but not exported: {missing}"
    assert handle.b200_abi_version() >= 1


def test_prototype_table_matches_header(pkg):
    assert sorted(list(pkg.capi.PROTOTYPES) + list(pkg.engine.ENGINE_PROTOTYPES)) == declared_symbols()


def test_ops_fail_loudly_without_cuda(pkg):
    import torch
    if torch.cuda.is_available():
        pytest.skip("CUDA present")
    x = torch.zeros(2, 8)
    with pytest.raises(RuntimeError):
        pkg.ops.launchRMSNorm(x, None, torch.ones(8), 1e-5)
