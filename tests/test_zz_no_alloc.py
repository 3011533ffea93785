"""b200_decode_attention_static (the allocation-free entry point: split-KV scratch in a file-scope
device array, include/b200_llama.h) against b200_decode_attention with a caller-owned workspace,
through the C ABI, on two calls in a row of different geometry (the second call sees the first
call's leftovers in the shared array). Same kernels, same split count -> the outputs must be
identical; the fp16 / fp32 tolerances of the other parity tests are asserted as the stated bar.
Collected last on purpose."""
import ctypes as C

import numpy as np
import pytest
import torch

from conftest import assert_close

pytestmark = pytest.mark.gpu


def _vp(t):
    return C.c_void_p(t.data_ptr())


def _case(lib, dtype, B, h, hk, d, max_seq, step, seed):
    g = torch.Generator(device="cpu").manual_seed(seed)
    qkv = (torch.rand(B, h + 2 * hk, d, generator=g) * 2 - 1).to(dtype).cuda()
    kc = (torch.rand(B, hk, max_seq, d, generator=g) * 2 - 1).to(dtype).cuda()
    vc = (torch.rand(B, hk, max_seq, d, generator=g) * 2 - 1).to(dtype).cuda()
    code = 1 if dtype == torch.float16 else 0
    stream = C.c_void_p(torch.cuda.current_stream().cuda_stream)

    # existing entry point, caller-owned workspace
    k1, v1 = kc.clone(), vc.clone()
    want = torch.zeros(B, h * d, dtype=dtype, device="cuda")
    nbytes = lib.b200_decode_attention_workspace_bytes(B, h, d, max_seq)
    ws = torch.empty(nbytes, dtype=torch.uint8, device="cuda")
    rc = lib.b200_decode_attention(_vp(qkv), _vp(k1), _vp(v1), _vp(want), _vp(ws), nbytes, B, h, hk, d,
                                   max_seq, step, 1, code, stream)
    assert rc == 0, f"b200_decode_attention rc={rc}"

    # new entry point: no workspace argument at all
    k2, v2 = kc.clone(), vc.clone()
    got = torch.full((B, h * d), float("nan"), dtype=dtype, device="cuda")
    rc = lib.b200_decode_attention_static(_vp(qkv), _vp(k2), _vp(v2), _vp(got), B, h, hk, d, max_seq, step,
                                          1, code, stream)
    assert rc == 0, f"b200_decode_attention_static rc={rc}"
    torch.cuda.synchronize()
    kind = "fp16" if dtype == torch.float16 else "fp32"
    what = f"static vs workspace B={B} h={h} hk={hk} d={d} step={step} {kind}"
    assert_close(got.float().cpu().numpy(), want.float().cpu().numpy(), kind, what)
    assert torch.equal(got, want), what + ": same kernels and split count must agree bit for bit"
    assert torch.equal(k1, k2) and torch.equal(v1, v2), what + ": appended K/V rows differ"


def test_static_scratch_matches_workspace_entry_on_two_calls_in_a_row(pkg):
    lib = pkg.capi.lib()
    torch.cuda.set_device(0)
    # call 1: 7B head geometry at cfg2's context (4 splits per head); call 2: GQA 8x, odd length,
    # another split count and partial layout over the same array; call 3: generic path (score rows)
    _case(lib, torch.float16, 1, 32, 32, 128, 640, 513, seed=1)
    _case(lib, torch.float16, 2, 64, 8, 128, 1100, 1031, seed=2)
    _case(lib, torch.float32, 2, 6, 3, 20, 40, 33, seed=3)
    # and the first geometry again after the array has held other layouts
    _case(lib, torch.float16, 1, 32, 32, 128, 640, 513, seed=4)


def test_static_scratch_rejects_shapes_beyond_its_maximum(pkg):
    lib = pkg.capi.lib()
    torch.cuda.set_device(0)
    # generic path: batch * heads * max_seq_len fp32 score rows = 9.6 MB > the 8 MiB array.
    # The call is refused before anything is launched, so small stand-in buffers are enough.
    B, h, hk, d, max_seq = 4, 6, 3, 20, 100000
    t = torch.zeros(B * (h + 2 * hk) * d, device="cuda")
    out = torch.zeros(B * h * d, device="cuda")
    rc = lib.b200_decode_attention_static(_vp(t), _vp(t), _vp(t), _vp(out), B, h, hk, d, max_seq, 7, 0, 0,
                                          None)
    assert rc == -3, f"expected B200_EWORKSPACE (-3), got {rc}"
    assert float(out.abs().sum()) == 0.0
