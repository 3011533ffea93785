"""GPU parity against the REFERENCE'S OWN CUDA kernels (oracle/_ref/libref.so = src/kernels/*.cu
compiled for sm_100a from where they lie under /root/reference, `make -C oracle ref`), on the
kernels that are sound at the tested shape: RMSNorm output, SwiGLU fp32, AddResidual
fp32, fused add+norm with one warp per row (D8), ConcatKV, RepeatKV, transpose-remove-pad,
padding offset, causal mask at klen==qlen (D17), cuBLAS fp32 linear, decode RoPE fp32 at B=1 (D5),
scale-mask-softmax with S_k <= 32 (D8). Copies / masks / offsets bit-exact, floats rel <= 1e-5."""
import ctypes as C
import os

import numpy as np
import pytest
import torch

from conftest import assert_close

pytestmark = pytest.mark.gpu
REF = os.path.join(os.path.dirname(__file__), "..", "oracle", "_ref", "libref.so")


@pytest.fixture(scope="module")
def ref():
    if not os.path.exists(REF):
        pytest.skip("oracle/_ref/libref.so not built (needs /root/reference at build time)")
    return C.CDLL(REF)


@pytest.fixture(scope="module")
def ops(pkg):
    assert torch.cuda.is_available()
    return pkg.ops


def P(t):
    return C.c_void_p(t.data_ptr())


def rnd(shape, seed, lo=-1.0, hi=1.0, dtype=torch.float32):
    g = np.random.default_rng(seed).uniform(lo, hi, shape).astype(np.float32)
    return torch.from_numpy(g).to(dtype).cuda()


def ints(vals):
    return torch.tensor(vals, dtype=torch.int32, device="cuda")


@pytest.mark.parametrize("nt,h", [(64, 4096), (3, 4096), (1, 2048)])
def test_rmsnorm_vs_reference_kernel(ref, ops, nt, h):
    # the reference launches H/4 threads per block (rmsnorm_kernel.cu:187): H <= 4096 only
    x, g = rnd((nt, h), 1), rnd((h,), 2, 0.5, 1.5)
    want, scratch = x.clone(), torch.zeros_like(x)
    assert ref.ref_gpu_rmsnorm_f32(P(want), P(scratch), P(g), nt, h, C.c_float(1e-5)) == 0
    got = x.clone()
    ops.launchRMSNorm(got, torch.zeros_like(x), g, 1e-5)
    assert_close(got.cpu().numpy(), want.cpu().numpy(), "fp32", "rmsnorm vs reference kernel")


def test_rmsnorm_f16_vs_reference_kernel(ref, ops):
    nt, h = 16, 4096
    x, g = rnd((nt, h), 3, dtype=torch.float16), rnd((h,), 4, 0.5, 1.5, torch.float16)
    want, scratch = x.clone(), torch.zeros_like(x)
    assert ref.ref_gpu_rmsnorm_f16(P(want), P(scratch), P(g), nt, h, C.c_float(1e-5)) == 0
    got = x.clone()
    ops.launchRMSNorm(got, torch.zeros_like(x), g, 1e-5)
    assert_close(got.float().cpu().numpy(), want.float().cpu().numpy(), "fp16", "rmsnorm f16")


def test_fused_add_norm_vs_reference_kernel_one_warp_rows(ref, ops):
    nt, h = 7, 128   # 32 threads x float4: the shape at which the reference block reduce is sound (D8)
    res, out, bias, g = rnd((nt, h), 5), rnd((nt, h), 6), rnd((h,), 7), rnd((h,), 8, 0.5, 1.5)
    r1, o1 = res.clone(), out.clone()
    assert ref.ref_gpu_fused_add_norm_f32(P(r1), P(o1), P(bias), P(g), nt, h, C.c_float(1e-5)) == 0
    r2, o2 = res.clone(), out.clone()
    ops.launchFusedAddBiasResidualRMSNorm(r2, o2, bias, g, 1e-5)
    assert_close(r2.cpu().numpy(), r1.cpu().numpy(), "fp32", "fused add+norm residual")
    assert_close(o2.cpu().numpy(), o1.cpu().numpy(), "fp32", "fused add+norm out")


def test_swiglu_and_add_residual_vs_reference_kernels(ref, ops):
    nt, inter = 16, 11008
    gu = rnd((nt, 2, inter), 9, -4, 4)
    want, got = torch.zeros(nt, inter, device="cuda"), torch.zeros(nt, inter, device="cuda")
    assert ref.ref_gpu_act_f32(P(gu), P(want), nt, inter) == 0
    ops.launchAct(gu, got)
    assert_close(got.cpu().numpy(), want.cpu().numpy(), "fp32", "swiglu")
    r, o = rnd((nt, 4096), 10), rnd((nt, 4096), 11)
    o1, o2 = o.clone(), o.clone()
    assert ref.ref_gpu_add_residual_f32(P(r), P(o1), nt, 4096) == 0
    ops.launchAddResidual(r, o2)
    assert torch.equal(o1, o2), "add residual must be bit-exact vs the reference kernel"


@pytest.mark.parametrize("dtype,sfx", [(torch.float32, "f32"), (torch.float16, "f16")])
def test_concat_and_repeat_kv_bit_exact_vs_reference_kernels(ref, ops, dtype, sfx):
    layers, layer, b, h, hk, s, max_seq, d = 3, 1, 2, 8, 2, 16, 48, 128
    ks, vs = rnd((b, hk, s, d), 12, dtype=dtype), rnd((b, hk, s, d), 13, dtype=dtype)
    cur, hist = ints([16, 9]), ints([5, 0])
    k1 = rnd((layers, b, hk, max_seq, d), 14, dtype=dtype)
    v1 = rnd((layers, b, hk, max_seq, d), 15, dtype=dtype)
    k2, v2 = k1.clone(), v1.clone()
    assert getattr(ref, f"ref_gpu_concat_{sfx}")(P(ks), P(vs), P(k1), P(v1), P(cur), P(hist), layers, layer,
                                                 b, hk, s, max_seq, d) == 0
    ops.launchConcatKVCache(ks, vs, layer, cur, hist, k2, v2)
    assert torch.equal(k1, k2) and torch.equal(v1, v2)
    ctx, max_k = ints([21, 9]), 24
    a1 = torch.zeros(b, h, max_k, d, dtype=dtype, device="cuda")
    b1, a2, b2 = a1.clone(), a1.clone(), a1.clone()
    assert getattr(ref, f"ref_gpu_repeat_{sfx}")(P(k1), P(v1), P(a1), P(b1), P(ctx), layers, layer, b, h, hk,
                                                 max_seq, max_k, d) == 0
    ops.launchRepeatKVCache(k2, v2, ctx, layer, a2, b2)
    assert torch.equal(a1, a2) and torch.equal(b1, b2)


def test_padding_offset_mask_transpose_bit_exact_vs_reference_kernels(ref, ops):
    lens, b, max_q = [5, 4, 7, 6], 4, 8
    nt = sum(lens)
    p1 = torch.zeros(b, max_q, dtype=torch.int32, device="cuda")
    c1 = torch.zeros(b + 1, dtype=torch.int32, device="cuda")
    p2, c2, dl = p1.clone(), c1.clone(), ints(lens)
    assert ref.ref_gpu_padding_offset(P(p1), P(c1), P(dl), b, max_q) == 0
    ops.launchCalPaddingoffset(p2, c2, dl)
    assert torch.equal(p1.flatten()[:nt], p2.flatten()[:nt]) and torch.equal(c1, c2)
    m1 = torch.zeros(b, max_q, max_q, device="cuda")
    m2 = m1.clone()
    assert ref.ref_gpu_causal_mask_f32(P(m1), P(dl), P(dl), b, max_q, max_q) == 0   # klen == qlen (D17)
    ops.launchBuildCausalMasks(m2, dl, dl)
    assert torch.equal(m1, m2)
    h, d = 4, 64
    src = rnd((b, h, max_q, d), 16)
    d1, d2 = torch.zeros(nt, h * d, device="cuda"), torch.zeros(nt, h * d, device="cuda")
    assert ref.ref_gpu_transpose_f32(P(src), P(p1), P(d1), nt, b, h, max_q, d) == 0
    ops.launchTransposeOutRemovePadding(src, p2.flatten()[:nt].contiguous(), d2)
    assert torch.equal(d1, d2)


@pytest.mark.parametrize("m,n,k,trans_b", [(13, 4096, 4096, 0), (1, 12288, 4096, 1), (5, 4096, 11008, 1),
                                           (64, 512, 1024, 1)])
def test_linear_f32_vs_reference_cublas(ref, ops, m, n, k, trans_b):
    x = rnd((m, k), 17)
    w = rnd((n, k) if trans_b else (k, n), 18, -0.05, 0.05)
    want, got = torch.zeros(m, n, device="cuda"), torch.zeros(m, n, device="cuda")
    assert ref.ref_gpu_linear_f32(P(x), P(w), P(want), m, n, k, trans_b) == 0
    ops.launchLinearGemm(x, w, got, trans_b=bool(trans_b))
    assert_close(got.cpu().numpy(), want.cpu().numpy(), "fp32", "linear vs cuBLAS fp32")


def test_rope_decode_f32_vs_reference_kernel_b1(ref, ops):
    h, d, step = 32, 128, 77
    qkv = rnd((1, 3 * h, d), 19)
    want, got = qkv.clone(), qkv.clone()
    assert ref.ref_gpu_rope_decode_f32(P(want), 1, 3 * h, d, step, d, C.c_float(10000.0)) == 0
    ops.launchRoPE(got, step, d, 10000.0)
    # the reference evaluates base^(2i/d) with device powf (~2 ulp): allow 1e-4 here, the oracle
    # (correctly rounded power) is held to 1e-5 in test_parity_gpu.py
    e = float((got - want).abs().max() / want.abs().max())
    assert e <= 1e-4, e


def test_softmax_f32_vs_reference_kernel_small_rows(ref, ops):
    b, h, q, k = 2, 3, 8, 32     # one warp per row: sound in the reference (D8)
    qk = rnd((b, h, q, k), 20, -3, 3)
    mask = torch.tril(torch.ones(q, k, device="cuda"), diagonal=k - q).repeat(b, 1, 1).contiguous()
    want, got = torch.zeros_like(qk), torch.zeros_like(qk)
    assert ref.ref_gpu_softmax_f32(P(qk), P(mask), P(want), b, h, q, k, C.c_float(0.125)) == 0
    ops.launchScaleMaskAndSoftmax(qk, mask, got, 0.125)
    assert_close(got.cpu().numpy(), want.cpu().numpy(), "fp32", "scale-mask-softmax")


@pytest.mark.parametrize("rows,vocab", [(1, 32000), (3, 32000), (2, 1031)])
def test_topk_indices_vs_reference_kernel(ref, ops, rows, vocab):
    """launchTopKforBeamSearch (topK.cu:147, K = 5, cub BlockReduce) on tie-free POSITIVE data (its
    heap starts at 1e-20 / id -1, topK.h:13-18): our b200_topk must return the same indices, in the
    same (descending) order, and the same values; our top-1 pick = its first index."""
    rng = np.random.default_rng(17 + rows)
    probs = np.stack([rng.permutation(vocab).astype(np.float32) + 1.0 for _ in range(rows)]) / vocab   # all distinct
    p = torch.from_numpy(probs).cuda()
    tmp_i = torch.zeros(rows * 8 * 5, dtype=torch.int32, device="cuda")
    tmp_v = torch.zeros(rows * 8 * 5, device="cuda")
    fin_i = torch.zeros(rows, 5, dtype=torch.int32, device="cuda")
    fin_v = torch.zeros(rows, 5, device="cuda")
    assert ref.ref_gpu_topk_f32(P(p), P(tmp_i), P(tmp_v), P(fin_i), P(fin_v), rows, vocab) == 0
    ids = torch.zeros(rows, 5, dtype=torch.int32, device="cuda")
    vals = torch.zeros(rows, 5, device="cuda")
    ops.launchTopK(p, ids, vals, 5)
    top1 = torch.zeros(rows, dtype=torch.int32, device="cuda")
    ops.argmax(p, top1)
    torch.cuda.synchronize()
    want = np.argsort(-probs, axis=1)[:, :5]
    assert np.array_equal(fin_i.cpu().numpy(), want), "the reference kernel itself disagrees with a sort"
    assert np.array_equal(ids.cpu().numpy(), fin_i.cpu().numpy()), "top-K indices differ from the reference kernel"
    assert np.array_equal(vals.cpu().numpy(), fin_v.cpu().numpy())
    assert np.array_equal(top1.cpu().numpy(), fin_i.cpu().numpy()[:, 0])
