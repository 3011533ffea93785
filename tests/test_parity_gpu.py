"""GPU parity: every C-ABI op (through ctypes -> libb200llama.so) against the CPU oracle on the
same seeded inputs. Shapes: the reference unit tests' own (tests/unittests/*.cu) plus
LLaMA-2-7B / 70B shaped cases and ragged / empty edge cases.
Tolerance (north_star): normalised max error <= 1e-5 fp32, <= 2e-2 fp16; copies, masks, padding
offsets and top-K indices are bit-exact."""
import numpy as np
import pytest
import torch

from conftest import assert_close
from oracle import cpu

pytestmark = pytest.mark.gpu

DT = {"fp32": torch.float32, "fp16": torch.float16}
KINDS = ["fp32", "fp16"]


def dev(a, kind=None):
    t = torch.from_numpy(np.ascontiguousarray(a))
    if kind is not None:
        t = t.to(DT[kind])
    return t.cuda()


def host(t):
    return t.detach().float().cpu().numpy()


def widen(a, kind):
    """the values the GPU actually sees (fp16 rounding applied), as float32"""
    return np.ascontiguousarray(a.astype(np.float16).astype(np.float32) if kind == "fp16" else a.astype(np.float32))


@pytest.fixture(scope="module")
def ops(pkg):
    assert torch.cuda.is_available()
    return pkg.ops


# --------------------------------------------------------------------------- norm family
@pytest.mark.parametrize("kind", KINDS)
@pytest.mark.parametrize("nt,h", [(64, 4096), (1, 4096), (3, 8192), (2, 32), (5, 11008), (2, 40960)])
def test_rmsnorm(ops, kind, nt, h):
    rng = np.random.default_rng(nt * 131 + h)
    x = widen(rng.uniform(-1, 1, (nt, h)), kind)
    g = widen(1 + 0.02 * rng.standard_normal(h), kind)
    want, res = x.copy(), np.zeros_like(x)
    cpu.rmsnorm(want, res, g, 1e-5)
    dx, dres = dev(x, kind), torch.zeros(nt, h, dtype=DT[kind], device="cuda")
    ops.launchRMSNorm(dx, dres, dev(g, kind), 1e-5)
    assert_close(host(dx), want, kind, "rmsnorm out")
    assert np.array_equal(host(dres), x), "residual must be the pre-norm input, bit-exact"


def test_rmsnorm_reference_test_data(ops):
    # tests/unittests/test_rmsnorm.cu:139-175 with its own abs 1e-6 check
    nt, h = 64, 4096
    x = (np.arange(nt * h) % 2 + 1).astype(np.float32).reshape(nt, h)
    g = (np.arange(h) % 2 + 1).astype(np.float32)
    want = x.copy()
    cpu.rmsnorm(want, None, g, 1e-6)
    dx = dev(x)
    ops.launchRMSNorm(dx, None, dev(g), 1e-6)
    assert np.max(np.abs(host(dx) - want)) <= 1e-6


@pytest.mark.parametrize("kind", KINDS)
@pytest.mark.parametrize("nt,h,bias", [(2, 32, True), (1, 4096, False), (64, 4096, True), (3, 8192, False)])
def test_fused_add_bias_residual_rmsnorm(ops, kind, nt, h, bias):
    rng = np.random.default_rng(nt + h)
    r = widen(rng.uniform(-1, 1, (nt, h)), kind)
    o = widen(rng.uniform(-1, 1, (nt, h)), kind)
    b = widen(0.1 * rng.standard_normal(h), kind) if bias else None
    g = widen(1 + 0.02 * rng.standard_normal(h), kind)
    wr, wo = r.copy(), o.copy()
    cpu.fused_add_bias_residual_rmsnorm(wr, wo, b, g, 1e-5)
    dr, do = dev(r, kind), dev(o, kind)
    ops.launchFusedAddBiasResidualRMSNorm(dr, do, dev(b, kind) if bias else None, dev(g, kind), 1e-5)
    assert_close(host(dr), wr, kind, "residual")
    assert_close(host(do), wo, kind, "normed out")


def test_fused_add_norm_reference_test_data(ops):
    # tests/unittests/test_fused_addresidual_norm.cu:69-141 -> 1/sqrt(1.5)
    dr, do = torch.zeros(2, 32, device="cuda"), torch.ones(2, 32, device="cuda")
    ops.launchFusedAddBiasResidualRMSNorm(dr, do, torch.zeros(32, device="cuda"), torch.ones(32, device="cuda"), 0.5)
    assert np.max(np.abs(host(do) - 1 / np.sqrt(1.5))) <= 1e-6


@pytest.mark.parametrize("kind", KINDS)
@pytest.mark.parametrize("nt,h", [(1, 4096), (64, 4096), (7, 24)])
def test_add_residual(ops, kind, nt, h):
    rng = np.random.default_rng(5)
    r, o = widen(rng.standard_normal((nt, h)), kind), widen(rng.standard_normal((nt, h)), kind)
    want = o.copy()
    cpu.add_residual(r, want)
    do = dev(o, kind)
    ops.launchAddResidual(dev(r, kind), do)
    assert_close(host(do), want, kind, "add_residual")


@pytest.mark.parametrize("kind", KINDS)
@pytest.mark.parametrize("nt,inter", [(16, 11008), (1, 11008), (3, 28672), (2, 37)])
def test_swiglu(ops, kind, nt, inter):
    rng = np.random.default_rng(inter)
    gi = widen(rng.uniform(-4, 4, (nt, 2, inter)), kind)
    want = np.zeros((nt, inter), np.float32)
    cpu.swiglu(gi, want)
    out = torch.zeros(nt, inter, dtype=DT[kind], device="cuda")
    ops.launchAct(dev(gi, kind), out)
    assert_close(host(out), want, kind, "swiglu")


def test_swiglu_reference_test_data(ops):
    out = torch.zeros(16, 11008, device="cuda")
    ops.launchAct(torch.ones(16, 2, 11008, device="cuda"), out)
    assert np.max(np.abs(host(out) - 1 / (1 + np.exp(-1.0)))) <= 1e-6
    with pytest.raises(RuntimeError):
        ops.launchAct(torch.ones(16, 3, 8, device="cuda"), out)   # LLM_CHECK(shape[1]==2), act_kernel.cu:62


# --------------------------------------------------------------------------- RoPE
@pytest.mark.parametrize("kind", KINDS)
@pytest.mark.parametrize("b,h,hk,d,step,rot", [(1, 32, 32, 128, 65, 128), (4, 64, 8, 128, 513, 128),
                                               (2, 4, 2, 16, 3, 8), (1, 32, 32, 128, 4096, 128)])
def test_rope_decode(ops, kind, b, h, hk, d, step, rot):
    rng = np.random.default_rng(step)
    qkv = widen(rng.uniform(-1, 1, (b, h + 2 * hk, d)), kind)
    want = qkv.copy()
    cpu.rope_decode(want, h, hk, step, rot, 10000.0)
    dq = dev(qkv, kind)
    ops.launchRoPE(dq, step, rot, 10000.0, kv_head_num=hk)
    # cos/sin of large angles differ by a few ulp between libm and CUDA: tolerance per north_star
    assert_close(host(dq), want, "fp16" if kind == "fp16" else "fp32", "rope decode")
    assert np.array_equal(host(dq)[:, h + hk:], qkv[:, h + hk:]), "v heads untouched"


@pytest.mark.parametrize("kind", KINDS)
@pytest.mark.parametrize("lens,hist,S,h,hk,d,bias", [
    ([7], [0], 32, 32, 32, 128, False),          # tests/unittests/test_bias_and_RoPE.cu shape
    ([5, 0, 3], [2, 9, 0], 6, 8, 2, 64, True),   # ragged + empty + history + GQA + bias
    ([64, 64], [0, 0], 64, 32, 32, 128, False)])
def test_qkv_bias_rope_transpose(ops, kind, lens, hist, S, h, hk, d, bias):
    rng = np.random.default_rng(len(lens) + S)
    B, nt = len(lens), int(sum(lens))
    lens_a, hist_a = np.array(lens, np.int32), np.array(hist, np.int32)
    pad = np.zeros((B, S), np.int32)
    cum = np.zeros(B + 1, np.int32)
    cpu.cal_padding_offset(pad, cum, lens_a)
    pad_flat = np.ascontiguousarray(pad.reshape(-1)[:max(nt, 1)])
    QKV = widen(rng.uniform(-1, 1, (nt, h + 2 * hk, d)), kind)
    bs = widen(0.1 * rng.standard_normal((h + 2 * hk) * d), kind) if bias else None
    q = np.full((B, h, S, d), 9, np.float32)
    k = np.full((B, hk, S, d), 9, np.float32)
    v = np.full((B, hk, S, d), 9, np.float32)
    cpu.qkv_bias_rope_transpose(q, k, v, QKV, bs, pad_flat, hist_a, d, 10000.0)
    dq, dk, dv = dev(np.full_like(q, 9), kind), dev(np.full_like(k, 9), kind), dev(np.full_like(v, 9), kind)
    ops.launchAddFusedQKVBiasTransposeAndRoPE(dq, dk, dv, dev(QKV, kind), dev(bs, kind) if bias else None,
                                              dev(pad_flat), dev(hist_a), dev(lens_a), d, 10000.0)
    assert_close(host(dq), q, kind, "q")
    assert_close(host(dk), k, kind, "k")
    assert_close(host(dv), v, kind, "v")


# --------------------------------------------------------------------------- copies (bit-exact)
@pytest.mark.parametrize("kind", KINDS)
@pytest.mark.parametrize("layers,B,hk,h,S,max_seq,d,cur,hist", [
    (1, 1, 2, 2, 16, 32, 8, [16], [1]),                 # tests/unittests/test_concat_kv.cu
    (2, 1, 2, 2, 2, 4, 2, [2], [0]),                    # tests/unittests/test_repeat_kv.cu (d=2: scalar path)
    (3, 3, 2, 8, 5, 9, 128, [5, 0, 2], [1, 3, 4]),      # ragged, empty, GQA 4x
    (2, 2, 8, 64, 64, 128, 128, [64, 33], [0, 64])])    # 70B-shaped GQA 8x
def test_concat_and_repeat_kv(ops, kind, layers, B, hk, h, S, max_seq, d, cur, hist):
    npdt = np.float16 if kind == "fp16" else np.float32
    rng = np.random.default_rng(B * 7 + d)
    ks = rng.standard_normal((B, hk, S, d)).astype(npdt)
    vs = rng.standard_normal((B, hk, S, d)).astype(npdt)
    kc = rng.standard_normal((layers, B, hk, max_seq, d)).astype(npdt)
    vc = rng.standard_normal((layers, B, hk, max_seq, d)).astype(npdt)
    cur_a, hist_a = np.array(cur, np.int32), np.array(hist, np.int32)
    layer = layers - 1
    wk, wv = kc.copy(), vc.copy()
    cpu.concat_kv(ks, wk, layer, cur_a, hist_a)
    cpu.concat_kv(vs, wv, layer, cur_a, hist_a)
    dk, dv = dev(kc), dev(vc)
    ops.launchConcatKVCache(dev(ks), dev(vs), layer, dev(cur_a), dev(hist_a), dk, dv)
    assert np.array_equal(dk.cpu().numpy(), wk) and np.array_equal(dv.cpu().numpy(), wv)

    ctx = (cur_a + hist_a).astype(np.int32)
    max_k = int(ctx.max())
    rk = np.full((B, h, max_k, d), 3, npdt)
    rv = np.full((B, h, max_k, d), 3, npdt)
    cpu.repeat_kv(wk, rk, layer, ctx)
    cpu.repeat_kv(wv, rv, layer, ctx)
    drk, drv = dev(np.full_like(rk, 3)), dev(np.full_like(rv, 3))
    ops.launchRepeatKVCache(dk, dv, dev(ctx), layer, drk, drv)
    assert np.array_equal(drk.cpu().numpy(), rk) and np.array_equal(drv.cpu().numpy(), rv)


@pytest.mark.parametrize("kind", KINDS)
@pytest.mark.parametrize("lens,S,h,d", [([3, 2], 4, 2, 2), ([5, 0, 3], 6, 8, 64), ([128, 77], 128, 32, 128)])
def test_transpose_remove_padding(ops, kind, lens, S, h, d):
    # first case: tests/unittests/test_fused_trans_remv_pad.cu (B=2,h=2,S=4,d=2,Nt=5)
    npdt = np.float16 if kind == "fp16" else np.float32
    B, nt = len(lens), int(sum(lens))
    pad = np.zeros((B, S), np.int32)
    cpu.cal_padding_offset(pad, np.zeros(B + 1, np.int32), np.array(lens, np.int32))
    pad_flat = np.ascontiguousarray(pad.reshape(-1)[:nt])
    src = (np.arange(B * h * S * d) % 1021).astype(npdt).reshape(B, h, S, d)
    want = np.zeros((nt, h * d), npdt)
    cpu.transpose_remove_padding(src, want, pad_flat)
    out = torch.zeros(nt, h * d, dtype=DT[kind], device="cuda")
    ops.launchTransposeOutRemovePadding(dev(src), dev(pad_flat), out)
    assert np.array_equal(out.cpu().numpy(), want)


@pytest.mark.parametrize("lens,max_q", [([3, 3, 3], 5), ([1], 1), ([0, 4, 0, 2], 4), ([2048] * 4, 2048),
                                        (list(range(1, 1300)), 1300)])
def test_cal_padding_offset(ops, lens, max_q):
    B = len(lens)
    la = np.array(lens, np.int32)
    pad = np.full((B, max_q), -1, np.int32)
    cum = np.zeros(B + 1, np.int32)
    cpu.cal_padding_offset(pad, cum, la)
    dp, dc = dev(np.full((B, max_q), -1, np.int32)), torch.zeros(B + 1, dtype=torch.int32, device="cuda")
    ops.launchCalPaddingoffset(dp, dc, dev(la))
    assert np.array_equal(dp.cpu().numpy(), pad) and np.array_equal(dc.cpu().numpy(), cum)
    if lens == [3, 3, 3]:   # golden vector, tests/unittests/test_cal_paddingoffset.cu:54-56
        assert dp.cpu().numpy().reshape(-1)[:9].tolist() == [0, 0, 0, 2, 2, 2, 4, 4, 4]
        assert dc.cpu().numpy().tolist() == [0, 3, 6, 9]


@pytest.mark.parametrize("kind", KINDS)
@pytest.mark.parametrize("ql,kl,mq,mk", [([5], [5], 5, 5), ([3, 4], [6, 4], 4, 7), ([0, 2], [0, 9], 3, 9),
                                         ([2048, 1000], [2048, 1500], 2048, 2048)])
@pytest.mark.parametrize("hv", [True, False])
def test_build_causal_mask(ops, kind, ql, kl, mq, mk, hv):
    B = len(ql)
    want = np.zeros((B, mq, mk), np.float32)
    cpu.build_causal_mask(want, np.array(ql, np.int32), np.array(kl, np.int32), hv)
    m = torch.full((B, mq, mk), 7, dtype=DT[kind], device="cuda")
    ops.launchBuildCausalMasks(m, dev(np.array(ql, np.int32)), dev(np.array(kl, np.int32)), hv)
    assert np.array_equal(host(m), want)


# --------------------------------------------------------------------------- softmax
@pytest.mark.parametrize("kind", KINDS)
@pytest.mark.parametrize("B,h,ql,kl", [(1, 2, 8, 8), (2, 4, 33, 40), (1, 2, 16, 2048), (1, 1, 4, 5000), (2, 3, 5, 7)])
def test_scale_mask_softmax(ops, kind, B, h, ql, kl):
    if kind == "fp16" and kl % 2:
        pytest.skip("reference requires even k_len in half (attn_softmax_kernel.cu:347)")
    rng = np.random.default_rng(kl)
    qk = widen(rng.uniform(-8, 8, (B, h, ql, kl)), kind)
    mask = np.zeros((B, ql, kl), np.float32)
    cpu.build_causal_mask(mask, np.full(B, ql, np.int32), np.full(B, kl, np.int32), True)
    want = np.zeros_like(qk)
    cpu.scale_mask_softmax(qk, mask, want, 0.125)
    dq = dev(qk, kind)
    ops.launchScaleMaskAndSoftmax(dq, dev(mask, kind), dq, 0.125)   # in place, as context_attention.cpp:150
    assert_close(host(dq), want, kind, "softmax")


# --------------------------------------------------------------------------- linear
@pytest.mark.parametrize("kind", KINDS)
@pytest.mark.parametrize("m,n,k,tb", [(1, 12288, 4096, True), (1, 4096, 11008, True), (1, 22016, 4096, True),
                                      (2, 4096, 4096, True), (3, 1000, 512, True), (8, 4096, 4096, True),
                                      (5, 4097, 4096, True), (1, 32000, 4096, True),
                                      (13, 256, 4096, False), (40, 300, 520, True), (9, 64, 100, False)])
def test_linear(ops, kind, m, n, k, tb):
    rng = np.random.default_rng(m * 3 + n)
    x = widen(rng.uniform(-1, 1, (m, k)), kind)
    W = widen(0.02 * rng.standard_normal((n, k) if tb else (k, n)), kind)
    want = np.zeros((m, n), np.float32)
    cpu.linear(x, W, want, trans_b=tb)
    y = torch.zeros(m, n, dtype=DT[kind], device="cuda")
    ops.launchLinearGemm(dev(x, kind), dev(W, kind), y, trans_b=tb)
    assert_close(host(y), want, kind, f"linear {m}x{n}x{k}")


def test_linear_reference_test_data(ops):
    # tests/unittests/test_linear.cu:39-60: 13x4096 . 4096x4096, i%3 integer data -> exact
    m, k, n = 13, 4096, 4096
    a = (np.arange(m * k) % 3).astype(np.float32).reshape(m, k)
    w = (np.arange(k * n) % 3).astype(np.float32).reshape(k, n)
    want = np.zeros((m, n), np.float32)
    cpu.linear(a, w, want, trans_b=False)
    y = torch.zeros(m, n, device="cuda")
    ops.launchLinearGemm(dev(a), dev(w), y)
    assert np.array_equal(host(y), want)
    with pytest.raises(RuntimeError):
        ops.launchLinearGemm(dev(a), dev(w[:100]), y)   # "2nd dim of input MUST = 1st dim of weight"


@pytest.mark.parametrize("kind", KINDS)
def test_batched_gemm(ops, kind):
    rng = np.random.default_rng(0)
    B, h, s, l, d = 2, 3, 17, 29, 32
    q = widen(rng.standard_normal((B, h, s, d)), kind)
    kk = widen(rng.standard_normal((B, h, l, d)), kind)
    want = np.zeros((B, h, s, l), np.float32)
    cpu.batched_gemm(q, kk, want, True)
    out = torch.zeros(B, h, s, l, dtype=DT[kind], device="cuda")
    ops.launchLinearStridedBatchGemm(dev(q, kind), dev(kk, kind), out, trans_b=True)
    assert_close(host(out), want, kind, "QK^T")
    p = widen(rng.uniform(0, 1, (B, h, s, l)), kind)
    v = widen(rng.standard_normal((B, h, l, d)), kind)
    want2 = np.zeros((B, h, s, d), np.float32)
    cpu.batched_gemm(p, v, want2, False)
    out2 = torch.zeros(B, h, s, d, dtype=DT[kind], device="cuda")
    ops.launchLinearStridedBatchGemm(dev(p, kind), dev(v, kind), out2, trans_b=False)
    assert_close(host(out2), want2, kind, "PV")
