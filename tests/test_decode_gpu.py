"""GPU parity for the decode hot path proper: split-KV decode attention, the fused
weight-streaming GEMV (norm prologue / residual / SwiGLU / RoPE+KV-append epilogues) and the
sampling-side ops, each against the CPU oracle; plus fused == unfused consistency."""
import numpy as np
import pytest
import torch

from conftest import assert_close
from oracle import cpu
from test_parity_gpu import DT, KINDS, dev, host, widen

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def ops(pkg):
    return pkg.ops


@pytest.fixture(scope="module")
def capi(pkg):
    return pkg.capi


def test_decode_attention_reference_test_data(ops):
    """tests/unittests/test_fused_decoder_attention.cu:269-349: B=1, h=hk=2, d=4, max_seq=4, step=4,
    ramp data (v taken from its declared slot in qkv, SURVEY §4 note on the test's h_v offset)."""
    B, h, hk, d, max_seq, step = 1, 2, 2, 4, 4, 4
    n = B * (h + 2 * hk) * d
    qkv = np.zeros(n, np.float32)
    for i in range(n):
        if i < B * h * d:
            qkv[i] = i + 1 if i < B * h * d // 2 else (i - 3) / 10
        elif i < B * (h + hk) * d:
            qkv[i] = i + 5 if i < B * (h + hk // 2) * d else (i + 1) / 10
        else:
            qkv[i] = i - 3 if i < B * (h + hk + hk // 2) * d else (i - 7) / 10
    qkv = qkv.reshape(B, h + 2 * hk, d)
    size = max_seq * B * hk * d
    kc = np.array([(i + 1) if i < size // 2 else (i - size // 2 + 1) / 10 for i in range(size)],
                  np.float32).reshape(1, B, hk, max_seq, d)
    vc = kc.copy()
    wk, wv = kc[0].copy(), vc[0].copy()
    want = np.zeros((B, h * d), np.float32)
    cpu.decode_attention(qkv, wk, wv, want, h, hk, step, True)
    dk, dv = dev(kc), dev(vc)
    out = torch.zeros(B, h * d, device="cuda")
    ops.launchDecoderMaskedMHA(dev(qkv), 0, dk, dv, step, out)
    assert_close(host(out), want, "fp32", "mha out")
    assert np.array_equal(host(dk)[0], wk) and np.array_equal(host(dv)[0], wv)


@pytest.mark.parametrize("kind", KINDS)
@pytest.mark.parametrize("B,h,hk,d,max_seq,step,layers", [
    (1, 32, 32, 128, 128, 65, 1),      # config 1: 64-token cache + the new token
    (1, 32, 32, 128, 640, 513, 2),     # config 2 shape
    (3, 32, 32, 128, 300, 1, 1),       # first token: only the appended row
    (2, 64, 8, 128, 1100, 1031, 1),    # 70B-shaped GQA 8x, odd length
    (4, 8, 2, 64, 257, 257, 1),        # d=64, GQA 4x, full cache
    (2, 6, 3, 20, 40, 33, 1),          # odd head size -> generic path
    (32, 32, 32, 128, 1024, 777, 1)])  # batch 32
def test_decode_attention(ops, kind, B, h, hk, d, max_seq, step, layers):
    rng = np.random.default_rng(B * 1000 + step)
    qkv = widen(rng.uniform(-1, 1, (B, h + 2 * hk, d)), kind)
    kc = widen(rng.uniform(-1, 1, (layers, B, hk, max_seq, d)), kind)
    vc = widen(rng.uniform(-1, 1, (layers, B, hk, max_seq, d)), kind)
    layer = layers - 1
    wk, wv = kc[layer].copy(), vc[layer].copy()
    want = np.zeros((B, h * d), np.float32)
    cpu.decode_attention(qkv, wk, wv, want, h, hk, step, True)
    dk, dv = dev(kc, kind), dev(vc, kind)
    out = torch.zeros(B, h * d, dtype=DT[kind], device="cuda")
    ops.launchDecoderMaskedMHA(dev(qkv, kind), layer, dk, dv, step, out)
    assert_close(host(out), want, kind, "mha out")
    assert np.array_equal(host(dk)[layer], wk), "k appended at row step-1, rest untouched"
    assert np.array_equal(host(dv)[layer], wv)
    if layers > 1:
        assert np.array_equal(host(dk)[0], kc[0]), "other layers untouched"
    # same answer when the row is already in the cache and step comes from device memory
    out2 = torch.zeros_like(out)
    ops.launchDecoderMaskedMHA(dev(qkv, kind), layer, dk, dv, max_seq, out2, append_kv=False,
                               step_dev=torch.tensor([step], dtype=torch.int32, device="cuda"))
    assert_close(host(out2), want, kind, "mha out (step from device)")


@pytest.mark.parametrize("kind", KINDS)
@pytest.mark.parametrize("nb", [1, 2, 3, 8])
@pytest.mark.parametrize("N,K", [(4096, 4096), (4096, 11008), (1001, 512)])
def test_gemv_plain_and_residual(ops, capi, kind, nb, N, K):
    rng = np.random.default_rng(N + K + nb)
    x = widen(rng.uniform(-1, 1, (nb, K)), kind)
    W = widen(0.02 * rng.standard_normal((N, K)), kind)
    bias = widen(0.1 * rng.standard_normal(N), kind)
    want = np.zeros((nb, N), np.float32)
    cpu.linear(x, W, want, True, bias)
    y = torch.zeros(nb, N, dtype=DT[kind], device="cuda")
    ops.gemv_fused(dev(x, kind), dev(W, kind), y, 0, bias=dev(bias, kind))
    assert_close(host(y), want, kind, "gemv+bias")
    # residual epilogue == linear then add_residual
    r = widen(rng.uniform(-1, 1, (nb, N)), kind)
    lin = np.zeros((nb, N), np.float32)
    cpu.linear(x, W, lin, True)
    want_r = widen(lin, kind) + r
    y2 = dev(r, kind)
    ops.gemv_fused(dev(x, kind), dev(W, kind), y2, capi.GEMV_RESIDUAL)
    assert_close(host(y2), want_r, kind, "gemv+residual")


@pytest.mark.parametrize("kind", KINDS)
@pytest.mark.parametrize("nb,H,I", [(1, 4096, 11008), (2, 512, 1000), (4, 4096, 1376)])
def test_gemv_norm_swiglu(ops, capi, kind, nb, H, I):
    """gate_up GEMV with fused RMSNorm prologue + SwiGLU epilogue == rmsnorm -> linear -> act"""
    rng = np.random.default_rng(H + I)
    x = widen(rng.uniform(-1, 1, (nb, H)), kind)
    g = widen(1 + 0.05 * rng.standard_normal(H), kind)
    W = widen(0.03 * rng.standard_normal((2 * I, H)), kind)
    xn = x.copy()
    cpu.rmsnorm(xn, None, g, 1e-5)
    xn = widen(xn, kind)
    gu = np.zeros((nb, 2 * I), np.float32)
    cpu.linear(xn, W, gu, True)
    gu = widen(gu, kind).reshape(nb, 2, I)
    want = np.zeros((nb, I), np.float32)
    cpu.swiglu(gu, want)
    dx = dev(x, kind)
    y = torch.zeros(nb, I, dtype=DT[kind], device="cuda")
    ops.gemv_fused(dx, dev(W, kind), y, capi.GEMV_NORM | capi.GEMV_SWIGLU, gamma=dev(g, kind), eps=1e-5)
    assert_close(host(y), want, kind, "norm+gate_up+swiglu")
    assert np.array_equal(host(dx), x), "input hidden state must stay untouched"
    # and bit-identical to the unfused GPU sequence modulo summation order (tight bound)
    xs = dev(x, kind)
    ops.launchRMSNorm(xs, None, dev(g, kind), 1e-5)
    gub = torch.zeros(nb, 2, I, dtype=DT[kind], device="cuda")
    ops.launchLinearGemm(xs, dev(W, kind), gub, trans_b=True)
    y_unfused = torch.zeros_like(y)
    ops.launchAct(gub, y_unfused)
    assert_close(host(y), host(y_unfused), kind, "fused vs unfused")


@pytest.mark.parametrize("kind", KINDS)
@pytest.mark.parametrize("nb,h,hk,d,H,step,max_seq", [(1, 32, 32, 128, 4096, 65, 128), (2, 16, 2, 128, 1024, 9, 16),
                                                      (1, 8, 8, 64, 512, 1, 4)])
def test_gemv_norm_rope_kv(ops, capi, kind, nb, h, hk, d, H, step, max_seq):
    """qkv GEMV with fused RMSNorm + RoPE + KV append == rmsnorm -> linear -> RoPE -> append"""
    rng = np.random.default_rng(h + step)
    N = (h + 2 * hk) * d
    x = widen(rng.uniform(-1, 1, (nb, H)), kind)
    g = widen(1 + 0.05 * rng.standard_normal(H), kind)
    W = widen(0.03 * rng.standard_normal((N, H)), kind)
    xn = x.copy()
    cpu.rmsnorm(xn, None, g, 1e-5)
    xn = widen(xn, kind)
    qkv = np.zeros((nb, N), np.float32)
    cpu.linear(xn, W, qkv, True)
    qkv = widen(qkv, kind).reshape(nb, h + 2 * hk, d)
    cpu.rope_decode(qkv, h, hk, step, d, 10000.0)
    kc0 = widen(rng.uniform(-1, 1, (nb, hk, max_seq, d)), kind)
    vc0 = widen(rng.uniform(-1, 1, (nb, hk, max_seq, d)), kind)
    wk, wv = kc0.copy(), vc0.copy()
    wk[:, :, step - 1] = widen(qkv[:, h:h + hk], kind)
    wv[:, :, step - 1] = widen(qkv[:, h + hk:], kind)
    dk, dv = dev(kc0, kind), dev(vc0, kind)
    y = torch.zeros(nb, h + 2 * hk, d, dtype=DT[kind], device="cuda")
    step_dev = torch.tensor([step], dtype=torch.int32, device="cuda")
    ops.gemv_fused(dev(x, kind), dev(W, kind), y, capi.GEMV_NORM | capi.GEMV_ROPE_KV, gamma=dev(g, kind),
                   eps=1e-5, k_cache_layer=dk, v_cache_layer=dv, head_num=h, kv_head_num=hk, head_size=d,
                   step=0, rotary_dim=d, rotary_base=10000.0, step_dev=step_dev)
    assert_close(host(y), qkv, kind, "qkv after rope")
    assert_close(host(dk), wk, kind, "k cache")
    assert_close(host(dv), wv, kind, "v cache")
    untouched = np.ones(max_seq, bool)
    untouched[step - 1] = False
    assert np.array_equal(host(dk)[:, :, untouched], kc0[:, :, untouched])


@pytest.mark.parametrize("kind", KINDS)
def test_embedding_and_topk(ops, kind):
    rng = np.random.default_rng(11)
    V, H, nt = 30000, 4096, 64         # tests/unittests/test_input_embedding.cu:67-69
    table = widen(rng.standard_normal((V, H)), kind)
    ids = rng.integers(0, V, nt).astype(np.int32)
    out = torch.zeros(nt, H, dtype=DT[kind], device="cuda")
    ops.launchInputEmbedding(dev(ids), out, dev(table, kind))
    assert np.array_equal(host(out), table[ids])
    # top-K: tests/unittests/test_topk.cu:15-19 (V=30000, K=5) with ties
    logits = rng.standard_normal((3, V))
    logits[1, 100] = logits[1, 20000] = logits[1].max() + 1     # tie -> lower index first
    logits = widen(logits, kind)
    want_i = np.zeros((3, 5), np.int32)
    want_v = np.zeros((3, 5), np.float32)
    cpu.topk(logits, want_i, want_v, 5)
    di = torch.zeros(3, 5, dtype=torch.int32, device="cuda")
    dv = torch.zeros(3, 5, dtype=DT[kind], device="cuda")
    ops.launchTopK(dev(logits, kind), di, dv, 5)
    assert np.array_equal(di.cpu().numpy(), want_i)
    assert np.array_equal(host(dv), want_v)
    am = torch.zeros(3, dtype=torch.int32, device="cuda")
    ops.argmax(dev(logits, kind), am.view(3, 1))
    assert np.array_equal(am.cpu().numpy(), want_i[:, 0])
