"""fp8 (e4m3) KV cache (csrc/kv_fp8.cu, b200_decode_attention_fp8, B200_ENGINE_KV_FP8).

The fp8 cache is an engine option, not bit-compatible with the reference's T-typed cache (llama.cpp:77-78), so its
contract is stated here: a K / V row is rounded ONCE, when it enters the cache (scale = amax(row) / 448, e4m3 round to
nearest), and attention is the reference's masked decode attention (fused_decoder_self_attention.cu:80-291) over the
rounded rows. The oracle is therefore run on rows rounded the same way (`emulate_fp8_rows`: torch's float8_e4m3fn
cast), and the usual tolerances apply: logits / attention output rel <= 2e-2 (fp16), cached rows equal to the
emulation up to one e4m3 code on the few elements whose fp16 source differs in its last bit. The oracle's decode loop
rounds the new token's rows before it attends (`OracleModel.kv_round`), as the GPU does."""
import ctypes as C

import numpy as np
import pytest
import torch

from conftest import assert_close, rel_err
from oracle import cpu, model as omodel
from test_engine_gpu import SMALL, download_weights

pytestmark = pytest.mark.gpu


def emulate_fp8_rows(rows):
    """rows [..., d] float32 -> what the cache holds after kv_fp8.cu's quantize_row + dequantisation"""
    t = torch.from_numpy(np.ascontiguousarray(rows, dtype=np.float32))
    amax = t.abs().amax(dim=-1, keepdim=True)
    scale = torch.where(amax > 0, amax * np.float32(1.0 / 448.0), torch.ones_like(amax))
    q = (t * (1.0 / scale)).to(torch.float8_e4m3fn).float()
    return (q * scale).numpy()


def _vp(t):
    return C.c_void_p(t.data_ptr())


def assert_fp8_rows_close(got, want, what):
    """Rows that went through DIFFERENT fp16 computations before the same rounding (GPU GEMM vs oracle) may land on
    neighbouring e4m3 codes where a value sits on a rounding boundary: at most ONE code step of the top binade
    (32 / 448 of the row maximum), on few elements. Everything else must agree to fp16 rounding."""
    got, want = np.asarray(got, np.float64), np.asarray(want, np.float64)
    amax = np.maximum(np.max(np.abs(want), axis=-1, keepdims=True), 1e-30)
    err = np.abs(got - want) / amax
    assert np.all(np.isfinite(got)), f"{what}: non-finite"
    # (the two rows also carry slightly different scales: 15 % slack on the one-step bound)
    assert err.max() <= 1.15 * 32.0 / 448.0, f"{what}: an element is off by more than one e4m3 step ({err.max():.3e} of its row max)"
    assert np.mean(err > 2e-2) < 0.05, f"{what}: {100 * np.mean(err > 2e-2):.2f}% of the elements sit on another code"


@pytest.mark.parametrize("B,h,hk,max_seq,step", [(2, 32, 32, 800, 777), (3, 8, 2, 320, 300), (1, 4, 2, 70, 1), (4, 8, 8, 4200, 4097)])
def test_append_and_attention_fp8_vs_oracle(pkg, B, h, hk, max_seq, step):
    lib = pkg.capi.lib()
    d = 128
    g = torch.Generator(device="cpu").manual_seed(step)
    hist = step - 1
    k16 = ((torch.rand(B, hk, max(hist, 1), d, generator=g) * 2 - 1) * torch.rand(B, hk, max(hist, 1), 1, generator=g)).half().cuda()
    v16 = ((torch.rand(B, hk, max(hist, 1), d, generator=g) * 2 - 1)).half().cuda()
    qkv = (torch.rand(B, h + 2 * hk, d, generator=g) * 2 - 1).half().cuda()
    k8 = torch.zeros(B, hk, max_seq, d, dtype=torch.uint8, device="cuda")
    v8 = torch.zeros_like(k8)
    ks = torch.zeros(B, hk, max_seq, device="cuda")
    vs = torch.zeros_like(ks)
    n_rows = torch.full((B,), hist, dtype=torch.int32, device="cuda")
    row0 = torch.zeros(B, dtype=torch.int32, device="cuda")
    if hist > 0:
        for src, dst, sc in ((k16, k8, ks), (v16, v8, vs)):
            assert lib.b200_kv_quantize_fp8(_vp(src), _vp(dst), _vp(sc), _vp(n_rows), _vp(row0), B, hk, hist, max_seq, d, hist, None) == 0
    assert lib.b200_kv_append_fp8(_vp(qkv), _vp(k8), _vp(v8), _vp(ks), _vp(vs), B, h, hk, d, max_seq, step, None, 0, None) == 0
    # the cache as the attention kernel sees it
    kd = torch.empty(B, hk, max_seq, d, dtype=torch.float16, device="cuda")
    vd = torch.empty_like(kd)
    assert lib.b200_kv_dequant_fp8(_vp(k8), _vp(ks), _vp(kd), B * hk * max_seq, d, None) == 0
    assert lib.b200_kv_dequant_fp8(_vp(v8), _vp(vs), _vp(vd), B * hk * max_seq, d, None) == 0
    torch.cuda.synchronize()
    # rows hold exactly the emulated rounding of their sources (history and the appended token)
    if hist > 0:
        assert_close(kd[:, :, :hist].float().cpu().numpy(), emulate_fp8_rows(k16.float().cpu().numpy()), "fp16", "quantised K history")
    new_k = qkv[:, h:h + hk].float().cpu().numpy()
    assert_close(kd[:, :, step - 1].float().cpu().numpy(), emulate_fp8_rows(new_k), "fp16", "appended K row")
    assert float(kd[:, :, step:].abs().max()) == 0.0 if step < max_seq else True
    # attention over the fp8 cache vs the oracle over the dequantised rows
    nbytes = lib.b200_decode_attention_workspace_bytes(B, h, d, max_seq)
    ws = torch.empty(nbytes, dtype=torch.uint8, device="cuda")
    out = torch.zeros(B, h * d, dtype=torch.float16, device="cuda")
    rc = lib.b200_decode_attention_fp8(_vp(qkv), _vp(k8), _vp(v8), _vp(ks), _vp(vs), _vp(out), _vp(ws), nbytes, B, h, hk, d,
                                       max_seq, step, None, 0, None)
    assert rc == 0
    torch.cuda.synchronize()
    want = np.zeros((B, h * d), np.float32)
    cpu.decode_attention(np.ascontiguousarray(qkv.float().cpu().numpy()), np.ascontiguousarray(kd.float().cpu().numpy()),
                         np.ascontiguousarray(vd.float().cpu().numpy()), want, h, hk, step, False)
    assert_close(out.float().cpu().numpy(), want, "fp16", f"fp8 attention B={B} h={h} hk={hk} step={step}")


def test_fp8_rejects_unsupported_shapes(pkg):
    lib = pkg.capi.lib()
    t = torch.zeros(4096, dtype=torch.float16, device="cuda")
    f = torch.zeros(4096, device="cuda")
    # head size 64 / group of 8 are outside the fp8 kernel: B200_EUNSUPPORTED (-2), nothing launched
    assert lib.b200_decode_attention_fp8(_vp(t), _vp(t), _vp(t), _vp(f), _vp(f), _vp(t), None, 0, 1, 4, 4, 64, 8, 4, None, 0, None) == -2
    assert lib.b200_decode_attention_fp8(_vp(t), _vp(t), _vp(t), _vp(f), _vp(f), _vp(t), None, 0, 1, 8, 1, 128, 8, 4, None, 0, None) == -2


def _quantise_oracle_rows(ora, lo, hi):
    ora.k[:, :, :, lo:hi] = emulate_fp8_rows(ora.k[:, :, :, lo:hi])
    ora.v[:, :, :, lo:hi] = emulate_fp8_rows(ora.v[:, :, :, lo:hi])


@pytest.mark.parametrize("B,n", [(2, 40), (5, 130)])
def test_engine_fp8_prefill_then_decode_vs_oracle(pkg, B, n):
    """B200_ENGINE_KV_FP8 on the small model: the prompt pass attends in fp16 and stores rounded rows, decode appends
    rounded rows and attends over them (batch 2: GEMV linears, batch 5: tcgen05 linears); oracle with emulated rounding."""
    E = pkg.engine
    max_seq, steps = n + 12, 5
    eng = E.Engine(B, max_seq, dtype=pkg.F16, kv_fp8=True, **SMALL)
    eng.init_random(seed=9, std=0.05)
    w = download_weights(E, eng, SMALL["num_layers"])
    ora = omodel.OracleModel(w, SMALL, B, max_seq, "fp16")
    ora.kv_round = emulate_fp8_rows
    rng = np.random.default_rng(n)
    ids = rng.integers(0, SMALL["vocab_size"], (B, n)).astype(np.int32)
    got_ids = eng.prefill(ids)
    H = SMALL["head_num"] * SMALL["head_size"]
    x = np.zeros((B * n, H), np.float32)
    cpu.embedding(np.ascontiguousarray(ids.reshape(-1)), w["embed"], x)
    x = ora.prefill_layers(x, [n] * B, [0] * B)
    ora.step = n
    _quantise_oracle_rows(ora, 0, n)
    last = np.ascontiguousarray(x.reshape(B, n, H)[:, -1])
    cpu.rmsnorm(last, None, w["final_norm"], 1e-5)
    last = ora.rt(last)
    want_logits = np.zeros((B, SMALL["vocab_size"]), np.float32)
    cpu.linear(last, w["lm_head"], want_logits, True)
    assert_close(eng.logits().float().cpu().numpy(), ora.rt(want_logits), "fp16", "fp8 engine: prefill logits")
    kc, vc = eng.kv_caches()   # dequantised copies
    assert_fp8_rows_close(kc[:, :, :, :n].float().cpu().numpy(), ora.k[:, :, :, :n], "fp8 engine: K rows after prefill")
    assert_fp8_rows_close(vc[:, :, :, :n].float().cpu().numpy(), ora.v[:, :, :, :n], "fp8 engine: V rows after prefill")
    assert float(kc[:, :, :, n:].abs().max()) == 0.0
    # From here on the oracle attends over the rows the ENGINE holds (checked against the emulation just above and
    # after every step below): elements that landed on a neighbouring code would otherwise add up over the steps.
    ora.k[:, :, :, :n] = kc[:, :, :, :n].float().cpu().numpy()
    ora.v[:, :, :, :n] = vc[:, :, :, :n].float().cpu().numpy()
    ids1 = np.argmax(want_logits, axis=1).astype(np.int32)
    agree = total = 0
    for s in range(steps):
        got = eng.decode_step_host(ids1)
        want_logits, want_ids = ora.decode_step(ids1)   # appends rounded rows (kv_round)
        kc, vc = eng.kv_caches()
        assert_fp8_rows_close(kc[:, :, :, n + s].float().cpu().numpy(), ora.k[:, :, :, n + s], f"fp8 engine: K row of step {s}")
        assert_fp8_rows_close(vc[:, :, :, n + s].float().cpu().numpy(), ora.v[:, :, :, n + s], f"fp8 engine: V row of step {s}")
        ora.k[:, :, :, n + s] = kc[:, :, :, n + s].float().cpu().numpy()
        ora.v[:, :, :, n + s] = vc[:, :, :, n + s].float().cpu().numpy()
        assert_close(eng.logits().float().cpu().numpy(), want_logits, "fp16", f"fp8 engine: decode step {s}")
        srt = np.sort(want_logits, axis=1)
        band = 4 * 2e-2 * np.max(np.abs(want_logits))
        for b in range(B):
            if srt[b, -1] - srt[b, -2] > band:
                total += 1
                agree += int(got[b] == want_ids[b])
        ids1 = want_ids
    assert agree == total
    assert eng.step == n + steps
    # a second prompt pass over the existing history is refused loudly, not run on the wrong cache
    with pytest.raises(RuntimeError):
        eng.prefill(ids[:, :4])
    eng.close()


def test_cfg4_shape_fp8_cache(pkg):
    """configs[3] geometry (7B heads, 32 sequences x 4096 cached keys, 2 layers) with the fp8 cache: two teacher-forced
    steps against the oracle on the dequantised rows; the cache is half the bytes (+ 4 B per row of scales)."""
    E = pkg.engine
    L7B = dict(head_num=32, kv_head_num=32, head_size=128, inter_size=11008, vocab_size=32000)
    cfg = dict(num_layers=2, **L7B)
    B, ctx, steps, max_seq = 32, 4096, 2, 4104
    eng = E.Engine(B, max_seq, dtype=pkg.F16, kv_fp8=True, **cfg)
    ref_bytes = 2 * 2 * B * 32 * max_seq * 128 * 2
    assert eng.kv_bytes < 0.53 * ref_bytes, "fp8 cache must be about half the fp16 cache"
    eng.init_random(seed=77, std=0.02)
    eng.fill_kv_random(ctx, seed=5, amp=0.5)
    ora = omodel.OracleModel(download_weights(E, eng, 2), cfg, B, max_seq, "fp16")
    ora.kv_round = emulate_fp8_rows
    kc, vc = eng.kv_caches()
    for l in range(2):
        ora.k[l] = kc[l].float().cpu().numpy()
        ora.v[l] = vc[l].float().cpu().numpy()
    del kc, vc
    ora.step = ctx
    ids = np.random.default_rng(3).integers(0, 32000, B).astype(np.int32)
    for s in range(steps):
        eng.decode_step_host(ids)
        want_logits, want_ids = ora.decode_step(ids)
        assert_close(eng.logits().float().cpu().numpy(), want_logits, "fp16", f"cfg4 fp8 cache: logits step {s}")
        ids = want_ids
    assert eng.step == ctx + steps
    eng.close()
