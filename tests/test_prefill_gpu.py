"""b200_engine_prefill (the prompt pass, Llama<T>::firstTokenGen llama.cpp:274-318: embedding ->
LlamaContextDecoder layers -> final norm -> LM head on the last token -> top-1) against the CPU
oracle's prefill loop (oracle/model.py prefill_layers = context_decoder.cpp:48-140), then decode
steps continuing from the prefilled cache against the oracle's decode loop. fp16 runs the tcgen05
GEMMs + the flash context attention; fp32 the reference-order GPU sequence. Tolerances: rel 1e-5
fp32 / 2e-2 fp16 (north_star); cache rows beyond the prompt stay bit-exact (untouched)."""
import numpy as np
import pytest

from conftest import assert_close
from oracle import cpu, model as omodel
from test_engine_gpu import SMALL, download_weights

pytestmark = pytest.mark.gpu


def oracle_prefill(ora, w, cfg, ids, hist):
    """ids [B, len] -> (logits of the last token [B, V], top-1 ids [B]); advances ora.step"""
    B, n = ids.shape
    H = cfg["head_num"] * cfg["head_size"]
    x = np.zeros((B * n, H), np.float32)
    cpu.embedding(np.ascontiguousarray(ids.reshape(-1), dtype=np.int32), w["embed"], x)
    x = ora.prefill_layers(x, [n] * B, [hist] * B)
    ora.step = hist + n
    last = np.ascontiguousarray(x.reshape(B, n, H)[:, -1])
    cpu.rmsnorm(last, None, w["final_norm"], 1e-5)
    last = ora.rt(last)
    logits = np.zeros((B, cfg["vocab_size"]), np.float32)
    cpu.linear(last, w["lm_head"], logits, True)
    logits = ora.rt(logits)
    top = np.zeros((B, 1), np.int32)
    cpu.topk(logits, top, None, 1)
    return logits, top.reshape(B)


@pytest.mark.parametrize("kind,B,lens", [("fp16", 2, [40]), ("fp16", 1, [300]), ("fp16", 2, [130, 21]),
                                         ("fp32", 2, [19]), ("fp32", 1, [9, 5])])
def test_prefill_then_decode_vs_oracle(pkg, kind, B, lens):
    E = pkg.engine
    max_seq = sum(lens) + 8
    eng = E.Engine(B, max_seq, dtype=pkg.F16 if kind == "fp16" else pkg.F32, **SMALL)
    eng.init_random(seed=9, std=0.05)
    w = download_weights(E, eng, SMALL["num_layers"])
    ora = omodel.OracleModel(w, SMALL, B, max_seq, kind)
    rng = np.random.default_rng(len(lens) * 100 + lens[0])
    hist = 0
    for n in lens:   # several rounds: the later ones see the earlier ones as history (multi-round chat)
        ids = rng.integers(0, SMALL["vocab_size"], (B, n)).astype(np.int32)
        got_ids = eng.prefill(ids)
        want_logits, want_ids = oracle_prefill(ora, w, SMALL, ids, hist)
        hist += n
        assert eng.step == hist
        assert_close(eng.logits().float().cpu().numpy(), want_logits, kind, f"prefill logits (round of {n})")
        kc, vc = eng.kv_caches()
        assert_close(kc.float().cpu().numpy(), ora.k, kind, "k cache after prefill")
        assert_close(vc.float().cpu().numpy(), ora.v, kind, "v cache after prefill")
        assert np.all(kc[:, :, :, hist:].float().cpu().numpy() == 0), "rows beyond the prompt must be untouched"
        srt = np.sort(want_logits, axis=1)
        band = 4 * (2e-2 if kind == "fp16" else 1e-5) * np.max(np.abs(want_logits))
        for b in range(B):
            if srt[b, -1] - srt[b, -2] > band:
                assert got_ids[b] == want_ids[b]
    # decode continues from the prefilled cache (teacher forced with the oracle's ids)
    ids = want_ids
    for s in range(3):
        got = eng.decode_step_host(ids)
        want_logits, want_ids = ora.decode_step(ids)
        assert_close(eng.logits().float().cpu().numpy(), want_logits, kind, f"decode step {s} after prefill")
        ids = want_ids
    assert eng.step == hist + 3
    eng.close()


def test_prefill_matches_token_by_token_decode(pkg):
    """size-independent property: prefilling n tokens leaves the same cache and the same next-token
    logits (within fp16 tolerance) as feeding the same n tokens through n decode steps."""
    E = pkg.engine
    B, n = 2, 150
    a = E.Engine(B, n + 4, dtype=pkg.F16, **SMALL)
    b = E.Engine(B, n + 4, dtype=pkg.F16, **SMALL)
    a.init_random(seed=3, std=0.05)
    b.init_random(seed=3, std=0.05)
    ids = np.random.default_rng(1).integers(0, SMALL["vocab_size"], (B, n)).astype(np.int32)
    a.prefill(ids)
    for t in range(n):
        b.decode_step_host(ids[:, t])
    assert a.step == b.step == n
    assert_close(a.logits().float().cpu().numpy(), b.logits().float().cpu().numpy(), "fp16", "prefill vs decode logits")
    ka, va = a.kv_caches()
    kb, vb = b.kv_caches()
    assert_close(ka.float().cpu().numpy(), kb.float().cpu().numpy(), "fp16", "k cache")
    assert_close(va.float().cpu().numpy(), vb.float().cpu().numpy(), "fp16", "v cache")
    a.close()
    b.close()


@pytest.mark.parametrize("kind", ["fp16", "fp32"])
def test_rope_to_cache_equals_transpose_then_concat(pkg, kind):
    """b200_qkv_bias_rope_to_cache = launchAddFusedQKVBiasTransposeAndRoPE + launchConcatKVCache in one pass:
    bit-exact against the two-kernel sequence (both ours; each is pinned by the oracle in test_parity_gpu)."""
    import ctypes as C
    import torch
    ops, capi = pkg.ops, pkg.capi
    dt = torch.float16 if kind == "fp16" else torch.float32
    B, h, hk, d, S, max_seq = 3, 4, 2, 128, 9, 40
    lens = np.array([5, 9, 2], np.int32)
    hist = np.array([7, 0, 30], np.int32)
    nt = int(lens.sum())
    g = torch.Generator(device="cuda").manual_seed(3)
    qkv = (torch.rand(nt, h + 2 * hk, d, device="cuda", generator=g) - 0.5).to(dt)
    bias = (torch.rand((h + 2 * hk) * d, device="cuda", generator=g) - 0.5).to(dt)
    dev = lambda a: torch.from_numpy(a).cuda()
    pad = torch.zeros(B * S, dtype=torch.int32, device="cuda")
    cum = torch.zeros(B + 1, dtype=torch.int32, device="cuda")
    ops.launchCalPaddingoffset(pad.view(B, S), cum, dev(lens))
    q1 = torch.zeros(B, h, S, d, dtype=dt, device="cuda")
    k1 = torch.zeros(B, hk, S, d, dtype=dt, device="cuda")
    v1 = torch.zeros(B, hk, S, d, dtype=dt, device="cuda")
    ops.launchAddFusedQKVBiasTransposeAndRoPE(q1, k1, v1, qkv, bias, pad, dev(hist), dev(lens), 128, 10000.0)
    kc1 = torch.full((1, B, hk, max_seq, d), 3.0, dtype=dt, device="cuda")
    vc1 = torch.full((1, B, hk, max_seq, d), 3.0, dtype=dt, device="cuda")
    ops.launchConcatKVCache(k1, v1, 0, dev(lens), dev(hist), kc1, vc1)
    q2 = torch.zeros_like(q1)
    kc2, vc2 = torch.full_like(kc1, 3.0), torch.full_like(vc1, 3.0)
    p = lambda t: C.c_void_p(t.data_ptr())
    hd = dev(hist)
    rc = capi.lib().b200_qkv_bias_rope_to_cache(p(q2), p(kc2), p(vc2), p(qkv), p(bias), p(pad), p(hd), B, S, nt, h, hk, d,
                                                max_seq, 128, 10000.0, capi.F16 if kind == "fp16" else capi.F32, None)
    torch.cuda.synchronize()
    assert rc == 0
    assert torch.equal(q1, q2) and torch.equal(kc1, kc2) and torch.equal(vc1, vc2)
