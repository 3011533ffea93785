"""Ragged batches: one `step` per sequence instead of the reference's single host scalar (llama.h).
The oracle has no ragged decode of its own, and needs none: sequences of a batch share nothing but the weights, so
the expected result of sequence b in a ragged batch is the oracle's result for that sequence run ALONE (batch 1).
  * b200_rope_decode_ragged / b200_decode_attention_ragged against oracle rope_decode / decode_attention per sequence;
  * b200_engine_prefill_ragged (+ a second ragged round on top of ragged histories) and the decode steps that follow
    against per-sequence OracleModels: logits, the written cache rows, untouched rows bit-exact zero.
Tolerances: rel 1e-5 fp32 / 2e-2 fp16 (north_star)."""
import numpy as np
import pytest
import torch

from conftest import assert_close
from oracle import cpu, model as omodel
from test_engine_gpu import SMALL, download_weights
from test_prefill_gpu import oracle_prefill

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def ops(pkg):
    assert torch.cuda.is_available()
    return pkg.ops


@pytest.mark.parametrize("kind,h,hk,d,steps", [("fp16", 8, 1, 128, [1, 700, 4096, 65]), ("fp16", 4, 4, 128, [33, 2]),
                                               ("fp32", 4, 2, 64, [65, 5, 300]), ("fp32", 2, 2, 4, [3, 9])])
def test_rope_and_decode_attention_per_sequence_steps(ops, kind, h, hk, d, steps):
    rng = np.random.default_rng(sum(steps))
    B, max_seq = len(steps), max(steps) + 3
    tdt = torch.float16 if kind == "fp16" else torch.float32
    rt = (lambda a: a.astype(np.float16).astype(np.float32)) if kind == "fp16" else (lambda a: a)
    qkv = rt(rng.uniform(-1, 1, (B, h + 2 * hk, d)).astype(np.float32))
    kc = rt(rng.uniform(-1, 1, (1, B, hk, max_seq, d)).astype(np.float32))
    vc = rt(rng.uniform(-1, 1, (1, B, hk, max_seq, d)).astype(np.float32))
    steps_dev = torch.tensor(steps, dtype=torch.int32, device="cuda")
    g_qkv = torch.from_numpy(qkv).cuda().to(tdt)
    g_k, g_v = torch.from_numpy(kc).cuda().to(tdt), torch.from_numpy(vc).cuda().to(tdt)
    g_out = torch.full((B, h * d), float("nan"), dtype=tdt, device="cuda")
    ops.rope_decode_ragged(g_qkv, steps_dev, d, 10000.0, hk)
    # oracle: each sequence alone, with its own step
    want_qkv = qkv.copy()
    for b, st in enumerate(steps):
        one = np.ascontiguousarray(want_qkv[b:b + 1])
        cpu.rope_decode(one, h, hk, st, d, 10000.0)
        want_qkv[b] = one[0]
    assert_close(g_qkv.float().cpu().numpy(), want_qkv, kind, "rope, per-sequence positions")
    roped = g_qkv.float().cpu().numpy()   # feed the attention oracle exactly what the GPU attention reads
    ops.decode_attention_ragged(g_qkv, 0, g_k, g_v, steps_dev, max(steps), g_out)
    torch.cuda.synchronize()
    want = np.zeros((B, h * d), np.float32)
    wk, wv = kc.copy(), vc.copy()
    for b, st in enumerate(steps):
        kb, vb = np.ascontiguousarray(wk[0, b:b + 1]), np.ascontiguousarray(wv[0, b:b + 1])
        ob = np.zeros((1, h * d), np.float32)
        cpu.decode_attention(np.ascontiguousarray(roped[b:b + 1]), kb, vb, ob, h, hk, st, True)
        want[b], wk[0, b], wv[0, b] = ob[0], kb[0], vb[0]
    assert_close(g_out.float().cpu().numpy(), want, kind, "decode attention, per-sequence key counts")
    # the append went to row step_b - 1 of each sequence and nowhere else: bit-exact copy
    assert np.array_equal(g_k.float().cpu().numpy(), wk) and np.array_equal(g_v.float().cpu().numpy(), wv)


def _check_against_single_sequence_oracles(pkg, kind, rounds, n_decode, fused=True):
    E = pkg.engine
    B = len(rounds[0])
    max_seq = max(sum(r[b] for r in rounds) for b in range(B)) + n_decode + 4
    eng = E.Engine(B, max_seq, dtype=pkg.F16 if kind == "fp16" else pkg.F32, **SMALL)
    eng.init_random(seed=11, std=0.05)
    eng.set_mode(fused=fused, graph=True, persistent=fused)   # a ragged batch never takes the persistent kernel
    w = download_weights(E, eng, SMALL["num_layers"])
    oras = [omodel.OracleModel(w, SMALL, 1, max_seq, kind) for _ in range(B)]
    rng = np.random.default_rng(B * 1000 + rounds[0][0])
    cached = np.zeros(B, np.int64)
    band = lambda lg: 4 * (2e-2 if kind == "fp16" else 1e-5) * np.max(np.abs(lg))
    for lens in rounds:
        prompts = [rng.integers(0, SMALL["vocab_size"], n).astype(np.int32) for n in lens]
        got_ids = eng.prefill_ragged(prompts)
        want_logits, want_ids = [], []
        for b in range(B):
            lg, top = oracle_prefill(oras[b], w, SMALL, prompts[b][None, :], int(cached[b]))
            want_logits.append(lg[0]); want_ids.append(top[0])
        cached += np.array(lens)
        want_logits = np.stack(want_logits)
        assert list(eng.seq_steps) == list(cached) and eng.step == cached.max()
        assert_close(eng.logits().float().cpu().numpy(), want_logits, kind, f"ragged prefill logits {lens}")
        kc, vc = (t.float().cpu().numpy() for t in eng.kv_caches())
        for b in range(B):
            assert_close(kc[:, b], oras[b].k[:, 0], kind, f"k cache of sequence {b}")
            assert_close(vc[:, b], oras[b].v[:, 0], kind, f"v cache of sequence {b}")
            assert np.all(kc[:, b, :, cached[b]:] == 0) and np.all(vc[:, b, :, cached[b]:] == 0), \
                "rows beyond a sequence's own length must be untouched"
        srt = np.sort(want_logits, axis=1)
        for b in range(B):
            if srt[b, -1] - srt[b, -2] > band(want_logits):
                assert got_ids[b] == want_ids[b]
    ids = np.array(want_ids, np.int32)
    for s in range(n_decode):   # teacher forced with the oracles' ids
        eng.decode_step_host(ids)
        rows = [oras[b].decode_step(ids[b:b + 1]) for b in range(B)]
        want_logits = np.stack([r[0][0] for r in rows])
        assert_close(eng.logits().float().cpu().numpy(), want_logits, kind, f"ragged decode step {s}")
        ids = np.array([r[1][0] for r in rows], np.int32)
        cached += 1
        assert list(eng.seq_steps) == list(cached)
    kc = eng.kv_caches()[0].float().cpu().numpy()
    for b in range(B):
        assert_close(kc[:, b], oras[b].k[:, 0], kind, f"k cache of sequence {b} after decode")
    eng.close()


@pytest.mark.parametrize("fused", [True, False], ids=["fused_gemv", "unfused"])
@pytest.mark.parametrize("kind,rounds", [("fp16", [[37, 5, 140]]), ("fp16", [[130, 1], [4, 150]]),
                                         ("fp16", [[3, 40, 7, 1, 22, 9]]),   # > 4 sequences: tcgen05 linears in both modes
                                         ("fp32", [[9, 3, 14]]), ("fp32", [[6, 2], [1, 11]])])
def test_ragged_prefill_then_decode_vs_single_sequence_oracles(pkg, kind, rounds, fused):
    """fused: the GEMV epilogue rotates and appends each token at its own position (b200_gemv_args.step_stride);
    unfused: b200_rope_decode_ragged + b200_decode_attention_ragged in the reference's launch order"""
    _check_against_single_sequence_oracles(pkg, kind, rounds, n_decode=3, fused=fused)


def test_equal_lengths_stay_on_the_shared_step_path(pkg):
    """prefill_ragged with equal lengths == prefill: same logits bit for bit, and the engine keeps the fused decode path"""
    E = pkg.engine
    B, n = 3, 40
    rng = np.random.default_rng(4)
    ids = rng.integers(0, SMALL["vocab_size"], (B, n)).astype(np.int32)
    outs = []
    for ragged in (False, True):
        eng = E.Engine(B, n + 8, dtype=pkg.F16, **SMALL)
        eng.init_random(seed=3, std=0.05)
        first = eng.prefill_ragged(list(ids)) if ragged else eng.prefill(ids)
        lg0 = eng.logits().float().cpu().numpy().copy()
        nxt = eng.decode_step_host(None)
        outs.append((first, lg0, nxt, eng.logits().float().cpu().numpy().copy(), eng.launches_per_step))
        eng.close()
    for a, b in zip(outs[0][:4], outs[1][:4]):
        assert np.array_equal(a, b)
    assert outs[0][4] == outs[1][4]


def test_reset_returns_to_one_step(pkg):
    E = pkg.engine
    eng = E.Engine(2, 64, dtype=pkg.F16, **SMALL)
    eng.init_random(seed=3, std=0.05)
    eng.prefill_ragged([np.arange(5, dtype=np.int32), np.arange(9, dtype=np.int32)])
    assert list(eng.seq_steps) == [5, 9] and eng.step == 9
    eng.reset()
    assert list(eng.seq_steps) == [0, 0] and eng.step == 0
    # a zero-length entry sits the pass out (continuous batching, tests/test_slot_refill_gpu.py); no tokens at all is an error
    eng.prefill_ragged([np.arange(5, dtype=np.int32), np.zeros(0, np.int32)])
    assert list(eng.seq_steps) == [5, 0]
    eng.reset()
    with pytest.raises(RuntimeError):
        eng.prefill_ragged([np.zeros(0, np.int32), np.zeros(0, np.int32)])
    with pytest.raises(RuntimeError):
        eng.prefill_ragged([np.arange(5, dtype=np.int32), np.arange(64, dtype=np.int32)])
    eng.close()


def test_generate_ragged_equals_each_prompt_generated_alone(pkg):
    """b200_engine_generate_ragged on three prompts of different lengths == b200_engine_generate on each prompt
    alone (batch 1): sequences share nothing but the weights. Token ids are compared where the single-sequence
    run's top-1 margin is clear of the fp16 band; a sequence is followed only up to its first ambiguous token."""
    E = pkg.engine
    rng = np.random.default_rng(21)
    prompts = [rng.integers(3, SMALL["vocab_size"], n).astype(np.int32) for n in (12, 45, 3)]
    n_new, max_seq = 12, 80
    eng = E.Engine(3, max_seq, dtype=pkg.F16, **SMALL)
    eng.init_random(seed=17, std=0.05)
    got, cnt = eng.generate_ragged(prompts, n_new, eos_id=-1)
    assert list(cnt) == [n_new] * 3 and list(eng.seq_steps) == [len(p) + n_new - 1 for p in prompts]
    eng.close()
    for b, p in enumerate(prompts):
        one = E.Engine(1, max_seq, dtype=pkg.F16, **SMALL)
        one.init_random(seed=17, std=0.05)
        first = one.prefill(p[None, :])
        toks = [int(first[0])]
        for t in range(n_new):
            lg = np.sort(one.logits().float().cpu().numpy()[0])
            if toks[-1] != got[b, t] :
                assert lg[-1] - lg[-2] <= 8e-2 * np.max(np.abs(lg)), f"sequence {b} token {t}: {toks[-1]} vs {got[b, t]}"
                break
            if t + 1 < n_new:
                toks.append(int(one.decode_step_host(None)[0]))
        one.close()
