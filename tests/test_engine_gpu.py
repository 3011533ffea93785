"""GPU parity of the decode-step driver (include/b200_engine.h) against the CPU oracle of the
whole decoder stack (oracle/model.py), in all three execution modes: fused + CUDA graph (the
product path), fused without graph, and the unfused reference launch sequence.
Covers BASELINE.json configs[0] (one LLaMA-2-7B-shaped fp32 layer, batch 1, 64-token cache) and a
teacher-forced greedy token-ID check on a small LLaMA-shaped model."""
import numpy as np
import pytest
import torch

from conftest import assert_close, rel_err
from oracle import model as omodel

pytestmark = pytest.mark.gpu

SMALL = dict(num_layers=3, head_num=4, kv_head_num=2, head_size=128, inter_size=768, vocab_size=1000)
# down_proj rows longer than one ring stage (K = 9216 -> chunks of 4096 + 4096 + 1024), h/hk = 1, odd vocab pairs
WIDE = dict(num_layers=2, head_num=4, kv_head_num=4, head_size=128, inter_size=9216, vocab_size=2999)


def download_weights(E, eng, n_layers):
    f = lambda a: np.ascontiguousarray(a.astype(np.float32))
    layers = []
    for l in range(n_layers):
        layers.append(dict(attn_norm=f(eng.get_weight(E.W_ATTN_NORM, l)), qkv=f(eng.get_weight(E.W_QKV, l)),
                           o=f(eng.get_weight(E.W_O, l)), ffn_norm=f(eng.get_weight(E.W_FFN_NORM, l)),
                           gate_up=f(eng.get_weight(E.W_GATE_UP, l)), down=f(eng.get_weight(E.W_DOWN, l))))
    return dict(layers=layers, embed=f(eng.get_weight(E.W_EMBED)), final_norm=f(eng.get_weight(E.W_FINAL_NORM)),
                lm_head=f(eng.get_weight(E.W_LM_HEAD)))


@pytest.mark.parametrize("kind", ["fp32", "fp16"])
@pytest.mark.parametrize("mode", ["persistent", "fused_graph", "fused", "unfused"])
def test_small_model_teacher_forced(pkg, kind, mode):
    if mode == "persistent" and kind == "fp32":
        pytest.skip("the persistent whole-step kernel is fp16 only (fp32 runs the multi-kernel path)")
    E = pkg.engine
    B, max_seq, steps = 2, 32, 8
    eng = E.Engine(B, max_seq, dtype=pkg.F16 if kind == "fp16" else pkg.F32, **SMALL)
    eng.init_random(seed=5, std=0.05)
    eng.set_mode(fused=mode != "unfused", graph=mode in ("fused_graph", "persistent"), persistent=mode == "persistent")
    ora = omodel.OracleModel(download_weights(E, eng, SMALL["num_layers"]), SMALL, B, max_seq, kind)
    rng = np.random.default_rng(0)
    ids = rng.integers(0, SMALL["vocab_size"], B).astype(np.int32)
    agree = total = 0
    for s in range(steps):
        got_ids = eng.decode_step_host(ids)
        want_logits, want_ids = ora.decode_step(ids)
        got_logits = eng.logits().float().cpu().numpy()
        assert_close(got_logits, want_logits, kind, f"logits step {s} ({mode})")
        # greedy pick must be identical wherever the oracle's top-2 margin exceeds the tolerance band
        srt = np.sort(want_logits, axis=1)
        margin = srt[:, -1] - srt[:, -2]
        band = 4 * (2e-2 if kind == "fp16" else 1e-5) * np.max(np.abs(want_logits))
        for b in range(B):
            if margin[b] > band:
                total += 1
                agree += int(got_ids[b] == want_ids[b])
        ids = want_ids   # teacher forcing: both sides continue from the oracle's token
    assert eng.step == steps
    if mode == "persistent":
        assert eng.launches_per_step == 1, "the persistent path must be ONE launch per step"
    assert agree == total and total > 0, f"greedy ids differ: {agree}/{total}"
    kc, vc = eng.kv_caches()
    assert_close(kc[:, :, :, :steps].float().cpu().numpy(), ora.k[:, :, :, :steps], kind, "K cache")
    assert_close(vc[:, :, :, :steps].float().cpu().numpy(), ora.v[:, :, :, :steps], kind, "V cache")
    eng.close()


def test_modes_agree_and_graph_replays(pkg):
    """Graph replay is bit-identical to stream launches over 150 free-running greedy steps (several
    split-KV buckets -> several graphs, device-side step keeps tracking), and the fused path agrees
    with the unfused reference launch sequence when both are fed the same tokens."""
    E = pkg.engine
    B, max_seq = 3, 200
    outs = {}
    for mode in ("fused_graph", "fused"):
        eng = E.Engine(B, max_seq, dtype=pkg.F16, **SMALL)
        eng.init_random(seed=9, std=0.05)
        eng.set_mode(fused=True, graph=mode == "fused_graph")
        ids = np.array([1, 2, 3], np.int32)
        toks = []
        for s in range(150):
            ids = eng.decode_step_host(ids if s == 0 else None)
            toks.append(ids.copy())
        outs[mode] = (np.stack(toks), eng.logits().float().cpu().numpy().copy(), eng.launches_per_step)
        eng.close()
    assert np.array_equal(outs["fused"][0], outs["fused_graph"][0]), "graph replay changed the tokens"
    assert rel_err(outs["fused"][1], outs["fused_graph"][1]) == 0.0, "graph replay must be bit-identical"
    # fused vs unfused on the trajectory the fused engine produced (teacher forcing)
    traj = outs["fused_graph"][0]
    eng = E.Engine(B, max_seq, dtype=pkg.F16, **SMALL)
    eng.init_random(seed=9, std=0.05)
    eng.set_mode(fused=False, graph=False)
    ids = np.array([1, 2, 3], np.int32)
    same = 0
    for s in range(150):
        out = eng.decode_step_host(ids)
        same += int(np.array_equal(out, traj[s]))
        ids = traj[s]
    assert_close(eng.logits().float().cpu().numpy(), outs["fused_graph"][1], "fp16", "fused vs unfused logits")
    assert same >= 140, f"fused and unfused greedy picks agree on only {same}/150 steps"
    assert outs["fused_graph"][2] < eng.launches_per_step, "fusion must cut the launches per step"
    eng.close()


def test_config0_single_7b_layer_fp32(pkg):
    """BASELINE.json configs[0]: one LLaMA-2-7B-shaped decoder layer (hidden 4096, 32 heads,
    head_dim 128), fp32 random weights, batch 1, one decode step on a 64-token KV cache, checked
    against the CPU oracle at rel err <= 1e-5."""
    E = pkg.engine
    shape = dict(num_layers=1, head_num=32, kv_head_num=32, head_size=128, inter_size=11008, vocab_size=32000)
    for mode in ("fused_graph", "unfused"):
        eng = E.Engine(1, 128, dtype=pkg.F32, **shape)
        eng.init_random(seed=1234, std=0.02)
        eng.fill_kv_random(64, seed=3, amp=0.5)
        eng.set_mode(fused=mode != "unfused", graph=mode == "fused_graph")
        ora = omodel.OracleModel(download_weights(E, eng, 1), shape, 1, 128, "fp32")
        kc, vc = eng.kv_caches()
        ora.k[:] = kc.cpu().numpy()
        ora.v[:] = vc.cpu().numpy()
        ora.step = 64
        ids = np.array([4242], np.int32)
        got_ids = eng.decode_step_host(ids)
        want_logits, want_ids = ora.decode_step(ids)
        assert_close(eng.hidden().cpu().numpy(), ora.hidden, "fp32", f"layer output ({mode})")
        assert_close(eng.logits().cpu().numpy(), want_logits, "fp32", f"logits ({mode})")
        assert got_ids[0] == want_ids[0]
        assert eng.step == 65
        eng.close()


@pytest.mark.parametrize("B,graph", [(12, True), (40, False), (32, True)])
def test_batched_decode_uses_tcgen05_linears(pkg, B, graph):
    """Batches above the GEMV tile (8) run the reference launch sequence with every linear on the
    tcgen05 weight-streaming GEMM (weights read once per step); teacher-forced against the oracle,
    with the KV cache pre-filled so the device-side step / split-KV geometry is exercised."""
    E = pkg.engine
    max_seq, steps, ctx = 96, 4, 70
    eng = E.Engine(B, max_seq, dtype=pkg.F16, **SMALL)
    eng.init_random(seed=21, std=0.05)
    eng.fill_kv_random(ctx, seed=4, amp=0.5)
    eng.set_mode(fused=True, graph=graph)
    ora = omodel.OracleModel(download_weights(E, eng, SMALL["num_layers"]), SMALL, B, max_seq, "fp16")
    kc, vc = eng.kv_caches()
    ora.k[:] = kc.float().cpu().numpy()
    ora.v[:] = vc.float().cpu().numpy()
    ora.step = ctx
    ids = np.random.default_rng(1).integers(0, SMALL["vocab_size"], B).astype(np.int32)
    before = pkg.capi.lib().b200_launch_count()
    for s in range(steps):
        eng.decode_step_host(ids)
        want_logits, want_ids = ora.decode_step(ids)
        assert_close(eng.logits().float().cpu().numpy(), want_logits, "fp16", f"logits step {s} (B={B})")
        ids = want_ids
    assert pkg.capi.lib().b200_launch_count() > before
    kc, vc = eng.kv_caches()
    assert_close(kc[:, :, :, ctx:ctx + steps].float().cpu().numpy(), ora.k[:, :, :, ctx:ctx + steps], "fp16", "K rows")
    eng.close()


@pytest.mark.parametrize("cfg_name,B,ctx", [("small", 1, 0), ("small", 1, 300), ("small", 3, 130), ("small", 4, 64),
                                            ("small", 2, 1000), ("wide", 1, 40), ("wide", 2, 200)])
def test_persistent_step_kernel_vs_oracle_with_context(pkg, cfg_name, B, ctx):
    """The persistent whole-step kernel (ONE cooperative launch per token) teacher-forced against
    the CPU oracle from a pre-filled KV cache, so that several split-KV geometries (1..4+ splits,
    merged in the o_proj prologue), the device-side step and the grid barriers are exercised."""
    E = pkg.engine
    SMALL = {"small": globals()["SMALL"], "wide": WIDE}[cfg_name]
    max_seq, steps = ctx + 16, 6
    eng = E.Engine(B, max_seq, dtype=pkg.F16, **SMALL)
    eng.init_random(seed=33, std=0.05)
    eng.set_mode(fused=True, graph=True, persistent=True)   # the default for batch 1; batches 2-4 on request
    if ctx:
        eng.fill_kv_random(ctx, seed=5, amp=0.5)
    ora = omodel.OracleModel(download_weights(E, eng, SMALL["num_layers"]), SMALL, B, max_seq, "fp16")
    kc, vc = eng.kv_caches()
    ora.k[:] = kc.float().cpu().numpy()
    ora.v[:] = vc.float().cpu().numpy()
    ora.step = ctx
    ids = np.random.default_rng(2).integers(0, SMALL["vocab_size"], B).astype(np.int32)
    agree = total = 0
    for s in range(steps):
        got_ids = eng.decode_step_host(ids)
        want_logits, want_ids = ora.decode_step(ids)
        got_logits = eng.logits().float().cpu().numpy()
        assert_close(got_logits, want_logits, "fp16", f"logits step {s} (B={B}, ctx={ctx})")
        srt = np.sort(want_logits, axis=1)
        band = 4 * 2e-2 * np.max(np.abs(want_logits))
        for b in range(B):
            if srt[b, -1] - srt[b, -2] > band:
                total += 1
                agree += int(got_ids[b] == want_ids[b])
        ids = want_ids
    assert eng.launches_per_step == 1
    assert eng.step == ctx + steps
    assert agree == total
    kc, vc = eng.kv_caches()
    assert_close(kc[:, :, :, ctx:ctx + steps].float().cpu().numpy(), ora.k[:, :, :, ctx:ctx + steps], "fp16", "K rows")
    assert_close(vc[:, :, :, ctx:ctx + steps].float().cpu().numpy(), ora.v[:, :, :, ctx:ctx + steps], "fp16", "V rows")
    eng.close()


@pytest.mark.parametrize("B,ctx", [(1, 0), (1, 300), (2, 130)])
def test_persistent_step_kernel_exchange_form_single_rank(pkg, B, ctx):
    """The tensor-parallel form of the persistent step kernel (kOpExchange: the K-split accumulators are pushed as
    low-latency lines into the ranks' receive buffers, summed, and the residual added) on ONE GPU: a world = 1
    communicator attached to an unsharded engine exercises the wire format, call tags, buffer parity and the device
    call counter over several steps; the result must match the oracle like the plain form does."""
    import ctypes as C
    E = pkg.engine
    lib = E._lib()
    max_seq, steps = ctx + 16, 6
    eng = E.Engine(B, max_seq, dtype=pkg.F16, **SMALL)
    eng.init_random(seed=33, std=0.05)
    H = SMALL["head_num"] * SMALL["head_size"]
    comm = lib.b200_p2p_create(1, 0, B * H * 2)
    assert comm
    hnd = C.create_string_buffer(int(lib.b200_p2p_handle_bytes()))
    assert lib.b200_p2p_local_handle(comm, hnd) == 0
    assert lib.b200_p2p_connect(comm, hnd.raw) == 0
    assert lib.b200_engine_set_p2p(eng.h, comm) == 0
    eng.set_mode(fused=True, graph=True, persistent=True)
    if ctx:
        eng.fill_kv_random(ctx, seed=5, amp=0.5)
    ora = omodel.OracleModel(download_weights(E, eng, SMALL["num_layers"]), SMALL, B, max_seq, "fp16")
    kc, vc = eng.kv_caches()
    ora.k[:] = kc.float().cpu().numpy()
    ora.v[:] = vc.float().cpu().numpy()
    ora.step = ctx
    ids = np.random.default_rng(2).integers(0, SMALL["vocab_size"], B).astype(np.int32)
    for s in range(steps):
        eng.decode_step_host(ids)
        want_logits, want_ids = ora.decode_step(ids)
        assert_close(eng.logits().float().cpu().numpy(), want_logits, "fp16", f"exchange form: logits step {s}")
        ids = want_ids
    assert eng.launches_per_step == 1
    kc, vc = eng.kv_caches()
    assert_close(kc[:, :, :, ctx:ctx + steps].float().cpu().numpy(), ora.k[:, :, :, ctx:ctx + steps], "fp16", "K rows")
    assert lib.b200_engine_set_p2p(eng.h, None) == 0
    eng.close()
    lib.b200_p2p_destroy(comm)


def test_persistent_step_kernel_70b_geometry(pkg):
    """hidden 8192 (four 16-byte packets per thread in the K-split consumer), 64 q / 8 kv heads (groups of 8): the
    geometry of BASELINE configs[4], two layers with a narrow FFN, plain and exchange form, vs the oracle."""
    import ctypes as C
    E = pkg.engine
    lib = E._lib()
    cfg = dict(num_layers=2, head_num=64, kv_head_num=8, head_size=128, inter_size=2368, vocab_size=1000)
    B, ctx, steps = 1, 200, 3
    for form in ("plain", "exchange"):
        eng = E.Engine(B, ctx + 8, dtype=pkg.F16, **cfg)
        eng.init_random(seed=3, std=0.02)
        comm = None
        if form == "exchange":
            comm = lib.b200_p2p_create(1, 0, B * 8192 * 2)
            hnd = C.create_string_buffer(int(lib.b200_p2p_handle_bytes()))
            assert lib.b200_p2p_local_handle(comm, hnd) == 0 and lib.b200_p2p_connect(comm, hnd.raw) == 0
            assert lib.b200_engine_set_p2p(eng.h, comm) == 0
        eng.set_mode(fused=True, graph=True, persistent=True)
        eng.fill_kv_random(ctx, seed=5, amp=0.5)
        ora = omodel.OracleModel(download_weights(E, eng, 2), cfg, B, ctx + 8, "fp16")
        kc, vc = eng.kv_caches()
        ora.k[:] = kc.float().cpu().numpy()
        ora.v[:] = vc.float().cpu().numpy()
        ora.step = ctx
        ids = np.array([17], np.int32)
        for s in range(steps):
            eng.decode_step_host(ids)
            want_logits, want_ids = ora.decode_step(ids)
            assert_close(eng.logits().float().cpu().numpy(), want_logits, "fp16", f"70B geometry ({form}): logits step {s}")
            ids = want_ids
        assert eng.launches_per_step == 1
        if comm:
            assert lib.b200_engine_set_p2p(eng.h, None) == 0
        eng.close()
        if comm:
            lib.b200_p2p_destroy(comm)
