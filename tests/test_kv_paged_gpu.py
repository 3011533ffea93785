"""Paged KV cache (SURVEY section 8 f4 "paged blocks"; csrc/kv_paged.cu, b200_decode_attention_paged,
B200_ENGINE_KV_PAGED).

The reference's cache is dense, [layers, B, hk, max_seq, d] (llama.cpp:77-78): max_seq rows reserved per sequence slot.
The paged cache keeps the same rows in pages [n_pages, hk, 128, d] named by one page list per sequence. Its contract is
the strongest one available: NOTHING about the arithmetic changes, so on the same launch order the paged path must
be BIT-IDENTICAL to the dense path (attention output, appended rows, logits, greedy ids) - and the dense path is pinned
to the oracle (tests/test_decode_gpu.py, test_engine_gpu.py, test_7b_shapes_gpu.py); the kernel-level test also checks
the oracle directly. On top: the allocator (pages taken as sequences grow, returned by reset / retirement, a clean B200_ENOMEM when the pool is empty) and a pool smaller than the dense cache serving a ragged batch."""
import ctypes as C

import numpy as np
import pytest
import torch

from conftest import assert_close
from oracle import cpu
from test_engine_gpu import SMALL

pytestmark = pytest.mark.gpu

P = 128   # B200_KV_PAGE_ROWS


def _vp(t):
    return C.c_void_p(t.data_ptr())


@pytest.mark.parametrize("B,h,hk,steps", [
    (3, 8, 2, [300, 1, 129]),              # GQA 4, a sequence on its first token, one just past a page boundary
    (2, 32, 32, [777, 512]),               # 7B heads; 512 = the new row is the last row of page 3
    (1, 8, 1, [1025]),                     # GQA 8, the new row opens page 8
    (4, 4, 2, [4097, 130, 4000, 128]),     # long context: many splits
])
def test_paged_attention_bit_identical_to_dense_and_vs_oracle(pkg, B, h, hk, steps):
    lib = pkg.capi.lib()
    d = 128
    max_step = max(steps)
    max_pages = (max_step + P - 1) // P + 1
    max_seq = max_pages * P
    g = torch.Generator(device="cpu").manual_seed(max_step + B)
    k_dense = ((torch.rand(B, hk, max_seq, d, generator=g) * 2 - 1)).half().cuda()
    v_dense = ((torch.rand(B, hk, max_seq, d, generator=g) * 2 - 1)).half().cuda()
    qkv = (torch.rand(B, h + 2 * hk, d, generator=g) * 2 - 1).half().cuda()
    steps_t = torch.tensor(steps, dtype=torch.int32, device="cuda")
    hist = steps_t - 1
    # pages handed out in a scrambled order, with pages nobody owns in between
    n_pages = B * max_pages + 5
    perm = torch.randperm(n_pages, generator=g).int()
    table = torch.full((B, max_pages), -1, dtype=torch.int32)
    nxt = 0
    for b in range(B):
        for i in range((steps[b] + P - 1) // P):
            table[b, i] = perm[nxt]
            nxt += 1
    table_d = table.clamp(min=0).cuda()
    k_pool = torch.full((n_pages, hk, P, d), float("nan"), dtype=torch.float16, device="cuda")
    v_pool = torch.full_like(k_pool, float("nan"))
    for src, pool in ((k_dense, k_pool), (v_dense, v_pool)):
        assert lib.b200_kv_scatter_pages(_vp(src), _vp(pool), _vp(table_d), max_pages, _vp(hist), None, B, hk, max_seq, d, 2,
                                         max_step, None) == 0
    nbytes = lib.b200_decode_attention_workspace_bytes(B, h, d, max_seq)
    ws = torch.empty(nbytes, dtype=torch.uint8, device="cuda")
    out_p = torch.zeros(B, h * d, dtype=torch.float16, device="cuda")
    out_d = torch.zeros_like(out_p)
    rc = lib.b200_decode_attention_paged(_vp(qkv), _vp(k_pool), _vp(v_pool), _vp(table_d), max_pages, _vp(out_p), _vp(ws), nbytes,
                                         B, h, hk, d, max_step, _vp(steps_t), 1, 1, None)
    assert rc == 0
    kd, vd = k_dense.clone(), v_dense.clone()
    rc = lib.b200_decode_attention_ragged(_vp(qkv), _vp(kd), _vp(vd), _vp(out_d), _vp(ws), nbytes, B, h, hk, d, max_seq, max_step,
                                          _vp(steps_t), 1, pkg.F16, None)
    assert rc == 0
    torch.cuda.synchronize()
    assert torch.equal(out_p, out_d), "paged attention must be bit-identical to the dense kernel"
    # the pages hold exactly the dense rows, the appended one included, and nothing else was written
    kg = torch.zeros_like(k_dense)
    vg = torch.zeros_like(v_dense)
    for pool, dst in ((k_pool, kg), (v_pool, vg)):
        assert lib.b200_kv_gather_pages(_vp(pool), _vp(dst), _vp(table_d), max_pages, _vp(steps_t), B, hk, max_seq, d, 2, max_step,
                                        None) == 0
    torch.cuda.synchronize()
    for b in range(B):
        assert torch.equal(kg[b, :, :steps[b]], kd[b, :, :steps[b]]) and torch.equal(vg[b, :, :steps[b]], vd[b, :, :steps[b]])
        assert torch.equal(kg[b, :, steps[b] - 1], qkv[b, h:h + hk]), "appended K row"
        assert float(kg[b, :, steps[b]:].abs().max()) == 0.0
    owned = set(int(x) for x in table.flatten() if x >= 0)
    for pg in range(n_pages):
        if pg not in owned:
            assert bool(torch.isnan(k_pool[pg]).all()), "a page nobody owns was written"
    # and the oracle, one sequence at a time (its decode attention takes one shared step)
    for b in range(B):
        want = np.zeros((1, h * d), np.float32)
        cpu.decode_attention(np.ascontiguousarray(qkv[b:b + 1].float().cpu().numpy()),
                             np.ascontiguousarray(kd[b:b + 1].float().cpu().numpy()),
                             np.ascontiguousarray(vd[b:b + 1].float().cpu().numpy()), want, h, hk, steps[b], False)
        assert_close(out_p[b:b + 1].float().cpu().numpy(), want, "fp16", f"paged attention vs oracle, sequence {b} step {steps[b]}")


def test_paged_rejects_unsupported_shapes(pkg):
    lib = pkg.capi.lib()
    t = torch.zeros(1 << 16, dtype=torch.float16, device="cuda")
    tb = torch.zeros(64, dtype=torch.int32, device="cuda")
    # head size 64 and a group of 16 are outside the paged kernel: B200_EUNSUPPORTED (-2); a step beyond the list: EINVAL
    assert lib.b200_decode_attention_paged(_vp(t), _vp(t), _vp(t), _vp(tb), 4, _vp(t), None, 0, 1, 4, 4, 64, 8, None, 0, 1, None) == -2
    assert lib.b200_decode_attention_paged(_vp(t), _vp(t), _vp(t), _vp(tb), 4, _vp(t), None, 0, 1, 16, 1, 128, 8, None, 0, 1, None) == -2
    assert lib.b200_decode_attention_paged(_vp(t), _vp(t), _vp(t), _vp(tb), 4, _vp(t), None, 0, 1, 4, 4, 128, 4 * P + 1, None, 0, 1, None) == -1
    assert lib.b200_kv_scatter_pages(_vp(t), _vp(t), _vp(tb), 4, _vp(tb), None, 1, 4, 8, 100, 2, 8, None) == -2   # 200-byte rows


def _pair(pkg, B, max_seq, cfg=SMALL, pages=0, seed=9, std=0.05):
    """a paged engine and a dense engine with the same weights, both on the reference launch order (graphed)"""
    E = pkg.engine
    pe = E.Engine(B, max_seq, dtype=pkg.F16, kv_paged=True, kv_pages=pages, **cfg)
    de = E.Engine(B, max_seq, dtype=pkg.F16, **cfg)
    for eng in (pe, de):
        eng.init_random(seed=seed, std=std)
    de.set_mode(fused=False, graph=True)
    return pe, de


def _same_rows(pe, de, lens):
    kp, vp = pe.kv_caches()   # gathered copies, dense layout
    kd, vd = de.kv_caches()
    for b, n in enumerate(lens):
        assert torch.equal(kp[:, b, :, :n], kd[:, b, :, :n]) and torch.equal(vp[:, b, :, :n], vd[:, b, :, :n]), f"rows of sequence {b}"


@pytest.mark.parametrize("B,lens", [(2, [40, 40]), (3, [130, 5, 128]), (6, [17, 200, 1, 64, 127, 129])])
def test_engine_paged_equals_dense_prefill_then_decode(pkg, B, lens):
    """prompt pass (ragged where the lengths differ) + decode steps that cross page boundaries: logits, ids and cached
    rows bit-identical to the dense engine (batch <= 4: GEMV linears, batch 6: tcgen05 linears)"""
    max_seq, steps = 300, 6
    pe, de = _pair(pkg, B, max_seq)
    rng = np.random.default_rng(sum(lens))
    prompts = [rng.integers(0, SMALL["vocab_size"], n).astype(np.int32) for n in lens]
    if len(set(lens)) == 1:
        ip, idd = pe.prefill(np.stack(prompts)), de.prefill(np.stack(prompts))
    else:
        ip, idd = pe.prefill_ragged(prompts), de.prefill_ragged(prompts)
    assert np.array_equal(ip, idd)
    assert torch.equal(pe.logits(), de.logits()), "prefill logits"
    _same_rows(pe, de, lens)
    assert pe.kv_pages_total - pe.kv_pages_free == sum((n + P - 1) // P for n in lens)
    for s in range(steps):
        ip, idd = pe.decode_step_host(None), de.decode_step_host(None)
        assert torch.equal(pe.logits(), de.logits()), f"decode step {s}"
        assert np.array_equal(ip, idd)
    _same_rows(pe, de, [n + steps for n in lens])
    assert pe.kv_pages_total - pe.kv_pages_free == sum((n + steps + P - 1) // P for n in lens)
    # a second prompt pass appends behind the history: the context attention reads the history rows through the page lists
    more = [rng.integers(0, SMALL["vocab_size"], 3 + 10 * b).astype(np.int32) for b in range(B)]
    assert np.array_equal(pe.prefill_ragged(more), de.prefill_ragged(more))
    assert torch.equal(pe.logits(), de.logits()), "second prompt pass over a paged history"
    _same_rows(pe, de, [n + steps + len(m) for n, m in zip(lens, more)])
    for s in range(2):
        assert np.array_equal(pe.decode_step_host(None), de.decode_step_host(None))
        assert torch.equal(pe.logits(), de.logits()), f"decode step {s} after the second prompt pass"
    pe.close(), de.close()


def test_small_pool_serves_a_ragged_batch_and_fails_cleanly_when_empty(pkg):
    """4 sequence slots x 1024 rows would be 32 pages dense; 10 pages serve prompts of 70 / 300 / 50 / 600 rows
    (1 + 3 + 1 + 5). The pool is exhausted exactly when a sequence needs a page nobody can give."""
    E = pkg.engine
    B, max_seq = 4, 1024
    lens = [70, 300, 50, 600]
    pe, de = _pair(pkg, B, max_seq, pages=10)
    dense_bytes = de.kv_bytes
    assert pe.kv_bytes * 3 < dense_bytes, "10 of 32 pages: under a third of the dense cache"
    rng = np.random.default_rng(1)
    prompts = [rng.integers(0, SMALL["vocab_size"], n).astype(np.int32) for n in lens]
    assert np.array_equal(pe.prefill_ragged(prompts), de.prefill_ragged(prompts))
    assert pe.kv_pages_free == 0
    tab = pe.kv_page_table().cpu().numpy()
    owned = np.concatenate([tab[b, :(n + P - 1) // P] for b, n in enumerate(lens)])
    assert sorted(owned.tolist()) == list(range(10)), "every page has exactly one owner"
    # sequence 3 (600 rows) fills its fifth page at row 639, the others have more room: 40 steps fit ...
    for s in range(40):
        assert np.array_equal(pe.decode_step_host(None), de.decode_step_host(None)), f"step {s}"
    assert torch.equal(pe.logits(), de.logits())
    # ... the 41st appends row 640 of sequence 3 = a sixth page: B200_ENOMEM, nothing launched, engine still usable
    with pytest.raises(RuntimeError, match="page pool exhausted"):
        pe.decode_step_host(None)
    assert pe.step == 640
    pe.reset()
    assert pe.kv_pages_free == 10 and pe.step == 0
    ids = np.arange(4).astype(np.int32)
    de.reset()
    assert np.array_equal(pe.decode_step_host(ids), de.decode_step_host(ids))
    # prompts that cannot fit the pool at all (4 x 4 pages): refused, and the engine is left empty and consistent
    pe.reset()
    with pytest.raises(RuntimeError, match="page pool exhausted"):
        pe.prefill(rng.integers(0, SMALL["vocab_size"], (B, 400)).astype(np.int32))
    assert pe.step == 0   # nothing ran; the pages the first sequences took before the failure go back with a reset
    pe.reset()
    assert pe.kv_pages_free == 10
    de.reset()
    assert np.array_equal(pe.prefill_ragged(prompts), de.prefill_ragged(prompts))
    assert torch.equal(pe.logits(), de.logits())
    pe.close(), de.close()


def test_retired_sequences_return_their_pages(pkg):
    """generate() with an eos that one sequence hits early: same tokens as the dense engine, and the retired sequence's
    pages are back in the pool while the others still decode"""
    B, max_seq, n_new = 3, 512, 30
    pe, de = _pair(pkg, B, max_seq, pages=3 * 4)
    rng = np.random.default_rng(4)
    prompt = rng.integers(0, SMALL["vocab_size"], (B, 120)).astype(np.int32)
    free_run, _ = de.generate(prompt, n_new, eos_id=-1)
    eos = int(free_run[1][3])   # sequence 1 retires after its 4th token
    de.reset()                  # (a prompt pass continues the history the engine holds)
    want, want_n = de.generate(prompt, n_new, eos_id=eos)
    got, got_n = pe.generate(prompt, n_new, eos_id=eos)
    assert np.array_equal(np.asarray(got), np.asarray(want)) and np.array_equal(got_n, want_n)
    assert want_n[1] <= 4
    used = pe.kv_pages_total - pe.kv_pages_free
    others = sum((120 + int(n) + P - 1) // P for b, n in enumerate(got_n) if b != 1)
    assert used <= others, "sequence 1 retired early: its page must be back in the pool"
    pe.close(), de.close()


def test_cfg4_shape_paged_equals_dense(pkg):
    """configs[3] geometry (7B heads, 32 sequences x 4096 cached keys, 2 layers): the same synthetic rows in both caches,
    three decode steps (the second one opens page 33 of every sequence), logits bit-identical; the dense engine at this
    shape is pinned to the oracle by tests/test_7b_shapes_gpu.py"""
    E = pkg.engine
    cfg = dict(num_layers=2, head_num=32, kv_head_num=32, head_size=128, inter_size=11008, vocab_size=32000)
    B, ctx, max_seq = 32, 4095, 4104
    pe, de = _pair(pkg, B, max_seq, cfg=cfg, seed=77, std=0.02)
    pe.fill_kv_random(ctx, seed=5, amp=0.5)
    de.fill_kv_random(ctx, seed=6, amp=0.5)
    # copy the paged engine's rows into the dense cache (the two fills draw different rows)
    kp, vp = pe.kv_caches()
    kd, vd = de.kv_caches()
    torch.cuda.synchronize()
    kd.copy_(kp), vd.copy_(vp)
    torch.cuda.synchronize()   # torch's stream, not the engines': the copies must have landed before the first step
    del kp, vp
    ids = np.random.default_rng(3).integers(0, 32000, B).astype(np.int32)
    for s in range(3):
        a, b = pe.decode_step_host(ids if s == 0 else None), de.decode_step_host(ids if s == 0 else None)
        lp, ld = pe.logits().float(), de.logits().float()
        assert torch.equal(lp, ld), (f"cfg4 paged: logits step {s}: max |diff| {float((lp - ld).abs().max()):.3e} of max "
                                     f"{float(ld.abs().max()):.3e}, rows that differ {(lp != ld).any(dim=1).nonzero().flatten().tolist()}")
        assert np.array_equal(a, b)
    assert pe.kv_pages_total - pe.kv_pages_free == B * 33
    pe.close(), de.close()


@pytest.mark.parametrize("B,lens", [(2, [120, 120]), (5, [130, 5, 128, 64, 127])])
def test_fp8_pages_equal_the_dense_fp8_cache(pkg, B, lens):
    """B200_ENGINE_KV_FP8 | B200_ENGINE_KV_PAGED: e4m3 pages + scale pools. Same rounding, same kernels, same order as the
    dense fp8 cache (which tests/test_kv_fp8_gpu.py pins to the oracle on equally rounded rows): prompt pass + decode steps
    across a page boundary must give bit-identical logits, ids and (dequantised) cached rows; the pool is half the bytes."""
    E = pkg.engine
    max_seq, steps = 300, 10
    pe = E.Engine(B, max_seq, dtype=pkg.F16, kv_paged=True, kv_fp8=True, **SMALL)
    de = E.Engine(B, max_seq, dtype=pkg.F16, kv_fp8=True, **SMALL)
    fp16_paged = E.Engine(B, max_seq, dtype=pkg.F16, kv_paged=True, **SMALL)
    assert pe.kv_bytes < 0.53 * fp16_paged.kv_bytes
    fp16_paged.close()
    for eng in (pe, de):
        eng.init_random(seed=21, std=0.05)
    rng = np.random.default_rng(sum(lens))
    prompts = [rng.integers(0, SMALL["vocab_size"], n).astype(np.int32) for n in lens]
    if len(set(lens)) == 1:
        ip, idd = pe.prefill(np.stack(prompts)), de.prefill(np.stack(prompts))
    else:
        ip, idd = pe.prefill_ragged(prompts), de.prefill_ragged(prompts)
    assert np.array_equal(ip, idd)
    assert torch.equal(pe.logits(), de.logits()), "prefill logits"
    for s in range(steps):
        ip, idd = pe.decode_step_host(None), de.decode_step_host(None)
        assert torch.equal(pe.logits(), de.logits()), f"decode step {s}"
        assert np.array_equal(ip, idd)
    kp, vp = pe.kv_caches()
    kd, vd = de.kv_caches()
    for b, n in enumerate(lens):
        assert torch.equal(kp[:, b, :, :n + steps], kd[:, b, :, :n + steps]) and torch.equal(vp[:, b, :, :n + steps], vd[:, b, :, :n + steps])
    assert pe.kv_pages_total - pe.kv_pages_free == sum((n + steps + P - 1) // P for n in lens)
    pe.close(), de.close()
