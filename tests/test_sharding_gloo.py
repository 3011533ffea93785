"""N>1 host logic on CPU, two `gloo` ranks:
  * request sharding (independent sequences, no collective) and the max-over-ranks step time the
    bench reports;
  * tensor parallel: every rank runs the CPU oracle on ITS shard of the weights
    (sharding.tp_shard_layer), partial [B,H] sums are all-reduced after o_proj and down_proj only,
    and the logits / picked tokens / KV rows equal the unsharded oracle's (fp32 rel <= 1e-5 — the
    split changes the summation order, nothing else; fp16 <= 2e-2).
The GPU counterpart (the engine's allreduce hook) is tests/test_tp_gpu.py."""
import os
import socket
import sys

import numpy as np
import pytest

from conftest import ROOT, assert_close, load_package


def free_port():
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        return s.getsockname()[1]


def make_weights(cfg, kind, seed):
    from oracle import model as omodel
    rt = omodel.rounder(kind)
    rng = np.random.default_rng(seed)
    h, hk, d, I, V = cfg["head_num"], cfg["kv_head_num"], cfg["head_size"], cfg["inter_size"], cfg["vocab_size"]
    H = h * d
    layers = [dict(attn_norm=rt(1 + 0.05 * rng.standard_normal(H)),
                   qkv=rt(0.05 * rng.standard_normal(((h + 2 * hk) * d, H))),
                   o=rt(0.05 * rng.standard_normal((H, H))),
                   ffn_norm=rt(1 + 0.05 * rng.standard_normal(H)),
                   gate_up=rt(0.05 * rng.standard_normal((2 * I, H))),
                   down=rt(0.05 * rng.standard_normal((H, I)))) for _ in range(cfg["num_layers"])]
    return dict(layers=layers, embed=rt(rng.standard_normal((V, H))),
                final_norm=rt(1 + 0.05 * rng.standard_normal(H)), lm_head=rt(0.05 * rng.standard_normal((V, H))))


CFG = dict(num_layers=2, head_num=8, kv_head_num=4, head_size=32, inter_size=96, vocab_size=257)


def _tp_worker(rank, world, port, kind, out_dir):
    sys.path.insert(0, ROOT)
    sys.path.insert(0, os.path.join(ROOT, "tests"))
    import torch
    import torch.distributed as dist
    from oracle import model as omodel
    pkg = load_package()
    sh = pkg.sharding
    dist.init_process_group("gloo", init_method=f"tcp://127.0.0.1:{port}", rank=rank, world_size=world)
    try:
        B, max_seq, ctx, steps = 3, 24, 5, 3
        full = make_weights(CFG, kind, 21)
        local = dict(full, layers=[sh.tp_shard_layer(w, CFG, world, rank) for w in full["layers"]])
        lcfg = sh.tp_local_cfg(CFG, world)
        calls = []

        def allreduce(a):
            t = torch.from_numpy(a)
            dist.all_reduce(t)
            calls.append(a.shape)

        om = omodel.OracleModel(local, lcfg, B, max_seq, kind, allreduce=allreduce)
        rng = np.random.default_rng(5)
        kfull = omodel.rounder(kind)(rng.uniform(-1, 1, (CFG["num_layers"], B, CFG["kv_head_num"], max_seq, CFG["head_size"])))
        vfull = omodel.rounder(kind)(rng.uniform(-1, 1, kfull.shape))
        hkl = lcfg["kv_head_num"]
        om.k[:] = kfull[:, :, rank * hkl:(rank + 1) * hkl]
        om.v[:] = vfull[:, :, rank * hkl:(rank + 1) * hkl]
        om.step = ctx
        ids = np.array([3, 250, 17], np.int32)
        logits, picked = [], []
        for _ in range(steps):
            lg, ids = om.decode_step(ids)
            logits.append(lg)
            picked.append(ids.copy())
        assert len(calls) == steps * CFG["num_layers"] * 2, "two allreduces per layer per step, no more"
        assert all(s == (B, CFG["head_num"] * CFG["head_size"]) for s in calls)
        # every rank holds the same logits after the last allreduce
        t = torch.from_numpy(np.stack(logits).copy())
        gathered = [torch.empty_like(t) for _ in range(world)]
        dist.all_gather(gathered, t)
        assert all(torch.equal(g, gathered[0]) for g in gathered)
        # the job's step time is the slowest rank's
        assert sh.max_over_ranks(1.0 + rank) == float(world)
        np.savez(os.path.join(out_dir, f"rank{rank}.npz"), logits=np.stack(logits), ids=np.stack(picked), k=om.k, v=om.v)
    finally:
        dist.destroy_process_group()


@pytest.mark.parametrize("kind", ["fp32", "fp16"])
def test_tensor_parallel_oracle_two_gloo_ranks_equals_unsharded(tmp_path, kind):
    import torch.multiprocessing as mp
    from oracle import model as omodel
    world = 2
    mp.spawn(_tp_worker, args=(world, free_port(), kind, str(tmp_path)), nprocs=world, join=True)
    B, max_seq, ctx, steps = 3, 24, 5, 3
    full = make_weights(CFG, kind, 21)
    om = omodel.OracleModel(full, CFG, B, max_seq, kind)
    rng = np.random.default_rng(5)
    om.k[:] = omodel.rounder(kind)(rng.uniform(-1, 1, om.k.shape))
    om.v[:] = omodel.rounder(kind)(rng.uniform(-1, 1, om.v.shape))
    om.step = ctx
    ids = np.array([3, 250, 17], np.int32)
    want_logits, want_ids = [], []
    for _ in range(steps):
        lg, ids = om.decode_step(ids)
        want_logits.append(lg)
        want_ids.append(ids.copy())
    hkl = CFG["kv_head_num"] // world
    for r in range(world):
        got = np.load(os.path.join(tmp_path, f"rank{r}.npz"))
        assert_close(got["logits"], np.stack(want_logits), kind, f"rank {r} logits")
        if kind == "fp32":
            assert np.array_equal(got["ids"], np.stack(want_ids)), "greedy tokens"
        assert_close(got["k"], om.k[:, :, r * hkl:(r + 1) * hkl], kind, f"rank {r} K shard")
        assert_close(got["v"], om.v[:, :, r * hkl:(r + 1) * hkl], kind, f"rank {r} V shard")


def test_tp_shards_tile_the_layer_exactly():
    sh = load_package().sharding
    full = make_weights(CFG, "fp32", 3)["layers"][0]
    h, hk, d, I = CFG["head_num"], CFG["kv_head_num"], CFG["head_size"], CFG["inter_size"]
    H = h * d
    for tp in (1, 2, 4):
        parts = [sh.tp_shard_layer(full, CFG, tp, r) for r in range(tp)]
        hl, hkl, Il = h // tp, hk // tp, I // tp
        q = np.concatenate([p["qkv"][:hl * d] for p in parts])
        k = np.concatenate([p["qkv"][hl * d:(hl + hkl) * d] for p in parts])
        v = np.concatenate([p["qkv"][(hl + hkl) * d:] for p in parts])
        assert np.array_equal(np.concatenate([q, k, v]), full["qkv"])
        assert np.array_equal(np.concatenate([p["o"] for p in parts], axis=1), full["o"])
        assert np.array_equal(np.concatenate([p["down"] for p in parts], axis=1), full["down"])
        gate = np.concatenate([p["gate_up"][:Il] for p in parts])
        up = np.concatenate([p["gate_up"][Il:] for p in parts])
        assert np.array_equal(np.concatenate([gate, up]), full["gate_up"])
        assert all(p["qkv"].shape == ((hl + 2 * hkl) * d, H) for p in parts)
    with pytest.raises(ValueError):
        sh.tp_shard_layer(full, CFG, 3, 0)


def test_request_sharding_rules():
    sh = load_package().sharding
    assert sh.shard_requests(1, 8, 3, "weak") == (1, 8)
    assert sh.shard_requests(32, 8, 7, "strong") == (4, 32)
    got = [sh.shard_requests(10, 4, r, "strong")[0] for r in range(4)]
    assert got == [3, 3, 2, 2] and sum(got) == 10
    spans = [sh.request_slice(10, 4, r) for r in range(4)]
    assert spans == [(0, 3), (3, 6), (6, 8), (8, 10)]
    with pytest.raises(ValueError):
        sh.shard_requests(2, 4, 0, "strong")
    with pytest.raises(ValueError):
        sh.shard_requests(2, 2, 2, "weak")


def _dp_worker(rank, world, port, out_dir):
    sys.path.insert(0, ROOT)
    sys.path.insert(0, os.path.join(ROOT, "tests"))
    import torch.distributed as dist
    from oracle import model as omodel
    sh = load_package().sharding
    dist.init_process_group("gloo", init_method=f"tcp://127.0.0.1:{port}", rank=rank, world_size=world)
    try:
        B_total, max_seq = 5, 16
        n_local, total = sh.shard_requests(B_total, world, rank, "strong")
        lo, hi = sh.request_slice(B_total, world, rank)
        assert total == B_total and hi - lo == n_local
        full = make_weights(CFG, "fp32", 8)
        om = omodel.OracleModel(full, CFG, n_local, max_seq, "fp32")
        ids = np.arange(10, 10 + B_total, dtype=np.int32)[lo:hi]
        outs = []
        for _ in range(3):
            _, ids = om.decode_step(ids)
            outs.append(ids.copy())
        np.save(os.path.join(out_dir, f"dp{rank}.npy"), np.stack(outs, axis=1))
        assert sh.max_over_ranks(10.0 - rank) == 10.0
    finally:
        dist.destroy_process_group()


def test_independent_requests_two_gloo_ranks_need_no_collective(tmp_path):
    """sequences partitioned over ranks decode to the same tokens as the one-rank batch"""
    import torch.multiprocessing as mp
    from oracle import model as omodel
    world = 2
    mp.spawn(_dp_worker, args=(world, free_port(), str(tmp_path)), nprocs=world, join=True)
    full = make_weights(CFG, "fp32", 8)
    om = omodel.OracleModel(full, CFG, 5, 16, "fp32")
    ids = np.arange(10, 15, dtype=np.int32)
    outs = []
    for _ in range(3):
        _, ids = om.decode_step(ids)
        outs.append(ids.copy())
    want = np.stack(outs, axis=1)
    got = np.concatenate([np.load(os.path.join(tmp_path, f"dp{r}.npy")) for r in range(world)])
    assert np.array_equal(got, want)
