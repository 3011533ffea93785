"""Tensor-parallel decode through the engine's allreduce hook (include/b200_engine.h:
b200_engine_set_allreduce), two ranks, against the UNSHARDED CPU oracle.
Each rank owns head_num/2 query heads, kv_head_num/2 kv heads and inter_size/2 FFN columns
(sharding.tp_shard_layer); the hook sums the partial [B,H] after o_proj and after down_proj.
With >= 2 GPUs the ranks sit on different devices and the sum is an NCCL allreduce enqueued on the
engine's stream, captured in the step's CUDA graph; on a one-GPU box both ranks share cuda:0 and
the sum is staged through host memory over gloo (graph capture off). fp32 rel <= 1e-5, fp16 <= 2e-2."""
import os
import socket
import sys

import numpy as np
import pytest
import torch

from conftest import ROOT, assert_close, load_package
from oracle import model as omodel
from test_sharding_gloo import make_weights

pytestmark = pytest.mark.gpu
CFG = dict(num_layers=2, head_num=8, kv_head_num=4, head_size=128, inter_size=1536, vocab_size=1031)
B, MAX_SEQ, CTX, STEPS = 2, 48, 9, 4


def free_port():
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        return s.getsockname()[1]


def _worker(rank, world, port, kind, fused, out_dir, comm="hook"):
    sys.path.insert(0, ROOT)
    sys.path.insert(0, os.path.join(ROOT, "tests"))
    import torch.distributed as dist
    pkg = load_package()
    E, sh = pkg.engine, pkg.sharding
    multi = torch.cuda.device_count() >= world
    dev = rank if multi else 0
    torch.cuda.set_device(dev)
    backend = "nccl" if multi else "gloo"
    kw = dict(device_id=torch.device("cuda", dev)) if multi else {}
    dist.init_process_group(backend, init_method=f"tcp://127.0.0.1:{port}", rank=rank, world_size=world, **kw)
    try:
        full = make_weights(CFG, kind, 33)
        eng = E.Engine(B, MAX_SEQ, dtype=pkg.F16 if kind == "fp16" else pkg.F32, tp_size=world, tp_rank=rank, **CFG)
        for l, w in enumerate(full["layers"]):
            s = sh.tp_shard_layer(w, CFG, world, rank)
            for slot, key in ((E.W_ATTN_NORM, "attn_norm"), (E.W_QKV, "qkv"), (E.W_O, "o"),
                              (E.W_FFN_NORM, "ffn_norm"), (E.W_GATE_UP, "gate_up"), (E.W_DOWN, "down")):
                eng.set_weight(slot, l, s[key])
        eng.set_weight(E.W_EMBED, 0, full["embed"])
        eng.set_weight(E.W_FINAL_NORM, 0, full["final_norm"])
        eng.set_weight(E.W_LM_HEAD, 0, full["lm_head"])
        rng = np.random.default_rng(5)
        rt = omodel.rounder(kind)
        kfull = rt(rng.uniform(-1, 1, (CFG["num_layers"], B, CFG["kv_head_num"], MAX_SEQ, CFG["head_size"])))
        vfull = rt(rng.uniform(-1, 1, kfull.shape))
        hkl = CFG["kv_head_num"] // world
        eng.fill_kv_random(CTX)      # step <- CTX; the rows are overwritten with this rank's shard below
        kc, vc = eng.kv_caches()
        kc.copy_(torch.from_numpy(kfull[:, :, rank * hkl:(rank + 1) * hkl]).to(kc.dtype))
        vc.copy_(torch.from_numpy(vfull[:, :, rank * hkl:(rank + 1) * hkl]).to(vc.dtype))
        torch.cuda.synchronize()
        assert eng.step == CTX
        eng.set_mode(fused=fused, graph=multi)
        n_calls = [0]
        inner = sh.make_allreduce()

        def hook(ptr, count, dtype, stream):
            n_calls[0] += 1
            assert count == B * CFG["head_num"] * CFG["head_size"]
            return inner(ptr, count, dtype, stream)

        p2p = None
        if comm in ("p2p", "p2p_persistent"):   # the fused one-shot NVLink allreduce + residual add (csrc/p2p_allreduce.cu)
            p2p = sh.P2PComm(B * CFG["head_num"] * CFG["head_size"] * 4)
            eng.set_p2p(p2p)
            if comm == "p2p_persistent":   # ONE launch per token: the exchange ops of decode_step_kernel do the allreduce
                eng.set_mode(fused=True, graph=True, persistent=True)
        else:
            eng.set_allreduce(hook)
        ids = np.array([7, 1000], np.int32)
        logits, picked = [], []
        for _ in range(STEPS):
            ids = eng.decode_step_host(ids)
            logits.append(eng.logits().float().cpu().numpy().copy())
            picked.append(ids.copy())
        if not multi and comm == "hook":
            assert n_calls[0] == STEPS * CFG["num_layers"] * 2, "two allreduces per layer per step"
        if comm == "p2p_persistent":
            assert eng.launches_per_step == 1, "tensor-parallel persistent step kernel not selected"
        kc, vc = eng.kv_caches()
        np.savez(os.path.join(out_dir, f"rank{rank}.npz"), logits=np.stack(logits), ids=np.stack(picked),
                 k=kc.float().cpu().numpy(), v=vc.float().cpu().numpy())
        eng.close()
        if p2p is not None:
            p2p.close()
    finally:
        dist.destroy_process_group()


def _p2p_unit_worker(rank, world, port, out_dir):
    """b200_p2p_allreduce_add / b200_p2p_allreduce against a torch sum of the same partials, many
    back-to-back calls (buffer parity, epochs), eagerly and replayed from a CUDA graph"""
    sys.path.insert(0, ROOT)
    sys.path.insert(0, os.path.join(ROOT, "tests"))
    import torch.distributed as dist
    pkg = load_package()
    torch.cuda.set_device(rank)
    dist.init_process_group("nccl", init_method=f"tcp://127.0.0.1:{port}", rank=rank, world_size=world,
                            device_id=torch.device("cuda", rank))
    try:
        comm = pkg.sharding.P2PComm(32 * 8192 * 2)
        ok = True
        for dt, n in ((torch.float16, 8192), (torch.float16, 32 * 8192), (torch.float32, 4096), (torch.float16, 8)):
            for it in range(6):
                g = [torch.Generator(device="cuda").manual_seed(100 * it + r) for r in range(world)]
                parts = [(torch.rand(n, device="cuda", generator=g[r]) - 0.5).to(dt) for r in range(world)]
                hidden = torch.full((n,), 0.25, dtype=dt, device="cuda")
                want = hidden.float()
                acc = torch.zeros(n, dtype=torch.float32, device="cuda")
                for r in range(world):
                    acc += parts[r].float()
                want_sum = acc.to(dt)
                want = (want + want_sum.float()).to(dt)
                comm.allreduce_add(parts[rank], hidden)
                buf = parts[rank].clone()
                comm.allreduce(buf)
                torch.cuda.synchronize()
                ok &= bool(torch.equal(hidden, want)) and bool(torch.equal(buf, want_sum))
        # CUDA-graph replay: the call counters live on the device
        n = 8192
        part = torch.full((n,), float(rank + 1), dtype=torch.float16, device="cuda")
        hidden = torch.zeros(n, dtype=torch.float16, device="cuda")
        side = torch.cuda.Stream()
        with torch.cuda.stream(side):
            comm.allreduce_add(part, hidden)
            torch.cuda.synchronize()
            graph = torch.cuda.CUDAGraph()
            with torch.cuda.graph(graph, stream=side):
                comm.allreduce_add(part, hidden)
                comm.allreduce_add(part, hidden)
            for _ in range(5):
                graph.replay()
        torch.cuda.synchronize()
        total = sum(range(1, world + 1))
        ok &= bool(torch.all(hidden == 11 * total))
        open(os.path.join(out_dir, f"p2p{rank}.txt"), "w").write("ok" if ok else "mismatch")
        dist.barrier()
        comm.close()
    finally:
        dist.destroy_process_group()


def test_p2p_oneshot_allreduce_two_gpus(tmp_path):
    if torch.cuda.device_count() < 2:
        pytest.skip("needs two GPUs with peer access (verified with gpurun --gpus 2, profiles/r01_tp2_*.json)")
    import torch.multiprocessing as mp
    world = 2
    mp.spawn(_p2p_unit_worker, args=(world, free_port(), str(tmp_path)), nprocs=world, join=True)
    for r in range(world):
        assert open(os.path.join(tmp_path, f"p2p{r}.txt")).read() == "ok"


@pytest.mark.parametrize("comm", ["hook", "p2p", "p2p_persistent"])
@pytest.mark.parametrize("fused", [True, False])
@pytest.mark.parametrize("kind", ["fp32", "fp16"])
def test_engine_tensor_parallel_two_ranks_vs_unsharded_oracle(tmp_path, kind, fused, comm):
    import torch.multiprocessing as mp
    world = 2
    if comm in ("p2p", "p2p_persistent") and torch.cuda.device_count() < world:
        pytest.skip("the one-shot NVLink allreduce needs one GPU per rank")
    if comm == "p2p_persistent" and (kind != "fp16" or not fused):
        pytest.skip("the persistent step kernel is fp16 and has one launch order")
    mp.spawn(_worker, args=(world, free_port(), kind, fused, str(tmp_path), comm), nprocs=world, join=True)
    full = make_weights(CFG, kind, 33)
    om = omodel.OracleModel(full, CFG, B, MAX_SEQ, kind)
    rng = np.random.default_rng(5)
    rt = omodel.rounder(kind)
    om.k[:] = rt(rng.uniform(-1, 1, om.k.shape))
    om.v[:] = rt(rng.uniform(-1, 1, om.v.shape))
    om.step = CTX
    got = [np.load(os.path.join(tmp_path, f"rank{r}.npz")) for r in range(world)]
    assert np.array_equal(got[0]["ids"], got[1]["ids"]), "ranks must pick the same tokens"
    assert np.array_equal(got[0]["logits"], got[1]["logits"]), "ranks must hold identical logits"
    ids = np.array([7, 1000], np.int32)
    for s in range(STEPS):
        want_logits, want_ids = om.decode_step(ids)
        assert_close(got[0]["logits"][s], want_logits, kind, f"logits step {s}")
        ids = got[0]["ids"][s]      # follow the engine's tokens so both sides see the same inputs
        if kind == "fp32":
            assert np.array_equal(ids, want_ids)
    hkl = CFG["kv_head_num"] // world
    for r in range(world):
        assert_close(got[r]["k"], om.k[:, :, r * hkl:(r + 1) * hkl], kind, f"rank {r} K shard")
        assert_close(got[r]["v"], om.v[:, :, r * hkl:(r + 1) * hkl], kind, f"rank {r} V shard")


def test_p2p_single_rank_gemv_push_and_reduce(pkg):
    """One-GPU coverage of the fused GEMV-push + reduce path (world = 1: the only slot is this rank's own, the
    wire format, tags, buffer parity and call counter are exercised exactly as with peers): b200_gemv_fused with
    args.p2p followed by b200_p2p_reduce_add must equal the GEMV with the residual epilogue, bit for bit, over
    several calls; b200_p2p_allreduce_add / b200_p2p_allreduce with one rank are the identity sum."""
    import ctypes as C
    E, ops, capi = pkg.engine, pkg.ops, pkg.capi
    lib = E._lib()
    B, N, K = 2, 4096, 1024
    comm = lib.b200_p2p_create(1, 0, B * N * 2)
    assert comm
    h = C.create_string_buffer(int(lib.b200_p2p_handle_bytes()))
    assert lib.b200_p2p_local_handle(comm, h) == 0
    assert lib.b200_p2p_connect(comm, h.raw) == 0
    g = torch.Generator(device="cuda").manual_seed(1)
    W = ((torch.rand(N, K, device="cuda", generator=g) - 0.5) * 0.1).half()
    for it in range(5):
        x = (torch.rand(B, K, device="cuda", generator=g) - 0.5).half()
        hidden = (torch.rand(B, N, device="cuda", generator=g) - 0.5).half()
        want = hidden.clone()
        ops.gemv_fused(x, W, want, capi.GEMV_RESIDUAL)
        got = hidden.clone()
        scratch = torch.zeros(B, N, dtype=torch.float16, device="cuda")
        ops.gemv_fused(x, W, scratch, 0, p2p=comm)
        st = C.c_void_p(torch.cuda.current_stream().cuda_stream)
        assert lib.b200_p2p_reduce_add(comm, C.c_void_p(got.data_ptr()), B * N, capi.F16, st) == 0
        torch.cuda.synchronize()
        assert torch.equal(got, want), f"call {it}"
        assert float(scratch.abs().max()) == 0.0, "with p2p the GEMV must not write y"
        part = (torch.rand(B, N, device="cuda", generator=g) - 0.5).half()
        acc = hidden.clone()
        assert lib.b200_p2p_allreduce_add(comm, C.c_void_p(part.data_ptr()), C.c_void_p(acc.data_ptr()), B * N, capi.F16, st) == 0
        buf = part.clone()
        assert lib.b200_p2p_allreduce(comm, C.c_void_p(buf.data_ptr()), B * N, capi.F16, st) == 0
        torch.cuda.synchronize()
        assert torch.equal(acc, (hidden.float() + part.float()).half()) and torch.equal(buf, part)
    lib.b200_p2p_destroy(comm)


RAGGED_LENS, RAGGED_STEPS = [11, 4], 3


def _ragged_worker(rank, world, port, kind, out_dir, comm):
    """tensor-parallel engine, prompts of different lengths: b200_engine_prefill_ragged (row-split o / down summed
    over the ranks for sum(lens) tokens), then free-running decode steps with per-sequence positions"""
    sys.path.insert(0, ROOT)
    sys.path.insert(0, os.path.join(ROOT, "tests"))
    import torch.distributed as dist
    pkg = load_package()
    E, sh = pkg.engine, pkg.sharding
    multi = torch.cuda.device_count() >= world
    dev = rank if multi else 0
    torch.cuda.set_device(dev)
    kw = dict(device_id=torch.device("cuda", dev)) if multi else {}
    dist.init_process_group("nccl" if multi else "gloo", init_method=f"tcp://127.0.0.1:{port}", rank=rank,
                            world_size=world, **kw)
    try:
        full = make_weights(CFG, kind, 33)
        eng = E.Engine(B, MAX_SEQ, dtype=pkg.F16 if kind == "fp16" else pkg.F32, tp_size=world, tp_rank=rank, **CFG)
        for l, w in enumerate(full["layers"]):
            s = sh.tp_shard_layer(w, CFG, world, rank)
            for slot, key in ((E.W_ATTN_NORM, "attn_norm"), (E.W_QKV, "qkv"), (E.W_O, "o"),
                              (E.W_FFN_NORM, "ffn_norm"), (E.W_GATE_UP, "gate_up"), (E.W_DOWN, "down")):
                eng.set_weight(slot, l, s[key])
        eng.set_weight(E.W_EMBED, 0, full["embed"])
        eng.set_weight(E.W_FINAL_NORM, 0, full["final_norm"])
        eng.set_weight(E.W_LM_HEAD, 0, full["lm_head"])
        eng.set_mode(fused=True, graph=multi)
        p2p = None
        if comm == "p2p":
            p2p = sh.P2PComm(64 * CFG["head_num"] * CFG["head_size"] * 4)   # the prompt pass sums sum(lens) rows at once
            eng.set_p2p(p2p)
        else:
            eng.set_allreduce(sh.make_allreduce())
        rng = np.random.default_rng(77)
        prompts = [rng.integers(0, CFG["vocab_size"], n).astype(np.int32) for n in RAGGED_LENS]
        ids = eng.prefill_ragged(prompts)
        logits, picked = [eng.logits().float().cpu().numpy().copy()], [ids.copy()]
        for _ in range(RAGGED_STEPS):
            ids = eng.decode_step_host(None)
            logits.append(eng.logits().float().cpu().numpy().copy())
            picked.append(ids.copy())
        assert list(eng.seq_steps) == [n + RAGGED_STEPS for n in RAGGED_LENS]
        kc, _ = eng.kv_caches()
        np.savez(os.path.join(out_dir, f"ragged{rank}.npz"), logits=np.stack(logits), ids=np.stack(picked),
                 k=kc.float().cpu().numpy(), prompts=np.concatenate(prompts))
        eng.close()
        if p2p is not None:
            p2p.close()
    finally:
        dist.destroy_process_group()


@pytest.mark.parametrize("kind,comm", [("fp16", "p2p"), ("fp16", "hook"), ("fp32", "hook")])
def test_engine_tensor_parallel_ragged_batch(tmp_path, kind, comm):
    """each sequence of a ragged tensor-parallel batch against the UNSHARDED oracle run on that sequence alone"""
    import torch.multiprocessing as mp
    from test_prefill_gpu import oracle_prefill
    world = 2
    if comm == "p2p" and torch.cuda.device_count() < world:
        pytest.skip("the one-shot NVLink allreduce needs one GPU per rank")
    mp.spawn(_ragged_worker, args=(world, free_port(), kind, str(tmp_path), comm), nprocs=world, join=True)
    got = [np.load(os.path.join(tmp_path, f"ragged{r}.npz")) for r in range(world)]
    assert np.array_equal(got[0]["ids"], got[1]["ids"]) and np.array_equal(got[0]["logits"], got[1]["logits"])
    full = make_weights(CFG, kind, 33)
    hkl = CFG["kv_head_num"] // world
    off = 0
    for b, n in enumerate(RAGGED_LENS):
        om = omodel.OracleModel(full, CFG, 1, MAX_SEQ, kind)
        prompt = got[0]["prompts"][off:off + n]
        off += n
        want, _ = oracle_prefill(om, full, CFG, prompt[None, :], 0)
        assert_close(got[0]["logits"][0][b], want[0], kind, f"prefill logits of sequence {b}")
        for s in range(RAGGED_STEPS):
            want, _ = om.decode_step(got[0]["ids"][s][b:b + 1])   # follow the engine's tokens
            assert_close(got[0]["logits"][s + 1][b], want[0], kind, f"sequence {b} decode step {s}")
        for r in range(world):
            assert_close(got[r]["k"][:, b], om.k[:, 0, r * hkl:(r + 1) * hkl], kind, f"rank {r} K shard of sequence {b}")
