#!/usr/bin/env python
"""Per-op timing (CTA 0 stamps, B200_STEP_TRACE=1) of the persistent step kernel in its tensor-parallel (exchange)
form on ONE GPU: a world = 1 communicator attached to an engine with 70B-like per-rank shapes.
usage: step_trace_tp.py [layers] [inter_size] [context]"""
import ctypes as C
import os
import sys

import numpy as np

os.environ["B200_STEP_TRACE"] = "1"
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from __graft_entry__ import load_package  # noqa: E402

pkg = load_package()
E, capi = pkg.engine, pkg.capi
layers = int(sys.argv[1]) if len(sys.argv) > 1 else 8
inter = int(sys.argv[2]) if len(sys.argv) > 2 else 14336
ctx = int(sys.argv[3]) if len(sys.argv) > 3 else 512
cfg = dict(num_layers=layers, head_num=64, kv_head_num=8, head_size=128, inter_size=inter, vocab_size=32000)
eng = E.Engine(1, ctx + 64, dtype=capi.F16, **cfg)
eng.init_random(seed=1, std=0.02)
lib = E._lib()
comm = lib.b200_p2p_create(1, 0, 8192 * 2)
hnd = C.create_string_buffer(int(lib.b200_p2p_handle_bytes()))
assert lib.b200_p2p_local_handle(comm, hnd) == 0 and lib.b200_p2p_connect(comm, hnd.raw) == 0
assert lib.b200_engine_set_p2p(eng.h, comm) == 0
eng.set_mode(fused=True, graph=True, persistent=True)
eng.fill_kv_random(ctx, seed=2, amp=0.5)
ids = np.zeros(1, np.int32)
for _ in range(5):
    ids = eng.decode_step_host(ids)
t = eng.step_trace().astype(np.int64)
per_layer = (len(t) - 2) // layers
names = {5: ["qkv", "attn_o", "exch_o", "ffn", "exch_dn"]}.get(per_layer, [f"op{i}" for i in range(per_layer)])
print("ops:", len(t), "per layer:", per_layer, "total us:", (t[-1, 0] - t[0, 0]) / 1e3)
for j, nm in enumerate(names):
    idx = np.arange(j, layers * per_layer, per_layer)
    idx = idx[idx > 0]
    op = (t[idx, 0] - t[idx - 1, 1]).mean() / 1e3
    bar = (t[idx, 1] - t[idx, 0]).mean() / 1e3
    print(f"{nm:8s} op {op:7.2f} us   barrier wait {bar:6.2f} us")
w = eng.weight_bytes / 1e9
print(f"weights {w:.2f} GB -> {w / 7.07:.3f} ms at 7.07 TB/s")
