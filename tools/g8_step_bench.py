#!/usr/bin/env python
"""One GPU, 70B-shaped decoder slice (64 heads / 8 kv heads = the G = 8 attention of the tensor-parallel ranks, H 8192,
inter 28672, `layers` layers), batch 1: ms per step of the persistent kernel and of the graphed fused-GEMV order."""
import os
import sys
import time

import numpy as np
import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from __graft_entry__ import load_package  # noqa: E402

pkg = load_package()
E = pkg.engine
layers = int(sys.argv[1]) if len(sys.argv) > 1 else 8
ctx = int(sys.argv[2]) if len(sys.argv) > 2 else 512
cfg = dict(num_layers=layers, head_num=64, kv_head_num=8, head_size=128, inter_size=28672, vocab_size=32000)
for mode in ("persistent", "fused_graph"):
    eng = E.Engine(1, ctx + 200, dtype=pkg.F16, **cfg)
    eng.init_random(seed=1, std=0.02)
    eng.fill_kv_random(ctx, seed=3, amp=0.5)
    eng.set_mode(fused=True, graph=True, persistent=mode == "persistent")
    ids = np.array([5], np.int32)
    eng.decode_step_host(ids)
    for _ in range(5):
        eng.decode_step()
    torch.cuda.synchronize()
    stream = torch.cuda.ExternalStream(eng.stream)
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    with torch.cuda.stream(stream):
        e0.record()
        for _ in range(50):
            eng.decode_step()
        e1.record()
    torch.cuda.synchronize()
    ms = e0.elapsed_time(e1) / 50
    gb = (eng.weight_bytes - 2 * 32000 * 8192) / 1e9
    print(f"{mode:12s} {ms:7.3f} ms/step  launches {eng.launches_per_step:4d}  {gb / ms * 1e3:7.1f} GB/s of weights", flush=True)
    eng.close()
