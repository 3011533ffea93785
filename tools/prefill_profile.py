#!/usr/bin/env python
"""One prompt pass (cfg3 geometry: 4 x 2048 tokens) of an N-layer LLaMA-2-7B-shaped engine between
cudaProfilerStart/Stop, for `ncu --profile-from-start off ...` (launch list / --set full captures
of gemm_tcgen05_kernel and context_attn_tcgen05_kernel). Prints the device time of the pass."""
import argparse
import os
import sys

import numpy as np
import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from __graft_entry__ import load_package  # noqa: E402

ap = argparse.ArgumentParser()
ap.add_argument("--layers", type=int, default=2)
ap.add_argument("--batch", type=int, default=4)
ap.add_argument("--seq", type=int, default=2048)
a = ap.parse_args()
pkg = load_package()
E = pkg.engine
cfg = dict(num_layers=a.layers, head_num=32, kv_head_num=32, head_size=128, inter_size=11008, vocab_size=32000)
eng = E.Engine(a.batch, a.seq + 8, dtype=pkg.F16, **cfg)
eng.init_random(seed=1, std=0.02)
ids = np.random.default_rng(0).integers(0, 32000, (a.batch, a.seq)).astype(np.int32)
eng.prefill(ids)
eng.reset()
torch.cuda.synchronize()
stream = torch.cuda.ExternalStream(eng.stream)
e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
torch.cuda.profiler.start()
with torch.cuda.stream(stream):
    e0.record()
    eng.prefill(ids, sync=False)
    e1.record()
torch.cuda.synchronize()
torch.cuda.profiler.stop()
print(f"prefill {a.batch}x{a.seq} tokens, {a.layers} layers: {e0.elapsed_time(e1):.3f} ms, {eng.prefill_launches} launches")
