#!/usr/bin/env python
"""A few decode steps of an N-layer LLaMA-2-7B-shaped engine at a given batch / context between
cudaProfilerStart/Stop, for `ncu --profile-from-start off` launch lists of the batched decode path."""
import argparse
import os
import sys

import numpy as np
import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from __graft_entry__ import load_package  # noqa: E402

ap = argparse.ArgumentParser()
ap.add_argument("--layers", type=int, default=2)
ap.add_argument("--batch", type=int, default=32)
ap.add_argument("--context", type=int, default=4096)
ap.add_argument("--steps", type=int, default=2)
ap.add_argument("--unfused", action="store_true", help="the reference's launch order instead of the fused GEMV order")
a = ap.parse_args()
pkg = load_package()
E = pkg.engine
cfg = dict(num_layers=a.layers, head_num=32, kv_head_num=32, head_size=128, inter_size=11008, vocab_size=32000)
eng = E.Engine(a.batch, a.context + 16, dtype=pkg.F16, **cfg)
eng.init_random(seed=1, std=0.02)
eng.fill_kv_random(a.context)
eng.set_mode(fused=not a.unfused, graph=False)
ids = np.zeros(a.batch, np.int32)
eng.decode_step_host(ids)
torch.cuda.synchronize()
torch.cuda.profiler.start()
for _ in range(a.steps):
    eng.decode_step()
torch.cuda.synchronize()
torch.cuda.profiler.stop()
print("launches per step", eng.launches_per_step)
