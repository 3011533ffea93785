#!/usr/bin/env python
"""us per allreduce(+residual add) of a [B, 8192] fp16 partial: the one-shot NVLink kernel
(csrc/p2p_allreduce.cu) vs ncclAllReduce + b200_add_residual, each replayed from a CUDA graph of 200
calls. Run under torchrun, one rank per GPU:
  python -m torch.distributed.run --nproc-per-node 2 --master-addr 127.0.0.1 tools/allreduce_bench.py"""
import json
import os
import sys

import torch
import torch.distributed as dist

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from __graft_entry__ import load_package  # noqa: E402

rank, world, lr = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"]), int(os.environ["LOCAL_RANK"])
torch.cuda.set_device(lr)
dist.init_process_group("nccl", device_id=torch.device("cuda", lr))
pkg = load_package()
ops, sh = pkg.ops, pkg.sharding
comm = sh.P2PComm(32 * 8192 * 2)
res = {}
for B in (1, 32):
    n = B * 8192
    part = torch.full((B, 8192), 0.001 * (rank + 1), dtype=torch.float16, device="cuda")
    hidden = torch.zeros(B, 8192, dtype=torch.float16, device="cuda")
    side = torch.cuda.Stream()

    def run(fn, calls=200, reps=5):
        with torch.cuda.stream(side):
            fn()
            torch.cuda.synchronize()
            dist.barrier()
            g = torch.cuda.CUDAGraph()
            with torch.cuda.graph(g, stream=side):
                for _ in range(calls):
                    fn()
            g.replay()
            torch.cuda.synchronize()
            dist.barrier()
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record()
            for _ in range(reps):
                g.replay()
            e1.record()
            torch.cuda.synchronize()
        t = torch.tensor([e0.elapsed_time(e1) * 1e3 / (calls * reps)], device="cuda")
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        return round(float(t[0]), 2)

    def f_p2p():
        comm.allreduce_add(part, hidden)

    def f_nccl():
        dist.all_reduce(part)
        ops.launchAddResidual(part, hidden)

    res[f"B{B}"] = {"bytes": n * 2, "p2p_allreduce_add_us": run(f_p2p), "nccl_allreduce_plus_add_us": run(f_nccl)}
if rank == 0:
    print(json.dumps({"world": world, "what": "us per call inside a 200-call CUDA graph, max over ranks", **res}))
dist.barrier()
comm.close()
dist.destroy_process_group()
