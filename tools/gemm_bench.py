#!/usr/bin/env python
"""Time the tcgen05 GEMM (through launchLinearGemm / the C ABI) on the cfg3 prefill shapes and the
batched-decode shapes; CUDA events around a torch CUDA graph of the launches; prints TFLOP/s and
GB/s of algorithmic weight bytes. torch.matmul (cuBLAS) is timed beside it as a yardstick only."""
import json
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from __graft_entry__ import load_package  # noqa: E402

ops = load_package().ops


def timed(fn, reps):
    fn()
    torch.cuda.synchronize()
    s = torch.cuda.Stream()
    with torch.cuda.stream(s):
        g = torch.cuda.CUDAGraph()
        with torch.cuda.graph(g, stream=s):
            for _ in range(reps):
                fn()
        g.replay()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        g.replay()
        e1.record()
    torch.cuda.synchronize()
    return e0.elapsed_time(e1) / reps


def main():
    shapes = [(8192, 12288, 4096), (8192, 4096, 4096), (8192, 22016, 4096), (8192, 4096, 11008),
              (32, 12288, 4096), (32, 4096, 4096), (32, 22016, 4096), (32, 4096, 11008), (32, 32000, 4096),
              (16, 22016, 4096), (64, 22016, 4096), (128, 22016, 4096), (256, 22016, 4096), (512, 22016, 4096)]
    if len(sys.argv) > 1:   # e.g. 128,22016,4096 256,12288,4096
        shapes = [tuple(int(v) for v in a.split(',')) for a in sys.argv[1:]]
    out = []
    for m, n, k in shapes:
        nw = max(1, int(300e6 // (n * k * 2)) + 1)   # cycle through > 300 MB of weights so HBM, not L2, is measured
        ws = [(torch.randn(n, k, device="cuda") * 0.02).half() for _ in range(nw)]
        x = (torch.rand(m, k, device="cuda") - 0.5).half()
        y = torch.empty(m, n, dtype=torch.float16, device="cuda")
        it = [0]

        def ours():
            ops.launchLinearGemm(x, ws[it[0] % nw], y, trans_b=True)
            it[0] += 1

        def cublas():
            torch.matmul(x, ws[it[0] % nw].t(), out=y)
            it[0] += 1

        reps = 2 * nw if m <= 256 else 4
        t = timed(ours, reps)
        tc = timed(cublas, reps)
        rec = dict(m=m, n=n, k=k, ms=round(t, 4), tflops=round(2.0 * m * n * k / t / 1e9, 1),
                   weight_GBps=round(n * k * 2 / t / 1e6, 1), cublas_ms=round(tc, 4),
                   cublas_tflops=round(2.0 * m * n * k / tc / 1e9, 1))
        print(json.dumps(rec), flush=True)
        out.append(rec)
        del ws


if __name__ == "__main__":
    main()
