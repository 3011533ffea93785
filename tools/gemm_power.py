#!/usr/bin/env python
"""SM clock and board power while one GEMM runs back to back for a few seconds: gemm2_tcgen05_kernel vs the cuBLAS kernel
on the same 7B prefill shape (both are power-capped on a B200: the clock they settle at is part of the comparison)."""
import os
import subprocess
import sys
import threading
import time

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from __graft_entry__ import load_package  # noqa: E402

ops = load_package().ops
m, n, k = 8192, int(sys.argv[1]) if len(sys.argv) > 1 else 12288, 4096
ws = [(torch.randn(n, k, device="cuda") * 0.02).half() for _ in range(4)]
x = (torch.rand(m, k, device="cuda") - 0.5).half()
y = torch.empty(m, n, dtype=torch.float16, device="cuda")


def sample(stop, out):
    while not stop.is_set():
        r = subprocess.run(["nvidia-smi", "-i", "0", "--query-gpu=clocks.sm,power.draw,temperature.gpu",
                            "--format=csv,noheader,nounits"], capture_output=True, text=True).stdout.strip()
        try:
            out.append([float(v) for v in r.split(",")])
        except ValueError:
            pass
        time.sleep(0.05)


def run(name, fn, secs=4.0):
    for i in range(8):
        fn(i)
    torch.cuda.synchronize()
    stop, out = threading.Event(), []
    th = threading.Thread(target=sample, args=(stop, out))
    th.start()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    t0, it = time.time(), 0
    e0.record()
    while time.time() - t0 < secs:
        for i in range(50):
            fn(i)
        it += 50
        torch.cuda.synchronize()
    e1.record()
    torch.cuda.synchronize()
    stop.set()
    th.join()
    ms = e0.elapsed_time(e1) / it
    out = out[len(out) // 3:]
    med = lambda j: sorted(o[j] for o in out)[len(out) // 2] if out else float("nan")
    print(f"{name:28s} {ms * 1e3:8.1f} us  {2 * m * n * k / ms / 1e9:7.1f} TFLOP/s  sm {med(0):6.0f} MHz  power {med(1):6.0f} W  temp {med(2):4.0f} C"
          f"  -> {2 * m * n * k / ms / 1e9 / med(0) * 1000:7.1f} TFLOP/s per GHz", flush=True)


def mine(group_m, quiet):
    def fn(i):
        ops.launchLinearGemm(x, ws[i % 4], y, trans_b=True)
    def go(secs=3.0):
        os.environ["B200_GEMM2_GROUP_M"] = str(group_m)
        os.environ["B200_GEMM2_QUIET"] = str(quiet)
        run(f"gemm2 group_m={group_m} quiet={quiet}", fn, secs)
    return go


ref = torch.nn.functional.linear(x, ws[0]).float()
for gm, q in ((16, 7), (16, 22), (8, 10)):
    os.environ["B200_GEMM2_GROUP_M"], os.environ["B200_GEMM2_QUIET"] = str(gm), str(q)
    ops.launchLinearGemm(x, ws[0], y, trans_b=True)
    torch.cuda.synchronize()
    print("check", gm, q, float((y.float() - ref).abs().max()), flush=True)
run("cuBLAS (torch linear)", lambda i: torch.nn.functional.linear(x, ws[i % 4]), 3.0)
for gm, q in ((16, 0), (16, 1), (16, 16), (16, 2), (16, 4), (16, 8 + 2), (16, 1 + 2 + 4), (16, 16 + 2 + 4), (8, 0), (32, 0)):
    mine(gm, q)()
run("cuBLAS (torch linear)", lambda i: torch.nn.functional.linear(x, ws[i % 4]), 3.0)
