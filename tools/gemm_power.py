#!/usr/bin/env python
"""usage: gemm_power.py [N] [VARIANT ...], VARIANT = comma-separated B200_GEMM2_* knobs without the prefix, e.g. WIDE=1 or QUAD=1,GROUP_M=8
SM clock and board power while one GEMM runs back to back for a few seconds: gemm2_tcgen05_kernel vs the cuBLAS kernel
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


variants = [dict(kv.split("=") for kv in a.split(",") if kv) for a in sys.argv[2:]] or [{}, {"WIDE": "1"}, {"QUAD": "1"}]
variants = [{("B200_GEMM2_" + k_): v for k_, v in d.items()} for d in variants]
# the board drifts (temperature, power-cap controller): every variant is timed in ROUNDS interleaved rounds, means are printed last
ROUNDS, SECS = 3, 2.0
names = ["cuBLAS"] + [(" ".join(f"{k_[11:]}={v}" for k_, v in env.items()) or "default") for env in variants]
acc = {nm: [] for nm in names}
_run = run


def run(name, fn, secs=3.0):   # noqa: F811 - keeps the one-line report, collects TFLOP/s and clock
    import io, contextlib
    buf = io.StringIO()
    with contextlib.redirect_stdout(buf):
        _run(name, fn, secs)
    line = buf.getvalue().strip()
    print(line, flush=True)
    parts = line.split()
    return float(parts[parts.index("TFLOP/s") - 1]), float(parts[parts.index("MHz") - 1])


for r in range(ROUNDS):
    acc["cuBLAS"].append(run("cuBLAS (torch linear)", lambda i: torch.nn.functional.linear(x, ws[i % 4]), SECS))
    for env, nm in zip(variants, names[1:]):
        for k_ in ("B200_GEMM2_WIDE", "B200_GEMM2_QUAD", "B200_GEMM2_GROUP_M"):
            os.environ.pop(k_, None)
        os.environ.update(env)
        acc[nm].append(run("gemm2 " + nm, lambda i: ops.launchLinearGemm(x, ws[i % 4], y, trans_b=True), SECS))
print("---- means over", ROUNDS, "interleaved rounds")
base = sum(t for t, _ in acc["cuBLAS"]) / ROUNDS
for nm in names:
    t = sum(v[0] for v in acc[nm]) / ROUNDS
    c = sum(v[1] for v in acc[nm]) / ROUNDS
    print(f"{nm:32s} {t:7.1f} TFLOP/s  {c:6.0f} MHz  {t / base:5.3f} x cuBLAS", flush=True)
