"""Per-shape timing of the weight-streaming GEMV (development aid; run under gpurun).
Cycles through enough distinct weight buffers that nothing is served from L2."""
import json
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from __graft_entry__ import load_package

pkg = load_package()
ops, capi = pkg.ops, pkg.capi
peak = json.load(open("MEASURED_PEAKS.json"))["hbm_gbs"] if os.path.exists("MEASURED_PEAKS.json") else 6650.0


def run(name, N, K, flags, nb=1, dtype=torch.float16, copies=None):
    rows = 2 * N if flags & capi.GEMV_SWIGLU else N
    bytes_w = rows * K * (2 if dtype == torch.float16 else 4)
    copies = copies or max(4, int(600e6 // bytes_w))
    Ws = [torch.randn(rows, K, device="cuda", dtype=dtype) * 0.02 for _ in range(copies)]
    x = torch.randn(nb, K, device="cuda", dtype=dtype) * 0.1
    y = torch.zeros(nb, N, device="cuda", dtype=dtype)
    gam = torch.ones(K, device="cuda", dtype=dtype)
    kw = {}
    if flags & capi.GEMV_NORM:
        kw.update(gamma=gam, eps=1e-5)
    if flags & capi.GEMV_ROPE_KV:
        kc = torch.zeros(nb, 32, 64, 128, device="cuda", dtype=dtype)
        den = torch.zeros(64, device="cuda")
        capi.lib().b200_rope_table(den.data_ptr(), 128, 10000.0, None)
        kw.update(k_cache_layer=kc, v_cache_layer=kc.clone(), head_num=32, kv_head_num=32, head_size=128,
                  rotary_dim=128, step=5, rope_den=den)
    for W in Ws:
        ops.gemv_fused(x, W, y, flags, **kw)
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    reps = 3
    e0.record()
    for _ in range(reps):
        for W in Ws:
            ops.gemv_fused(x, W, y, flags, **kw)
    e1.record()
    torch.cuda.synchronize()
    us = e0.elapsed_time(e1) * 1e3 / (reps * copies)
    gbs = bytes_w / (us * 1e-6) / 1e9
    print(f"{name:28s} nb={nb} N={N:6d} K={K:6d} {bytes_w/1e6:7.1f} MB  {us:8.2f} us  {gbs:7.1f} GB/s  {gbs/peak*100:5.1f}% of {peak:.0f}")
    return us


if __name__ == "__main__":
    nbs = [int(a) for a in sys.argv[1:]] or [1]
    for nb in nbs:
        t = 0
        t += 32 * run("qkv  norm+rope+kv", 12288, 4096, capi.GEMV_NORM | capi.GEMV_ROPE_KV, nb)
        t += 32 * run("o    residual", 4096, 4096, capi.GEMV_RESIDUAL, nb)
        t += 32 * run("g/u  norm+swiglu", 11008, 4096, capi.GEMV_NORM | capi.GEMV_NORM_GFIRST | capi.GEMV_SWIGLU, nb)
        t += 32 * run("down residual", 4096, 11008, capi.GEMV_RESIDUAL, nb)
        t += run("lm_head norm", 32000, 4096, capi.GEMV_NORM, nb)
        print(f"nb={nb}: sum of GEMV time per token = {t/1e3:.3f} ms  -> {1e6/t:.1f} tok/s ceiling from GEMVs alone")
    run("plain fp32 qkv", 12288, 4096, 0, 1, torch.float32)
