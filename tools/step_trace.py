#!/usr/bin/env python
"""Per-op timing of the persistent whole-step kernel (B200_STEP_TRACE=1): CTA 0's %globaltimer
stamps after each op and after each grid barrier, averaged per op kind over the 32 layers."""
import os
import sys

import numpy as np

os.environ["B200_STEP_TRACE"] = "1"
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from __graft_entry__ import load_package  # noqa: E402

pkg = load_package()
E, capi = pkg.engine, pkg.capi
L7B = dict(num_layers=32, head_num=32, kv_head_num=32, head_size=128, inter_size=11008, vocab_size=32000)
B = int(sys.argv[1]) if len(sys.argv) > 1 else 1
ctx = int(sys.argv[2]) if len(sys.argv) > 2 else 512
eng = E.Engine(B, ctx + 64, dtype=capi.F16, **L7B)
eng.init_random(seed=1, std=0.02)
eng.fill_kv_random(ctx, seed=2, amp=0.5)
ids = np.zeros(B, np.int32)
for _ in range(5):
    ids = eng.decode_step_host(ids)
t = eng.step_trace().astype(np.int64)
per_layer = (len(t) - 2) // 32
names = {5: ["qkv", "attn", "o", "gate_up", "down"], 4: ["qkv", "attn", "o", "ffn"], 3: ["qkv", "attn_o", "ffn"]}[per_layer]
n_layer_ops = 32 * per_layer
start = t[0, 0]
print("kernel start -> first op done: n/a; total ns:", t[-1, 0] - (t[0, 0]), "ops:", len(t))
prev = None
acc = {}
for i in range(len(t)):
    kind = names[i % per_layer] if i < n_layer_ops else ("lm_head" if i == n_layer_ops else "final")
    op_ns = t[i, 0] - (t[i - 1, 1] if i else t[i, 0])
    bar_ns = t[i, 1] - t[i, 0]
    x_ns = (t[i, 2] - t[i - 1, 1]) if (i and t[i, 2] > 0) else 0
    a = acc.setdefault(kind, [0, 0, 0, 0, 0])
    a[4] += t[i, 3]
    a[0] += op_ns
    a[1] += bar_ns
    a[2] += 1
    a[3] += x_ns
for k, (o, b, n, x, ld) in acc.items():
    print(f"{k:8s} n={n:3d} op avg {o / n / 1e3:8.2f} us (x staging {x / n / 1e3:6.2f} us, ring stages landed at start "
          f"{ld / n:4.1f})   barrier wait avg {b / n / 1e3:8.2f} us")
print("L2 prefetch stages issued by CTA 0:", eng.prefetch_stages)
# starvation: cycles warp 0 of the first / last CTA spent waiting for weights, per op kind, and as a share of the GEMV
st, dn = eng.trace_starved, eng.trace_done
for which, row in (("first CTA", st[0]), ("last CTA", st[1])):
    for j, kname in enumerate(names):
        if kname.startswith("attn"):
            continue
        cyc = row[j:n_layer_ops:per_layer].mean()
        print(f"{which}: {kname:8s} warp 0 starved {cyc / 1.965e3:7.2f} us per op")
# skew: op-done stamps over all CTAs
for j, kname in enumerate(names):
    d = dn[j:n_layer_ops:per_layer]                                   # [32 layers, grid]
    spread = d.max(axis=1) - d.min(axis=1)
    late = d.max(axis=1)[:, None] - d
    print(f"skew {kname:8s}: last-first CTA {spread.mean() / 1e3:6.2f} us, mean CTA finishes {late.mean() / 1e3:6.2f} us before the last; "
          f"slowest CTA ids (mode) {np.bincount(d.argmax(axis=1)).argmax()}")

print("first rows:", (t[:12] - start).tolist())

if per_layer == 3:   # anatomy of the attention + o projection op (CTA 0): attention items | wait + merge | o block
    idx = np.arange(1, n_layer_ops, 3)
    a_done, merged = eng.trace_starved[0][idx], eng.trace_starved[1][idx]
    start = t[idx - 1, 1]
    print("o block stages landed at the start of the o phase:", t[idx, 2].mean())
    print(f"attn_o anatomy: attention {np.mean(a_done - start) / 1e3:.2f} us, wait + merge {np.mean(merged - a_done) / 1e3:.2f} us, "
          f"o block stages {np.mean(t[idx, 3] - merged) / 1e3:.2f} us, reds + return {np.mean(t[idx, 0] - t[idx, 3]) / 1e3:.2f} us")

# per-CTA lateness (mean over layers of: op-done stamp minus the earliest CTA's), to see whether the skew is systematic
if len(sys.argv) > 3 and sys.argv[3] == "cta":
    for j, kname in enumerate(names):
        d = dn[j:n_layer_ops:per_layer].astype(np.float64)
        late = (d - d.min(axis=1, keepdims=True)).mean(axis=0) / 1e3
        order = np.argsort(-late)
        print(f"{kname}: lateness us by CTA, worst 12:", [(int(c), round(float(late[c]), 2)) for c in order[:12]])
        print(f"{kname}: best 8:", [(int(c), round(float(late[c]), 2)) for c in order[-8:]])
        print(f"{kname}: mean {late.mean():.2f} max {late.max():.2f}; by CTA index block of 16:",
              [round(float(late[i:i + 16].mean()), 2) for i in range(0, len(late), 16)])
        units = [(6144 * (c + 1)) // 148 - (6144 * c) // 148 for c in range(148)] if kname == "qkv" else None
        if units:
            print("qkv: mean lateness of CTAs with 42 row pairs:", round(float(np.mean([late[c] for c in range(148) if units[c] == 42])), 2),
                  "with 41:", round(float(np.mean([late[c] for c in range(148) if units[c] == 41])), 2))
