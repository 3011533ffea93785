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
names = ["qkv", "attn", "o", "gate_up", "down"]
start = t[0, 0]
print("kernel start -> first op done: n/a; total ns:", t[-1, 0] - (t[0, 0]), "ops:", len(t))
prev = None
acc = {}
for i in range(len(t)):
    kind = names[i % 5] if i < 160 else ("lm_head" if i == 160 else "final")
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
print("first rows:", (t[:12] - start).tolist())
ct = eng.cta_trace
if ct is not None:
    ct = ct.astype(np.int64)
    smid = ct[:, -1]
    done = ct[:, :-1]
    print("per-op finish-time skew across CTAs (us): op kind, median (max - min), median (max - median)")
    for k, kind in enumerate(names):
        cols = [i for i in range(5 * 8, 160) if i % 5 == k]
        rng = [(done[:, i].max() - done[:, i].min()) / 1e3 for i in cols]
        top = [(done[:, i].max() - np.median(done[:, i])) / 1e3 for i in cols]
        print(f"  {kind:8s} {np.median(rng):6.2f} {np.median(top):6.2f}")
    # is lateness systematic per SM? average rank of each CTA's finish time over the gate_up ops
    cols = [i for i in range(40, 160) if i % 5 == 3]
    late = np.stack([done[:, i] - np.median(done[:, i]) for i in cols], 1) / 1e3   # [cta, op] us vs median
    mean_late = late.mean(1)
    order = np.argsort(mean_late)
    print("gate_up: mean lateness vs median per CTA (us): min %.2f  p10 %.2f  median %.2f  p90 %.2f  max %.2f; per-CTA std over layers (median) %.2f"
          % (mean_late.min(), np.percentile(mean_late, 10), np.median(mean_late), np.percentile(mean_late, 90),
             mean_late.max(), np.median(late.std(1))))
    print("latest 12 CTAs (cta, smid, mean lateness us):", [(int(c), int(smid[c]), round(float(mean_late[c]), 2)) for c in order[-12:]])
    print("earliest 12 CTAs:", [(int(c), int(smid[c]), round(float(mean_late[c]), 2)) for c in order[:12]])
if ct is not None:
    for k, kind in enumerate(names):
        cols = [i for i in range(40, 160) if i % 5 == k]
        lt = np.stack([done[:, i] - np.median(done[:, i]) for i in cols], 1) / 1e3
        ml = lt.mean(1)
        od = np.argsort(ml)
        print(f"{kind}: latest 8 CTAs (cta, smid, mean lateness us, std):",
              [(int(c), int(smid[c]), round(float(ml[c]), 2), round(float(lt[c].std()), 2)) for c in od[-8:]],
              " p50/p90/p99 of all (cta, op) lateness: %.2f %.2f %.2f" % tuple(np.percentile(lt, [50, 90, 99])))
    by_sm = sorted((int(smid[c]), round(float(mean_late[c]), 2)) for c in range(len(smid)))
    print("gate_up mean lateness by smid:", by_sm)
