#!/usr/bin/env python
"""Timeline of one CTA of context_attn_tcgen05_kernel (the last query tile: 16 key tiles at S=2048):
per key tile, ns since the first stamp for S_j issued / P_j V_j issued / S_j visible / maxima
exchanged / P_j published / O_{j-1} folded (b200_context_attention_set_trace)."""
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from __graft_entry__ import load_package  # noqa: E402

pkg = load_package()
ops, capi = pkg.ops, pkg.capi
B, h, S, d = 4, 32, 2048, 128
q = (torch.rand(B, h, S, d, device="cuda") - 0.5).half()
k = (torch.rand(B, h, S, d, device="cuda") - 0.5).half()
v = (torch.rand(B, h, S, d, device="cuda") - 0.5).half()
out = torch.empty(B * S, h * d, dtype=torch.float16, device="cuda")
lens = torch.full((B,), S, dtype=torch.int32, device="cuda")
hist = torch.zeros(B, dtype=torch.int32, device="cuda")
pad = torch.zeros(B * S, dtype=torch.int32, device="cuda")
for _ in range(3):
    ops.context_attention(q, k, v, out, pad, lens, hist)
torch.cuda.synchronize()
tr = torch.zeros(8 * 16, dtype=torch.int64, device="cuda")
import ctypes as C
capi.lib().b200_context_attention_set_trace(C.c_void_p(tr.data_ptr()))
ops.context_attention(q, k, v, out, pad, lens, hist)
torch.cuda.synchronize()
capi.lib().b200_context_attention_set_trace(None)
t = tr.cpu().view(16, 8)
t0 = int(t[t > 0].min())
print("tile   S_iss  PV_iss  S_vis  maxx  P_pub  O_fold   (ns since first stamp)")
for j in range(16):
    print(f"{j:4d} " + " ".join(f"{(int(x) - t0) if x > 0 else -1:7d}" for x in t[j, :6]))
