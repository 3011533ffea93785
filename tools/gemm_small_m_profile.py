#!/usr/bin/env python
"""A few launches of the swapped (small-M, weight-streaming) tcgen05 GEMM on the batch-32 decode shapes, for
`ncu --set full -k regex:gemm_tcgen05` (weights cycle through > 300 MB so HBM, not L2, is measured)."""
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from __graft_entry__ import load_package  # noqa: E402

ops = load_package().ops
m = int(sys.argv[1]) if len(sys.argv) > 1 else 32
for n, k in ((12288, 4096), (22016, 4096), (4096, 11008)):
    nw = max(1, int(300e6 // (n * k * 2)) + 1)
    ws = [(torch.randn(n, k, device="cuda") * 0.02).half() for _ in range(nw)]
    x = (torch.rand(m, k, device="cuda") - 0.5).half()
    y = torch.empty(m, n, dtype=torch.float16, device="cuda")
    for i in range(2 * nw):
        ops.launchLinearGemm(x, ws[i % nw], y, trans_b=True)
    torch.cuda.synchronize()
    del ws
print("ok")
