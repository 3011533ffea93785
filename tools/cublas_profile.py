#!/usr/bin/env python
"""The cuBLAS fp16 GEMM the reference would call (cublas_utils.cc:52) on one 7B prefill shape, for an ncu capture next to
gemm2_tcgen05_kernel's: y[m, n] = x[m, k] @ W[n, k]^T, weights cycling through > 300 MB."""
import sys

import torch

m = int(sys.argv[1]) if len(sys.argv) > 1 else 8192
n, k = 12288, 4096
ws = [(torch.randn(n, k, device="cuda") * 0.02).half() for _ in range(4)]
x = (torch.rand(m, k, device="cuda") - 0.5).half()
for i in range(8):
    y = torch.nn.functional.linear(x, ws[i % 4])
torch.cuda.synchronize()
print("ok", float(y[0, 0]))
