"""Regenerates tests/golden/ref_cpu_golden.npz from the REFERENCE's own CPU test functions
(oracle/_ref/libref.so, built by `make -C oracle ref` from /root/reference/tests/unittests/*.cu)
run on the reference tests' own input data. Run in the build container only:

    python tests/golden/make_golden.py

The fixture pins the oracle (tests/test_oracle_golden.py) wherever /root/reference is absent.
"""
import ctypes as C
import os

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))
ref = C.CDLL(os.path.join(ROOT, "oracle", "_ref", "libref.so"))


def p(a):
    return a.ctypes.data_as(C.c_void_p)


out = {}

# test_rmsnorm.cu:139-192 — 64 x 4096, x = i%2+1, gamma = i%2+1, eps 1e-6
nt, H = 64, 4096
x = (np.arange(nt * H) % 2 + 1).astype(np.float32).reshape(nt, H)
g = (np.arange(H) % 2 + 1).astype(np.float32)
ref.ref_cpu_rmsnorm(p(x), p(g), C.c_float(1e-6), H, nt)
out["rmsnorm_row0"] = x[0].copy()
out["rmsnorm_row63"] = x[63].copy()

# test_fused_addresidual_norm.cu:69-141 — Nt=2, H=32, eps 0.5, out=1, resid=0, bias=0, gamma=1
nt, H = 2, 32
o = np.ones((nt, H), np.float32)
r = np.zeros((nt, H), np.float32)
b = np.zeros(H, np.float32)
gm = np.ones(H, np.float32)
ref.ref_cpu_fused_add_norm(p(r), p(o), p(b), p(gm), C.c_float(0.5), H, nt)
out["addnorm_out"] = o.copy()

# test_act.cu:46-57 — B=16, I=11008, all ones
B, I = 16, 11008
gi = np.ones((B, 2, I), np.float32)
go = np.zeros((B, I), np.float32)
ref.ref_cpu_swiglu(p(gi), p(go), B, I)
out["swiglu_first8"] = go.reshape(-1)[:8].copy()
# and a non-degenerate ramp through the same reference function
gi = np.linspace(-6, 6, 2 * 2 * 64, dtype=np.float32).reshape(2, 2, 64)
go = np.zeros((2, 64), np.float32)
ref.ref_cpu_swiglu(p(gi), p(go), 2, 64)
out["swiglu_ramp_in"] = gi.copy()
out["swiglu_ramp_out"] = go.copy()

# test_bias_and_RoPE.cu:99-160 — B=1, S=32, h=hk=32, d=128, QKV = 32.0, history 0
Bq, S, h, hk, d = 1, 32, 32, 32, 128
q = np.zeros((Bq, S, h, d), np.float32)
k = np.zeros((Bq, S, hk, d), np.float32)
v = np.zeros((Bq, S, hk, d), np.float32)
QKV = np.full((Bq * S, h + 2 * hk, d), 32.0, np.float32)
bias = np.zeros((h + 2 * hk) * d, np.float32)
pad = np.zeros(Bq * S, np.int32)
hist = np.zeros(Bq, np.int32)
ilen = np.full(Bq, 7, np.int32)
ref.ref_cpu_rope(p(q), p(k), p(v), p(QKV), p(bias), p(pad), p(hist), p(ilen), Bq, S, Bq * S, h, hk,
                 d, 128, C.c_float(10000.0))
out["rope_q_head0"] = q[0, :, 0, :].copy()   # [S, d]: identical for every head (constant data)
out["rope_k_head0"] = k[0, :, 0, :].copy()

# test_linear.cu:39-60 — 13x4096 . 4096x4096, data i%3, no transposition
m, kk, n = 13, 4096, 4096
a = (np.arange(m * kk) % 3).astype(np.float32).reshape(m, kk)
w = (np.arange(kk * n) % 3).astype(np.float32).reshape(kk, n)
y = np.zeros((m, n), np.float32)
ref.ref_cpu_linear(p(a), p(w), p(y), m, kk, n)
out["linear_cols64"] = y[:, :64].copy()

# test_casual_mask.cu:45-47 — B=1, 5x5, q = k lens; plus a history case through the same function
mask = np.zeros((1, 5, 5), np.float32)
ql = np.array([5], np.int32)
kl = np.array([5], np.int32)
ref.ref_cpu_causal_mask(p(mask), p(ql), p(kl), 5, 5, 1)
out["mask_5x5"] = mask.copy()
mask = np.zeros((2, 4, 7), np.float32)
ql = np.array([3, 4], np.int32)
kl = np.array([6, 4], np.int32)
ref.ref_cpu_causal_mask(p(mask), p(ql), p(kl), 4, 7, 2)
out["mask_hist"] = mask.copy()

# test_cal_paddingoffset.cu:54-56 — golden vector written in the reference's comment
out["padding_offset_golden"] = np.array([0, 0, 0, 2, 2, 2, 4, 4, 4], np.int32)
out["cum_seqlens_golden"] = np.array([0, 3, 6, 9], np.int32)

np.savez_compressed(os.path.join(HERE, "ref_cpu_golden.npz"), **out)
print("wrote", {k_: v_.shape for k_, v_ in out.items()})
