"""CPU oracle of one decode step of the whole LLaMA decoder stack, composed from oracle/cpu.py in
the reference's launch order (self_decoder.cpp:53-85, masked_self_attention.cpp:55-92,
ffn.cpp:53-93, llama.cpp:319-361 with D18's TensorMap key fixed). `rt` is the storage rounding the
GPU path applies between kernels: identity for fp32, round-to-half for fp16.
TEST INFRASTRUCTURE ONLY."""
import numpy as np

from . import cpu


def rounder(kind):
    if kind == "fp16":
        return lambda a: np.ascontiguousarray(a.astype(np.float16).astype(np.float32))
    return lambda a: np.ascontiguousarray(a, dtype=np.float32)


class OracleModel:
    def __init__(self, weights, cfg, batch, max_seq_len, kind):
        """weights: dict with 'layers' (list of dicts attn_norm,qkv,o,ffn_norm,gate_up,down),
        'embed','final_norm','lm_head' — float32 arrays holding the (already rounded) values."""
        self.w, self.cfg, self.B, self.kind = weights, cfg, batch, kind
        self.rt = rounder(kind)
        L, hk, d = cfg["num_layers"], cfg["kv_head_num"], cfg["head_size"]
        self.k = np.zeros((L, batch, hk, max_seq_len, d), np.float32)
        self.v = np.zeros((L, batch, hk, max_seq_len, d), np.float32)
        self.step = 0   # tokens already cached

    def decode_step(self, ids):
        c, rt, B = self.cfg, self.rt, self.B
        h, hk, d, I = c["head_num"], c["kv_head_num"], c["head_size"], c["inter_size"]
        H = h * d
        eps, base, rot = c.get("rms_eps", 1e-5), c.get("rotary_base", 10000.0), c.get("rotary_dim", d)
        ref_step = self.step + 1
        x = np.zeros((B, H), np.float32)
        cpu.embedding(np.ascontiguousarray(ids, dtype=np.int32), self.w["embed"], x)
        for l, w in enumerate(self.w["layers"]):
            resid = np.zeros_like(x)
            cpu.rmsnorm(x, resid, w["attn_norm"], eps)
            x = rt(x)
            qkv = np.zeros((B, (h + 2 * hk) * d), np.float32)
            cpu.linear(x, w["qkv"], qkv, True)
            qkv = rt(qkv).reshape(B, h + 2 * hk, d)
            cpu.rope_decode(qkv, h, hk, ref_step, rot, base)
            qkv = rt(qkv)
            att = np.zeros((B, H), np.float32)
            cpu.decode_attention(qkv, self.k[l], self.v[l], att, h, hk, ref_step, True)
            att = rt(att)
            o = np.zeros((B, H), np.float32)
            cpu.linear(att, w["o"], o, True)
            o = rt(o)
            cpu.fused_add_bias_residual_rmsnorm(resid, o, None, w["ffn_norm"], eps)
            resid, o = rt(resid), rt(o)
            gu = np.zeros((B, 2 * I), np.float32)
            cpu.linear(o, w["gate_up"], gu, True)
            gu = rt(gu).reshape(B, 2, I)
            if self.kind == "fp16":   # the half kernel rounds silu(gate) before the product (act_kernel.cu:10)
                g = gu[:, 0]
                s = rt(g / (1.0 + np.exp(-g, dtype=np.float32)))
                act = rt(s * gu[:, 1])
            else:
                act = np.zeros((B, I), np.float32)
                cpu.swiglu(gu, act)
            dn = np.zeros((B, H), np.float32)
            cpu.linear(np.ascontiguousarray(act), w["down"], dn, True)
            dn = rt(dn)
            cpu.add_residual(resid, dn)
            x = rt(dn)
        self.hidden = x.copy()
        xn = x.copy()
        cpu.rmsnorm(xn, None, self.w["final_norm"], eps)
        xn = rt(xn)
        logits = np.zeros((B, c["vocab_size"]), np.float32)
        cpu.linear(xn, self.w["lm_head"], logits, True)
        logits = rt(logits)
        ids_out = np.zeros((B, 1), np.int32)
        cpu.topk(logits, ids_out, None, 1)
        self.step += 1
        return logits, ids_out.reshape(B)
