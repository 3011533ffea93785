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
    def __init__(self, weights, cfg, batch, max_seq_len, kind, allreduce=None):
        """weights: dict with 'layers' (list of dicts attn_norm,qkv,o,ffn_norm,gate_up,down),
        'embed','final_norm','lm_head' — float32 arrays holding the (already rounded) values.
        Tensor parallel: cfg holds this rank's head counts / FFN columns plus
        'hidden_size' (the unsplit H), weights are this rank's shard, and `allreduce(a)` sums a
        [n,H] partial result in place over the ranks after o_proj and after down_proj."""
        self.w, self.cfg, self.B, self.kind = weights, cfg, batch, kind
        self.rt = rounder(kind)
        self.allreduce = allreduce
        L, hk, d = cfg["num_layers"], cfg["kv_head_num"], cfg["head_size"]
        self.k = np.zeros((L, batch, hk, max_seq_len, d), np.float32)
        self.v = np.zeros((L, batch, hk, max_seq_len, d), np.float32)
        self.step = 0   # tokens already cached
        self.kv_round = None   # optional rows [..., d] -> rows: the rounding a cache applies to a row when it is appended

    def _ffn(self, w, o, resid, I):
        """gate_up GEMM -> SwiGLU -> down GEMM -> + residual (ffn.cpp:53-93, self_decoder.cpp:79-81)"""
        rt, n = self.rt, o.shape[0]
        gu = np.zeros((n, 2 * I), np.float32)
        cpu.linear(o, w["gate_up"], gu, True)
        gu = rt(gu).reshape(n, 2, I)
        if self.kind == "fp16":   # the half kernel rounds silu(gate) before the product (act_kernel.cu:10)
            g = gu[:, 0]
            s = rt(g / (1.0 + np.exp(-g, dtype=np.float32)))
            act = rt(s * gu[:, 1])
        else:
            act = np.zeros((n, I), np.float32)
            cpu.swiglu(gu, act)
        dn = np.zeros((n, o.shape[1]), np.float32)
        cpu.linear(np.ascontiguousarray(act), w["down"], dn, True)
        dn = rt(dn)
        if self.allreduce is not None:
            self.allreduce(dn)
            dn = rt(dn)
        cpu.add_residual(resid, dn)
        return rt(dn)

    def decode_layers(self, x):
        """the layer loop of LlamaSelfDecoder::forward (self_decoder.cpp:53-85) on hidden x [B,H];
        appends k,v at row self.step and advances self.step"""
        c, rt, B = self.cfg, self.rt, self.B
        h, hk, d, I = c["head_num"], c["kv_head_num"], c["head_size"], c["inter_size"]
        H, Hl = c.get("hidden_size", h * d), h * d   # Hl: this rank's attention columns
        eps, base, rot = c.get("rms_eps", 1e-5), c.get("rotary_base", 10000.0), c.get("rotary_dim", d)
        ref_step = self.step + 1
        x = np.ascontiguousarray(x, dtype=np.float32).copy()
        for l, w in enumerate(self.w["layers"]):
            resid = np.zeros_like(x)
            cpu.rmsnorm(x, resid, w["attn_norm"], eps)
            x = rt(x)
            qkv = np.zeros((B, (h + 2 * hk) * d), np.float32)
            cpu.linear(x, w["qkv"], qkv, True)
            qkv = rt(qkv).reshape(B, h + 2 * hk, d)
            cpu.rope_decode(qkv, h, hk, ref_step, rot, base)
            qkv = rt(qkv)
            att = np.zeros((B, Hl), np.float32)
            if self.kv_round is not None:
                # a cache that rounds rows on entry (fp8 KV): the new token's k, v are rounded, stored, and THEN attended
                # over with everything else - the step that appends a row already sees the rounded row
                self.k[l][:, :, ref_step - 1] = self.kv_round(qkv[:, h:h + hk])
                self.v[l][:, :, ref_step - 1] = self.kv_round(qkv[:, h + hk:h + 2 * hk])
                cpu.decode_attention(qkv, self.k[l], self.v[l], att, h, hk, ref_step, False)
            else:
                cpu.decode_attention(qkv, self.k[l], self.v[l], att, h, hk, ref_step, True)
            att = rt(att)
            o = np.zeros((B, H), np.float32)
            cpu.linear(att, w["o"], o, True)
            o = rt(o)
            if self.allreduce is not None:
                self.allreduce(o)
                o = rt(o)
            cpu.fused_add_bias_residual_rmsnorm(resid, o, None, w["ffn_norm"], eps)
            resid, o = rt(resid), rt(o)
            x = self._ffn(w, o, resid, I)
        self.step += 1
        return x

    def prefill_layers(self, x, input_len, history_len, max_q_len=None):
        """the layer loop of LlamaContextDecoder::forward (context_decoder.cpp:48-140) with
        LLaMAContextAttentionLayer::forward (context_attention.cpp:86-170, D12-D15 resolved) on
        x [num_tokens, H] (no padding). Caches are updated at rows history_len[b] + t."""
        c, rt, B = self.cfg, self.rt, self.B
        h, hk, d, I = c["head_num"], c["kv_head_num"], c["head_size"], c["inter_size"]
        H = h * d
        eps, base, rot = c.get("rms_eps", 1e-5), c.get("rotary_base", 10000.0), c.get("rotary_dim", d)
        input_len = np.ascontiguousarray(input_len, dtype=np.int32)
        history_len = np.ascontiguousarray(history_len, dtype=np.int32)
        ctx_len = (input_len + history_len).astype(np.int32)
        Nt, Sq, Sk = int(input_len.sum()), int(max_q_len or input_len.max()), int(ctx_len.max())
        pad = np.zeros((B, Sq), np.int32)
        cum = np.zeros(B + 1, np.int32)
        cpu.cal_padding_offset(pad, cum, input_len)
        pad_flat = np.ascontiguousarray(pad.reshape(-1)[:Nt])
        mask = np.zeros((B, Sq, Sk), np.float32)
        cpu.build_causal_mask(mask, input_len, ctx_len, True)
        x = np.ascontiguousarray(x, dtype=np.float32).copy()
        scale = 1.0 / np.sqrt(np.float32(d))
        for l, w in enumerate(self.w["layers"]):
            resid = np.zeros_like(x)
            cpu.rmsnorm(x, resid, w["attn_norm"], eps)
            x = rt(x)
            qkv = np.zeros((Nt, (h + 2 * hk) * d), np.float32)
            cpu.linear(x, w["qkv"], qkv, True)
            qkv = rt(qkv).reshape(Nt, h + 2 * hk, d)
            q = np.zeros((B, h, Sq, d), np.float32)
            k = np.zeros((B, hk, Sq, d), np.float32)
            v = np.zeros((B, hk, Sq, d), np.float32)
            cpu.qkv_bias_rope_transpose(q, k, v, qkv, None, pad_flat, history_len, rot, base)
            q, k, v = rt(q), rt(k), rt(v)
            cpu.concat_kv(k, self.k, l, input_len, history_len)
            cpu.concat_kv(v, self.v, l, input_len, history_len)
            kr = np.zeros((B, h, Sk, d), np.float32)
            vr = np.zeros((B, h, Sk, d), np.float32)
            cpu.repeat_kv(self.k, kr, l, ctx_len)
            cpu.repeat_kv(self.v, vr, l, ctx_len)
            qk = np.zeros((B, h, Sq, Sk), np.float32)
            cpu.batched_gemm(q, kr, qk, True)
            qk = rt(qk)
            cpu.scale_mask_softmax(qk, mask, qk, float(scale))
            qk = rt(qk)
            pv = np.zeros((B, h, Sq, d), np.float32)
            cpu.batched_gemm(qk, vr, pv, False)
            pv = rt(pv)
            att = np.zeros((Nt, H), np.float32)
            cpu.transpose_remove_padding(pv, att, pad_flat)
            o = np.zeros((Nt, H), np.float32)
            cpu.linear(att, w["o"], o, True)
            o = rt(o)
            cpu.fused_add_bias_residual_rmsnorm(resid, o, None, w["ffn_norm"], eps)
            resid, o = rt(resid), rt(o)
            x = self._ffn(w, o, resid, I)
        return x

    def decode_step(self, ids):
        c, rt, B = self.cfg, self.rt, self.B
        H = c.get("hidden_size", c["head_num"] * c["head_size"])
        eps = c.get("rms_eps", 1e-5)
        x = np.zeros((B, H), np.float32)
        cpu.embedding(np.ascontiguousarray(ids, dtype=np.int32), self.w["embed"], x)
        x = self.decode_layers(x)
        self.hidden = x.copy()
        xn = x.copy()
        cpu.rmsnorm(xn, None, self.w["final_norm"], eps)
        xn = rt(xn)
        logits = np.zeros((B, c["vocab_size"]), np.float32)
        cpu.linear(xn, self.w["lm_head"], logits, True)
        logits = rt(logits)
        ids_out = np.zeros((B, 1), np.int32)
        cpu.topk(logits, ids_out, None, 1)
        return logits, ids_out.reshape(B)
