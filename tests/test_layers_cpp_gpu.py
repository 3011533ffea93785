"""The reference-shaped C++ layer classes (LlamaSelfDecoder / LlamaContextDecoder and the attention /
FFN layers under llm-inference_b200/include/src/layers, written over the header-only launch* shim
and libb200llama.so) driven by llm-inference_b200/bin/layer_driver on data written by this test,
compared with the CPU oracle's layer loops (oracle/model.py). fp32 rel <= 1e-5, fp16 <= 2e-2;
the KV-cache rows written by ConcatKV (prefill) are compared at the same tolerance (they hold
RoPE'd keys), untouched rows bit-exact."""
import os
import subprocess

import numpy as np
import pytest

from conftest import ROOT, assert_close
from oracle import model as omodel

pytestmark = pytest.mark.gpu
DRIVER = os.path.join(ROOT, "llm-inference_b200", "bin", "layer_driver")
NAMES = {"attn_norm": "input_layernorm.weight", "ffn_norm": "post_attention_layernorm.weight",
         "qkv": "self_attn.qkv.weight", "o": "self_attn.o_proj.weight",
         "gate_up": "mlp.gate_up_proj.weight", "down": "mlp.down_proj.weight"}


def widen(a, kind):
    a = np.asarray(a, dtype=np.float32)
    return np.ascontiguousarray(a.astype(np.float16).astype(np.float32) if kind == "fp16" else a)


def make_case(tmp, kind, cfg, B, max_seq, seed, extra):
    rng = np.random.default_rng(seed)
    L, h, hk, d, I = cfg["num_layers"], cfg["head_num"], cfg["kv_head_num"], cfg["head_size"], cfg["inter_size"]
    H = h * d
    layers = []
    for l in range(L):
        w = dict(attn_norm=widen(1 + 0.05 * rng.standard_normal(H), kind),
                 qkv=widen(0.05 * rng.standard_normal(((h + 2 * hk) * d, H)), kind),
                 o=widen(0.05 * rng.standard_normal((H, H)), kind),
                 ffn_norm=widen(1 + 0.05 * rng.standard_normal(H), kind),
                 gate_up=widen(0.05 * rng.standard_normal((2 * I, H)), kind),
                 down=widen(0.05 * rng.standard_normal((H, I)), kind))
        for key, name in NAMES.items():
            w[key].tofile(os.path.join(tmp, f"model.layers.{l}.{name}.bin"))
        layers.append(w)
    kc = widen(rng.uniform(-1, 1, (L, B, hk, max_seq, d)), kind)
    vc = widen(rng.uniform(-1, 1, (L, B, hk, max_seq, d)), kind)
    kc.tofile(os.path.join(tmp, "k_cache.bin"))
    vc.tofile(os.path.join(tmp, "v_cache.bin"))
    lines = dict(dtype=kind, layers=L, head_num=h, kv_head_num=hk, head_size=d, inter_size=I, batch=B,
                 max_seq=max_seq, **extra)
    with open(os.path.join(tmp, "cfg.txt"), "w") as f:
        f.write("".join(f"{k}={v}\n" for k, v in lines.items()))
    om = omodel.OracleModel(dict(layers=layers), cfg, B, max_seq, kind)
    om.k[:], om.v[:] = kc, vc
    return om, rng


def run_driver(tmp):
    if not os.path.exists(DRIVER):
        pytest.fail(f"{DRIVER} is missing: build it with __graft_entry__.build()")
    r = subprocess.run([DRIVER, str(tmp)], capture_output=True, text=True, timeout=600)
    assert r.returncode == 0, r.stdout + r.stderr
    assert "kernel launches through libb200llama.so" in r.stdout
    return r.stdout


CFGS = {"mha": dict(num_layers=2, head_num=4, kv_head_num=4, head_size=128, inter_size=688),
        "gqa": dict(num_layers=2, head_num=8, kv_head_num=2, head_size=128, inter_size=1024)}


@pytest.mark.parametrize("kind", ["fp32", "fp16"])
@pytest.mark.parametrize("name,B,context,steps", [("mha", 1, 64, 2), ("gqa", 3, 37, 2)])
def test_self_decoder_cpp_vs_oracle(tmp_path, kind, name, B, context, steps):
    cfg, max_seq = CFGS[name], 96
    H = cfg["head_num"] * cfg["head_size"]
    om, rng = make_case(str(tmp_path), kind, cfg, B, max_seq, 11, dict(mode="decode", steps=steps, context=context))
    xin = widen(rng.uniform(-1, 1, (steps, B, H)), kind)
    xin.tofile(os.path.join(tmp_path, "hidden_in.bin"))
    om.step = context
    want = np.stack([om.decode_layers(xin[s]) for s in range(steps)])
    run_driver(tmp_path)
    got = np.fromfile(os.path.join(tmp_path, "hidden_out.bin"), np.float32).reshape(steps, B, H)
    assert_close(got, want, kind, "LlamaSelfDecoder::forward")
    gk = np.fromfile(os.path.join(tmp_path, "k_cache_out.bin"), np.float32).reshape(om.k.shape)
    gv = np.fromfile(os.path.join(tmp_path, "v_cache_out.bin"), np.float32).reshape(om.v.shape)
    assert_close(gk, om.k, kind, "k cache after decode")
    assert_close(gv, om.v, kind, "v cache after decode")
    rows = np.ones(max_seq, bool)
    rows[context:context + steps] = False
    assert np.array_equal(gk[:, :, :, rows], om.k[:, :, :, rows]), "rows not appended to must be untouched"


@pytest.mark.parametrize("fused", [0, 1])
@pytest.mark.parametrize("kind", ["fp32", "fp16"])
@pytest.mark.parametrize("name,lens,hist", [("mha", [7], [0]), ("gqa", [5, 16, 9], [0, 0, 0]),
                                            ("gqa", [6, 3], [10, 4])])
def test_context_decoder_cpp_vs_oracle(tmp_path, kind, name, lens, hist, fused):
    cfg, max_seq, B = CFGS[name], 48, len(lens)
    H = cfg["head_num"] * cfg["head_size"]
    om, rng = make_case(str(tmp_path), kind, cfg, B, max_seq, 13, dict(mode="prefill", fused_attention=fused))
    nt = sum(lens)
    xin = widen(rng.uniform(-1, 1, (nt, H)), kind)
    xin.tofile(os.path.join(tmp_path, "hidden_in.bin"))
    np.asarray(lens, np.int32).tofile(os.path.join(tmp_path, "input_length.bin"))
    np.asarray(hist, np.int32).tofile(os.path.join(tmp_path, "history_length.bin"))
    want = om.prefill_layers(xin, lens, hist)
    run_driver(tmp_path)
    got = np.fromfile(os.path.join(tmp_path, "hidden_out.bin"), np.float32).reshape(nt, H)
    assert_close(got, want, kind, "LlamaContextDecoder::forward")
    gk = np.fromfile(os.path.join(tmp_path, "k_cache_out.bin"), np.float32).reshape(om.k.shape)
    gv = np.fromfile(os.path.join(tmp_path, "v_cache_out.bin"), np.float32).reshape(om.v.shape)
    assert_close(gk, om.k, kind, "k cache after prefill")
    assert_close(gv, om.v, kind, "v cache after prefill")
