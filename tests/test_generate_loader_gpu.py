"""b200_engine_load_weights (the reference's weight directory: raw fp32 .bin files named as in
llama_weights.cc:41-53 / layer_weights.cc:48-66, converted to the engine dtype, Megatron-sharded for
tensor parallel ranks) and b200_engine_generate (the loop of Llama<T>::Response, llama.cpp:363-448,
without the tokenizer). Loader: bit-exact against set_weight of the same arrays / of
sharding.tp_shard_layer. Generate: identical ids to prefill + decode_step_host driven by hand."""
import os

import numpy as np
import pytest

from test_engine_gpu import SMALL
from test_sharding_gloo import make_weights

pytestmark = pytest.mark.gpu
FILES = {"attn_norm": "input_layernorm.weight", "ffn_norm": "post_attention_layernorm.weight",
         "qkv": "self_attn.qkv.weight", "o": "self_attn.o_proj.weight",
         "gate_up": "mlp.gate_up_proj.weight", "down": "mlp.down_proj.weight"}


def write_dir(tmp, full):
    for l, w in enumerate(full["layers"]):
        for key, name in FILES.items():
            np.asarray(w[key], np.float32).tofile(os.path.join(tmp, f"model.layers.{l}.{name}.bin"))
    np.asarray(full["final_norm"], np.float32).tofile(os.path.join(tmp, "model.norm.weight.bin"))
    np.asarray(full["lm_head"], np.float32).tofile(os.path.join(tmp, "lm_head.weight.bin"))
    np.asarray(full["embed"], np.float32).tofile(os.path.join(tmp, "model.embed_tokens.weight.bin"))


@pytest.mark.parametrize("kind", ["fp16", "fp32"])
@pytest.mark.parametrize("tp,rank", [(1, 0), (2, 0), (2, 1)])
def test_load_reference_weight_directory(pkg, tmp_path, kind, tp, rank):
    E, sh = pkg.engine, pkg.sharding
    full = make_weights(SMALL, kind, 21)
    write_dir(str(tmp_path), full)
    eng = E.Engine(1, 16, dtype=pkg.F16 if kind == "fp16" else pkg.F32, tp_size=tp, tp_rank=rank, **SMALL)
    eng.load_weights(tmp_path)
    for l, w in enumerate(full["layers"]):
        want = sh.tp_shard_layer(w, SMALL, tp, rank) if tp > 1 else w
        for slot, key in ((E.W_ATTN_NORM, "attn_norm"), (E.W_QKV, "qkv"), (E.W_O, "o"), (E.W_FFN_NORM, "ffn_norm"),
                          (E.W_GATE_UP, "gate_up"), (E.W_DOWN, "down")):
            got = np.asarray(eng.get_weight(slot, l), np.float32)
            assert np.array_equal(got.reshape(-1), np.asarray(want[key], np.float32).reshape(-1)), (l, key)
    for slot, key in ((E.W_EMBED, "embed"), (E.W_FINAL_NORM, "final_norm"), (E.W_LM_HEAD, "lm_head")):
        assert np.array_equal(np.asarray(eng.get_weight(slot), np.float32).reshape(-1),
                              np.asarray(full[key], np.float32).reshape(-1)), key
    eng.close()
    # a truncated file must be reported, not silently accepted
    with open(os.path.join(tmp_path, "model.norm.weight.bin"), "ab") as f:
        f.write(b"\0\0\0\0")
    eng = E.Engine(1, 16, dtype=pkg.F16 if kind == "fp16" else pkg.F32, **SMALL)
    with pytest.raises(RuntimeError):
        eng.load_weights(tmp_path)
    eng.close()


def test_generate_equals_manual_loop(pkg):
    E = pkg.engine
    B, n, new = 2, 23, 7
    a = E.Engine(B, 64, dtype=pkg.F16, **SMALL)
    b = E.Engine(B, 64, dtype=pkg.F16, **SMALL)
    a.init_random(seed=4, std=0.05)
    b.init_random(seed=4, std=0.05)
    prompt = np.random.default_rng(2).integers(0, SMALL["vocab_size"], (B, n)).astype(np.int32)
    want = [b.prefill(prompt)]
    for _ in range(new - 1):
        want.append(b.decode_step_host(None))
    want = np.stack(want, axis=1)                       # [B, new]
    got, cnt = a.generate(prompt, new, eos_id=-1)       # no EOS: every sequence emits `new` tokens
    assert np.array_equal(got, want) and cnt.tolist() == [new, new]
    assert a.step == n + new - 1
    # EOS: sequence 0 stops where its third token would be the EOS id; the other runs on
    a.reset()
    eos = int(want[0, 2])
    got, cnt = a.generate(prompt, new, eos_id=eos)
    stop0 = 1 + next(i for i in range(1, new) if want[0, i] == eos) - 1   # first EOS at position >= 1
    assert cnt[0] == stop0 and np.array_equal(got[0, :stop0], want[0, :stop0]) and np.all(got[0, stop0:] == eos)
    if eos not in want[1, 1:].tolist():
        assert cnt[1] == new and np.array_equal(got[1], want[1])
    a.close()
    b.close()
