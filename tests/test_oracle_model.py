"""CPU checks of the oracle's layer loops (oracle/model.py): prefill of S tokens must equal S
single-token decode steps (a size-independent property tying the context path — padding offsets,
causal mask with visible history, ConcatKV / RepeatKV, batched GEMMs, masked softmax — to the
decode path: RoPE at step-1, KV append, split-free attention)."""
import numpy as np

from conftest import rel_err
from oracle import model as omodel


def _model(seed, B, max_seq, h=4, hk=2, d=32, I=96, L=2):
    rng = np.random.default_rng(seed)
    H = h * d
    layers = [dict(attn_norm=(1 + 0.05 * rng.standard_normal(H)).astype(np.float32),
                   qkv=(0.08 * rng.standard_normal(((h + 2 * hk) * d, H))).astype(np.float32),
                   o=(0.08 * rng.standard_normal((H, H))).astype(np.float32),
                   ffn_norm=(1 + 0.05 * rng.standard_normal(H)).astype(np.float32),
                   gate_up=(0.08 * rng.standard_normal((2 * I, H))).astype(np.float32),
                   down=(0.08 * rng.standard_normal((H, I))).astype(np.float32)) for _ in range(L)]
    cfg = dict(num_layers=L, head_num=h, kv_head_num=hk, head_size=d, inter_size=I)
    return omodel.OracleModel(dict(layers=layers), cfg, B, max_seq, "fp32"), rng, H


def test_prefill_equals_token_by_token_decode():
    S = 9
    om_p, rng, H = _model(3, 1, 16)
    om_d, _, _ = _model(3, 1, 16)
    x = rng.uniform(-1, 1, (S, H)).astype(np.float32)
    want = np.concatenate([om_d.decode_layers(x[t:t + 1]) for t in range(S)])
    got = om_p.prefill_layers(x, [S], [0])
    assert rel_err(got, want) <= 1e-5
    assert rel_err(om_p.k[:, :, :, :S], om_d.k[:, :, :, :S]) <= 1e-5
    assert rel_err(om_p.v[:, :, :, :S], om_d.v[:, :, :, :S]) <= 1e-5


def test_prefill_with_history_equals_continued_decode():
    S0, S1 = 5, 4
    om_p, rng, H = _model(4, 1, 16)
    om_d, _, _ = _model(4, 1, 16)
    x = rng.uniform(-1, 1, (S0 + S1, H)).astype(np.float32)
    want = np.concatenate([om_d.decode_layers(x[t:t + 1]) for t in range(S0 + S1)])
    om_p.prefill_layers(x[:S0], [S0], [0])
    got = om_p.prefill_layers(x[S0:], [S1], [S0])      # second turn: history visible
    assert rel_err(got, want[S0:]) <= 1e-5


def test_ragged_batch_prefill_equals_per_sequence_prefill():
    lens = [3, 7, 5]
    om_b, rng, H = _model(5, 3, 12)
    x = rng.uniform(-1, 1, (sum(lens), H)).astype(np.float32)
    got = om_b.prefill_layers(x, lens, [0, 0, 0])
    off = 0
    for n in lens:
        om_1, _, _ = _model(5, 1, 12)
        want = om_1.prefill_layers(x[off:off + n], [n], [0])
        assert rel_err(got[off:off + n], want) <= 1e-5
        off += n
