"""Sequence retirement - the reference's `finished` flags (src/kernels/sampling.cu:46,81-83: finished[b] = (id == end_id);
llama.cpp:189,327 carry them through the loop; launchDecoderMaskedMHA receives them and never reads them).
With b200_engine_set_eos a sequence that has produced eos_id is flagged on the device, keeps emitting eos_id, and from the
next step on neither appends to nor reads its KV cache. Checked: the other sequences decode exactly the tokens they
decode without retirement; the retired sequence emits eos_id; in the reference launch order (append inside the
attention kernel) its cache rows after retirement stay untouched."""
import numpy as np
import pytest

from test_engine_gpu import SMALL

pytestmark = pytest.mark.gpu


@pytest.mark.parametrize("mode,B", [("unfused", 3), ("fused_graph", 3), ("unfused", 6)])
def test_retired_sequence_stops_touching_its_cache(pkg, mode, B):
    E = pkg.engine
    max_seq, steps = 40, 10

    def run(eos):
        eng = E.Engine(B, max_seq, dtype=pkg.F16, **SMALL)
        eng.init_random(seed=11, std=0.05)
        eng.set_mode(fused=mode != "unfused", graph=mode == "fused_graph")
        if eos is not None:
            eng.set_eos(eos)
        ids = np.arange(5, 5 + B).astype(np.int32)
        toks, fin = [], []
        for s in range(steps):
            ids = eng.decode_step_host(ids if s == 0 else None)
            toks.append(ids.copy())
            fin.append(eng.finished().copy())
        kc, _ = eng.kv_caches()
        k = kc.float().cpu().numpy()
        eng.close()
        return np.stack(toks), np.stack(fin), k

    free, fin0, _ = run(None)
    assert not fin0.any()
    # retire sequence 1 with the token it produces at step 2 (and whoever else happens to produce it later)
    eos = int(free[2, 1])
    toks, fin, k = run(eos)
    hit = {b: int(np.argmax(free[:, b] == eos)) for b in range(B) if (free[:, b] == eos).any()}
    assert 1 in hit and hit[1] <= 2
    for b in range(B):
        if b in hit:
            t = hit[b]
            assert np.array_equal(toks[:t + 1, b], free[:t + 1, b]), "tokens up to and including eos are unchanged"
            assert np.all(toks[t:, b] == eos), "a retired sequence keeps emitting eos_id"
            assert not fin[:t, b].any() and fin[t:, b].all(), "flag set by the step that produced eos_id"
            if mode == "unfused":   # rows t+1.. would have been appended by the following steps
                assert np.all(k[:, b, :, t + 1:] == 0), "retired sequence still appended to its cache"
                assert np.any(k[:, b, :, t] != 0)
        else:
            assert np.array_equal(toks[:, b], free[:, b]), "an unretired sequence must decode the same tokens"
            assert not fin[:, b].any()
