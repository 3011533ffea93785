"""Continuous batching, the data-path half: b200_engine_reset_slot + b200_engine_prefill_ragged with zero-length entries.

The reference's loop (llama.cpp:319-361) runs one fixed batch to the end; a slot whose sequence has finished stays occupied.
Here a slot can be handed to the next request while the other sequences keep their caches, their positions and the token
they were about to decode. Checked, teacher-forced so that a near-tie cannot fork the chains:
  * the sequences that sat the refill out produce the logits of an UNDISTURBED batch run, step for step;
  * the refilled slot produces the logits of its prompt generated ALONE in a one-sequence engine;
  * (paged cache) the slot's pages go back to the pool on reset and are taken again by the new prompt."""
import numpy as np
import pytest

from conftest import assert_close
from test_engine_gpu import SMALL

pytestmark = pytest.mark.gpu

P = 128


def _logits(eng):
    return eng.logits().float().cpu().numpy()


@pytest.mark.parametrize("paged", [False, True])
def test_slot_refill_keeps_the_other_sequences_and_matches_a_solo_run(pkg, paged):
    E = pkg.engine
    B, max_seq, T1, T2 = 3, 420, 5, 6
    rng = np.random.default_rng(17)
    prompts = [rng.integers(0, SMALL["vocab_size"], n).astype(np.int32) for n in (140, 30, 50)]
    newp = rng.integers(0, SMALL["vocab_size"], 135).astype(np.int32)
    mk = lambda b, **kw: E.Engine(b, max_seq, dtype=pkg.F16, **kw, **SMALL)

    # (a) the undisturbed batch: first tokens + T1 + T2 greedy steps, logits recorded
    free = mk(B)
    free.init_random(seed=4, std=0.05)
    free.set_mode(fused=False, graph=True)
    ids = free.prefill_ragged(prompts)
    free_ids, free_logits = [ids.copy()], [_logits(free)]
    for _ in range(T1 + T2):
        ids = free.decode_step_host(ids)
        free_ids.append(ids.copy())
        free_logits.append(_logits(free))
    free.close()

    # (b) the new prompt alone
    solo = mk(1)
    solo.init_random(seed=4, std=0.05)
    solo.set_mode(fused=False, graph=True)
    ids = solo.prefill(newp[None, :])
    solo_ids, solo_logits = [ids.copy()], [_logits(solo)]
    for _ in range(T2):
        ids = solo.decode_step_host(ids)
        solo_ids.append(ids.copy())
        solo_logits.append(_logits(solo))
    solo.close()

    # (c) the scenario: T1 steps, slot 1 handed to the new prompt, T2 more steps - fed with the reference runs' tokens
    eng = mk(B, kv_paged=True, kv_pages=12) if paged else mk(B)
    eng.init_random(seed=4, std=0.05)
    if not paged:
        eng.set_mode(fused=False, graph=True)
    got = eng.prefill_ragged(prompts)
    assert_close(_logits(eng), free_logits[0], "fp16", "prompt pass")
    for t in range(T1):
        last_out = eng.decode_step_host(free_ids[t])
        assert_close(_logits(eng), free_logits[t + 1], "fp16", f"step {t} before the refill")
    if paged:
        used = eng.kv_pages_total - eng.kv_pages_free
        assert used == sum((len(p) + T1 + P - 1) // P for p in prompts)
    eng.reset_slot(1)
    if paged:
        assert eng.kv_pages_total - eng.kv_pages_free == used - (30 + T1 + P - 1) // P, "the slot's pages go back to the pool"
    first = eng.prefill_ragged([[], newp, []])
    assert first[0] == last_out[0] and first[2] == last_out[2], "sequences that sat the pass out keep their next token"
    assert_close(_logits(eng)[1:2], solo_logits[0], "fp16", "refilled slot: prompt-pass logits of the solo run")
    assert list(eng.seq_steps) == [140 + T1, 135, 50 + T1]   # rows cached per sequence
    if paged:
        assert eng.kv_pages_total - eng.kv_pages_free == used - 1 + 2, "135 prompt rows = two pages"
    for t in range(T2):
        feed = np.array([free_ids[T1 + t][0], solo_ids[t][0], free_ids[T1 + t][2]], np.int32)
        eng.decode_step_host(feed)
        lg = _logits(eng)
        assert_close(lg[[0, 2]], free_logits[T1 + t + 1][[0, 2]], "fp16", f"undisturbed sequences, step {t} after the refill")
        assert_close(lg[1:2], solo_logits[t + 1], "fp16", f"refilled slot, step {t}")
    eng.close()


def test_zero_length_entries_are_refused_where_they_cannot_work(pkg):
    E = pkg.engine
    eng = E.Engine(2, 64, dtype=pkg.F16, **SMALL)
    eng.init_random(seed=1, std=0.05)
    with pytest.raises(RuntimeError, match="no prompt tokens"):
        eng.prefill_ragged([[], []])
    with pytest.raises(RuntimeError, match="bad slot"):
        eng.reset_slot(2)
    eng.close()
    e32 = E.Engine(2, 64, dtype=pkg.F32, **SMALL)   # fp32 runs the reference-order attention: everybody takes part
    e32.init_random(seed=1, std=0.05)
    with pytest.raises(RuntimeError, match="sit a pass out"):
        e32.prefill_ragged([[1, 2, 3], []])
    e32.close()
