"""Continuous batching over the engine (dis_serving_*): more requests than engine slots, mixed lengths and
priorities; every request completes, gets one event per token in order, and produces the tokens it produces
when it runs alone (greedy; divergences only at sub-noise logit margins, which are reported)."""
import numpy as np
import pytest

from distributed_inference_server_b200 import ContinuousBatcher, Engine, InferenceParams, WorkerConfig
from distributed_inference_server_b200 import binding as B
from oracle import oracle as O
from tests._gpu_util import to_engine_shape

pytestmark = pytest.mark.gpu
SHAPE = O.ModelShape(2048, 512, 3, 8, 2, 64, 1408, 512)


def test_continuous_batching_serves_every_request():
    e = Engine(WorkerConfig(to_engine_shape(SHAPE), max_batch=6, kv_pages=96, page_size=16, max_step_tokens=256))
    e.load_synthetic(31); e.finalize()
    rng = np.random.default_rng(5)
    n_req = 23
    prompts = [rng.integers(0, SHAPE.vocab, int(rng.integers(1, 120))).tolist() for _ in range(n_req)]
    new = [int(rng.integers(1, 20)) for _ in range(n_req)]
    cb = ContinuousBatcher(e, max_active=6)
    for i in range(n_req):
        cb.submit(i, prompts[i], InferenceParams(new[i], 0.0), priority=int(rng.integers(0, 3)))
    first_seen, events = [], []
    for _ in range(10000):
        evs = cb.step()
        events += evs
        for ev in evs:
            if ev[1] == B.EVENT_TOKEN and ev[3] == 0:
                first_seen.append(ev[0])
        d = cb.depth()
        assert d.active <= 6
        if d.total == 0 and d.active == 0 and not evs:
            break
    # N requests in -> N results out, one DONE each, usage counts right (Usage, models.rs:16-22)
    done = {ev[0]: ev for ev in events if ev[1] == B.EVENT_DONE}
    assert sorted(done) == list(range(n_req))
    toks = {i: [ev[2] for ev in events if ev[0] == i and ev[1] == B.EVENT_TOKEN] for i in range(n_req)}
    for i in range(n_req):
        idx = [ev[3] for ev in events if ev[0] == i and ev[1] == B.EVENT_TOKEN]
        assert idx == list(range(new[i]))                      # one event per token, in order
        assert done[i][3] == new[i] and done[i][4] == B.FINISH_LENGTH and done[i][5] == len(prompts[i])
    assert e.status().kv_pages_free == e.status().kv_pages_total and e.status().active_batches == 0
    # each request matches the oracle run of that request alone
    mimic = O.Oracle(SHAPE, act_round=True); mimic.fill_synthetic(31, True)
    flipped = 0
    for i in range(n_req):
        w = mimic.generate([prompts[i]], new[i], n_logit_steps=new[i])
        if toks[i] != w["tokens"][0]:
            k = next(j for j in range(new[i]) if toks[i][j] != w["tokens"][0][j])
            margin, scale = float(w["margins"][0, k]), float(np.abs(w["logits"][0, k]).max())
            print(f"[margin report] request {i}: first difference at step {k}, margin {margin:.2e} of max|logit| {scale:.2f}")
            assert margin <= 5e-3 * scale
            flipped += 1
    assert flipped <= n_req // 3
    cb.close(); e.close()


def test_priority_decides_admission_order_and_bad_requests_fail_alone():
    e = Engine(WorkerConfig(to_engine_shape(SHAPE), max_batch=2, kv_pages=64, page_size=16, max_step_tokens=256))
    e.load_synthetic(31); e.finalize()
    cb = ContinuousBatcher(e, max_active=2)
    P = InferenceParams(3, 0.0)
    cb.submit(0, [1, 2, 3], P, priority=B.PRIORITY_LOW)
    cb.submit(1, [1, 2, 3], P, priority=B.PRIORITY_LOW)
    cb.submit(2, [1, 2, 3], P, priority=B.PRIORITY_HIGH)
    cb.submit(3, [999999], P, priority=B.PRIORITY_HIGH)       # out of vocabulary: fails alone
    cb.submit(4, [4, 5], P, priority=B.PRIORITY_NORMAL)
    res = cb.run_until_idle()
    assert res[3]["finish"] == B.FINISH_FAILED and res[3]["tokens"] == []
    for i in (0, 1, 2, 4):
        assert len(res[i]["tokens"]) == 3 and res[i]["finish"] == B.FINISH_LENGTH and res[i]["ttft_ms"] > 0
    assert res[0]["tokens"] == res[1]["tokens"] == res[2]["tokens"]
    assert res[2]["ttft_ms"] < res[0]["ttft_ms"]              # High was admitted before the earlier Low requests
    cb.close(); e.close()


def _matches_oracle(mimic, prompt, toks):
    w = mimic.generate([prompt], len(toks), n_logit_steps=len(toks))
    if toks == w["tokens"][0]:
        return True
    k = next(j for j in range(len(toks)) if toks[j] != w["tokens"][0][j])
    margin, scale = float(w["margins"][0, k]), float(np.abs(w["logits"][0, k]).max())
    print(f"[margin report] first difference at step {k}, margin {margin:.2e} of max|logit| {scale:.2f}")
    assert margin <= 5e-3 * scale
    return False


def test_engine_step_mixes_decode_rows_and_prompt_chunks():
    """dis_engine_step: one pass holds a running sequence's decode row and the next chunk of a new prompt.  The new
    prompt is read 24 tokens at a time (7 passes for 150 tokens), the running sequence gets a token in every one of
    them, and both end up with the tokens the oracle produces for each request alone."""
    e = Engine(WorkerConfig(to_engine_shape(SHAPE), max_batch=4, kv_pages=96, page_size=16, max_step_tokens=256))
    e.load_synthetic(31); e.finalize()
    rng = np.random.default_rng(8)
    pa, pb = rng.integers(0, SHAPE.vocab, 9).tolist(), rng.integers(0, SHAPE.vocab, 150).tolist()
    P = InferenceParams(max_tokens=40, temperature=0.0)
    a = e.seq_add(pa, P)
    ta = [int(e.prefill([a])[0][0])]
    for _ in range(3):
        ta.append(int(e.decode([a])[0][0]))
    b = e.seq_add(pb, P)
    tb, passes_until_first = [], 0
    while len(tb) < 10:
        toks, fin, prod = e.step([a, b], max_prefill_tokens=24)
        assert prod[0] == 1                                   # the running sequence never stalls
        ta.append(int(toks[0]))
        if prod[1]:
            tb.append(int(toks[1]))
        elif not tb:
            passes_until_first += 1
    assert passes_until_first == 150 // 24                   # 6 passes read 144 tokens, the 7th finishes the prompt and samples
    lp = e.last_logprobs([a, b])
    assert np.isfinite(lp).all() and (lp <= 0).all()
    # the three entry points mix: plain decode goes on from where the mixed passes left off
    toks, _ = e.decode([a, b])
    ta.append(int(toks[0])); tb.append(int(toks[1]))
    e.seq_remove(a); e.seq_remove(b)
    assert e.status().kv_pages_free == e.status().kv_pages_total
    mimic = O.Oracle(SHAPE, act_round=True); mimic.fill_synthetic(31, True)
    _matches_oracle(mimic, pa, ta)
    _matches_oracle(mimic, pb, tb)
    e.close()


def test_chunked_prefill_in_the_batcher_never_stalls_running_requests():
    """Requirements.md:82 wants tokens to keep flowing: with a prefill budget of 32 tokens per scheduler step, a
    request that is already generating receives one token in EVERY step while a 200-token prompt submitted later is
    being read; the late request still gets exactly the tokens it gets alone."""
    e = Engine(WorkerConfig(to_engine_shape(SHAPE), max_batch=4, kv_pages=128, page_size=16, max_step_tokens=256))
    e.load_synthetic(31); e.finalize()
    cb = ContinuousBatcher(e, max_active=4)
    cb.set_prefill_budget(32)
    rng = np.random.default_rng(12)
    early, late = rng.integers(0, SHAPE.vocab, 12).tolist(), rng.integers(0, SHAPE.vocab, 200).tolist()
    cb.submit("early", early, InferenceParams(30, 0.0))
    got = {"early": [], "late": []}
    for ev in cb.step():
        if ev[1] == B.EVENT_TOKEN:
            got[ev[0]].append(ev[2])
    assert len(got["early"]) == 1                             # a 12-token prompt fits the first pass
    cb.submit("late", late, InferenceParams(5, 0.0))
    steps_without_late = 0
    while not got["late"]:
        evs = cb.step()
        early_tokens = [ev for ev in evs if ev[0] == "early" and ev[1] == B.EVENT_TOKEN]
        assert len(early_tokens) == 1                         # one token per step, prompt chunk or not
        for ev in evs:
            if ev[1] == B.EVENT_TOKEN:
                got[ev[0]].append(ev[2])
        steps_without_late += not got["late"]
    assert steps_without_late == 200 // 32                    # 6 chunks of 32, the 7th pass ends the prompt
    res = cb.run_until_idle()
    for k in got:
        got[k] += res.get(k, {"tokens": []})["tokens"]
    assert len(got["early"]) == 30 and len(got["late"]) == 5
    mimic = O.Oracle(SHAPE, act_round=True); mimic.fill_synthetic(31, True)
    _matches_oracle(mimic, early, got["early"])
    _matches_oracle(mimic, late, got["late"])
    assert e.status().kv_pages_free == e.status().kv_pages_total
    cb.close(); e.close()


def test_short_prompts_are_read_before_a_long_one_that_arrived_first():
    """The order among requests that hold a slot (dis_serving_set_prefill_order): by default the shortest prompt is read
    first, so three 10-token requests submitted AFTER a 230-token one get their first token before it; in admission
    order (mode 0) the long prompt is read first.  Either way every request gets the tokens it gets alone."""
    rng = np.random.default_rng(21)
    long_p = rng.integers(0, SHAPE.vocab, 230).tolist()
    shorts = [rng.integers(0, SHAPE.vocab, 10).tolist() for _ in range(3)]
    mimic = O.Oracle(SHAPE, act_round=True); mimic.fill_synthetic(31, True)
    for shortest_first in (True, False):
        e = Engine(WorkerConfig(to_engine_shape(SHAPE), max_batch=4, kv_pages=128, page_size=16, max_step_tokens=256))
        e.load_synthetic(31); e.finalize()
        cb = ContinuousBatcher(e, max_active=4)
        cb.set_prefill_budget(64)
        cb.set_prefill_order(shortest_first)
        cb.submit("long", long_p, InferenceParams(4, 0.0))
        for i, p in enumerate(shorts):
            cb.submit(f"s{i}", p, InferenceParams(4, 0.0))
        first, toks = [], {}
        for _ in range(200):
            for ev in cb.step():
                if ev[1] == B.EVENT_TOKEN:
                    if ev[3] == 0:
                        first.append(ev[0])
                    toks.setdefault(ev[0], []).append(ev[2])
            d = cb.depth()
            if d.total == 0 and d.active == 0:
                break
        assert sorted(first) == ["long", "s0", "s1", "s2"]
        if shortest_first:
            assert first.index("long") == 3, first
        else:
            assert first.index("long") == 0, first
        _matches_oracle(mimic, long_p, toks["long"])
        for i, p in enumerate(shorts):
            _matches_oracle(mimic, p, toks[f"s{i}"])
        assert e.status().kv_pages_free == e.status().kv_pages_total
        cb.close(); e.close()
