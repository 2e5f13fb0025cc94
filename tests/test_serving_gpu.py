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
