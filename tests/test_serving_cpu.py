"""Admission queue of the serving layer (dis_serving_*, queue-only mode: no GPU needed) against the behaviour
the reference's PriorityQueueManager has (crates/core/src/queue.rs) and its spec properties 6-8
(.kiro/specs/distributed-inference-server/design.md:716-732).  The reference ships no tests for its queue."""
import pytest
from hypothesis import given, settings, strategies as st

from distributed_inference_server_b200 import ContinuousBatcher, InferenceParams, QueueFull
from distributed_inference_server_b200 import binding as B

P = InferenceParams(4, 0.0)


def mk(**kw):
    return ContinuousBatcher(None, **kw)


@settings(max_examples=100, deadline=None)
@given(st.lists(st.integers(0, 2), min_size=0, max_size=60), st.integers(1, 70))
def test_property6_priority_then_fifo(prios, take):
    q = mk()
    for i, p in enumerate(prios):
        q.submit(i, [1], P, priority=p, now=0.0)
    got = q.debug_pop(take)
    want = [i for p in (2, 1, 0) for i, x in enumerate(prios) if x == p][:take]
    assert got == want
    assert q.depth().total == len(prios) - len(got)


@settings(max_examples=100, deadline=None)
@given(st.lists(st.one_of(st.just("push"), st.integers(1, 6)), min_size=1, max_size=120),
       st.integers(2, 12), st.integers(0, 6))
def test_property7_backpressure_hysteresis(ops, high, low_gap):
    low = max(0, high - low_gap)
    q = mk(high_watermark=high, low_watermark=low, max_queue_size=10 * high)
    accepting, depth, nid = True, 0, 0
    for op in ops:
        if op == "push":
            nid += 1
            try:
                q.submit(nid, [1], P, now=0.0)
                assert accepting                       # a refused state never accepts
                depth += 1
            except QueueFull:
                assert not accepting
        else:
            depth -= len(q.debug_pop(op))
        # model of queue.rs:235-249
        if accepting and depth > high:
            accepting = False
        elif not accepting and depth < low:
            accepting = True
        d = q.depth()
        assert d.total == depth and bool(d.accepting) == accepting
        if depth > high:
            assert not d.accepting
        if depth < low:
            assert d.accepting


def test_capacity_limit_and_defaults():
    q = mk(high_watermark=100, low_watermark=50, max_queue_size=5)
    for i in range(5):
        q.submit(i, [1], P, now=0.0)
    with pytest.raises(QueueFull):
        q.submit(5, [1], P, now=0.0)                   # queue.rs:110-113
    d = ContinuousBatcher(None)                         # defaults: queue.rs:24-33
    for i in range(1001):
        d.submit(i, [1], P, now=0.0)
    assert not d.depth().accepting                      # 1001 > high watermark 1000
    d.debug_pop(501)
    assert not d.depth().accepting                      # 500 is not < 500
    d.debug_pop(1)
    assert d.depth().accepting
    with pytest.raises(ValueError):
        mk(high_watermark=1, low_watermark=2)


@settings(max_examples=50, deadline=None)
@given(st.lists(st.tuples(st.floats(0, 100), st.integers(0, 2)), min_size=1, max_size=40), st.floats(1, 50), st.floats(0, 200))
def test_property8_timeout(arrivals, timeout, now):
    q = mk(request_timeout_s=timeout)
    for i, (t, p) in enumerate(arrivals):
        q.submit(i, [1, 2], P, priority=p, now=t)
    events = q.step(now=now)                            # queue-only object: expiry only
    expired = sorted(e[0] for e in events if e[1] == B.EVENT_EXPIRED)
    want = sorted(i for i, (t, _) in enumerate(arrivals) if now - t > timeout)
    assert expired == want
    assert q.depth().total == len(arrivals) - len(want)
    for e in events:
        assert e[4] == B.FINISH_FAILED and e[5] == 2 and e[6] >= timeout * 1e3 - 1e-6


def test_request_ids_are_16_byte_uuids_and_cancel_removes_a_queued_request():
    """RequestID is a Uuid (crates/core/src/types.rs:7): ids cross the boundary as 16 bytes and come back unchanged;
    a client that disconnects while queued (requirements.md:85) leaves the queue and gets one DONE/Stop event."""
    import uuid
    q = mk()
    ids = [uuid.UUID(int=(0xfeedface << 96) | (k << 64) | 0x1234) for k in range(3)]   # differ only beyond the low 64 bits
    for rid in ids:
        q.submit(rid, [1, 2, 3], P, now=0.0)
    assert q.depth().total == 3
    assert q.cancel(ids[1]) and not q.cancel(uuid.uuid4())
    assert q.depth().total == 2
    ev = q.step(now=0.0)
    assert [(e[0], e[1], e[4], e[5]) for e in ev] == [(ids[1], B.EVENT_DONE, B.FINISH_STOP, 3)]
    assert q.debug_pop(5) == [ids[0], ids[2]]


@settings(max_examples=200, deadline=None)
@given(st.lists(st.integers(-3, 9000), min_size=0, max_size=80), st.integers(1, 9000), st.integers(1, 9000))
def test_mixed_step_plan_is_fair_and_bounded(remaining, budget, cap):
    """dis_serving_plan_step, the rule by which dis_engine_step shares one forward pass (chunked prefill): every running
    sequence gets its one token whatever the prompts need; prompt tokens never exceed the budget nor the room the pass
    has left; they are handed out in list (= admission) order, so a later prompt reads nothing unless every earlier one
    was served in full; and no room is wasted while a prompt still wants tokens."""
    import ctypes as C
    import numpy as np
    lib = B.load()
    n = len(remaining)
    rem = np.array(remaining, np.int32)
    take = np.full(n, -7, np.int32)
    assert lib.dis_serving_plan_step(n, rem.ctypes.data, budget, cap, take.ctypes.data) == B.OK
    dec = rem <= 0
    assert (take[dec] == 1).all()
    pre = take[~dec]
    assert (pre >= 0).all() and (pre <= rem[~dec]).all()
    room = max(0, min(budget, cap - int(dec.sum())))
    assert int(pre.sum()) <= room
    want = rem[~dec]
    short = np.nonzero(pre < want)[0]
    if len(short):
        assert (pre[short[0] + 1:] == 0).all()          # nobody behind a partly served prompt reads anything
        assert int(pre.sum()) == room                   # and the room was used up


@settings(max_examples=200, deadline=None)
@given(st.lists(st.tuples(st.integers(0, 2), st.integers(1, 8192), st.floats(0.0, 30.0)), min_size=0, max_size=60),
       st.floats(0.0, 10000.0))
def test_prompt_reading_order_is_shortest_first_within_priority_with_aging(reqs, aging):
    """dis_serving_order_prompts, the order in which the batcher reads the prompts of requests that already hold a slot:
    a permutation; a higher priority class is never behind a lower one; inside a class the effective length
    prompt_tokens - aging * waited never decreases; equal keys keep arrival (= list) order."""
    import numpy as np
    lib = B.load()
    n = len(reqs)
    pr = np.array([r[0] for r in reqs], np.int32)
    pt = np.array([r[1] for r in reqs], np.int32)
    wt = np.array([r[2] for r in reqs], np.float64)
    order = np.full(n, -1, np.int32)
    assert lib.dis_serving_order_prompts(n, pr.ctypes.data, pt.ctypes.data, wt.ctypes.data, float(aging), order.ctypes.data) == B.OK
    assert sorted(order.tolist()) == list(range(n))
    key = pt.astype(np.float64) - aging * wt
    for a, b in zip(order[:-1], order[1:]):
        assert pr[a] >= pr[b]
        if pr[a] == pr[b]:
            assert key[a] <= key[b]
            if key[a] == key[b]:
                assert a < b


def test_prompt_reading_order_examples():
    import numpy as np
    lib = B.load()

    def order(pr, pt, wt, aging):
        pr, pt, wt = np.array(pr, np.int32), np.array(pt, np.int32), np.array(wt, np.float64)
        out = np.zeros(len(pr), np.int32)
        assert lib.dis_serving_order_prompts(len(pr), pr.ctypes.data, pt.ctypes.data, wt.ctypes.data, aging, out.ctypes.data) == B.OK
        return out.tolist()
    N, H, L = B.PRIORITY_NORMAL, B.PRIORITY_HIGH, B.PRIORITY_LOW
    assert order([N, N, N], [8000, 128, 2000], [0, 0, 0], 4096.0) == [1, 2, 0]          # shortest first
    assert order([N, H, L], [100, 8000, 10], [0, 0, 0], 4096.0) == [1, 0, 2]            # priority before length
    assert order([N, N], [8000, 128], [2.0, 0.0], 4096.0) == [0, 1]                     # 2 s of waiting outweigh 7872 tokens
    assert order([N, N], [8000, 128], [1.0, 0.0], 4096.0) == [1, 0]                     # 1 s does not
    assert order([N, N, N], [500, 500, 500], [0, 0, 0], 0.0) == [0, 1, 2]               # ties: arrival order
    assert lib.dis_serving_order_prompts(2, None, None, None, 1.0, None) == B.ERR_INVALID_ARGUMENT
