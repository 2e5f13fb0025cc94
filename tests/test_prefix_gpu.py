"""Prefix KV reuse on the device page pool (spec Property 9 / 10,
.kiro/specs/distributed-inference-server/design.md:734-744), through the C ABI.

A request that starts with the tokens of an earlier one must (a) really reuse the pages -- the prefix counters say
how many tokens were not recomputed -- and (b) produce what an engine WITHOUT the cache produces: first-token
logits within the bf16 tolerance of north_star (2e-2 of max|logit|; in practice ~1e-3, the only difference being
the GEMM tile shapes the cached K/V were computed with) and the same greedy ids except at sub-noise margins."""
import numpy as np
import pytest

from distributed_inference_server_b200 import Engine, InferenceParams
from distributed_inference_server_b200 import binding as B
from oracle import oracle as O
from tests.test_engine_gpu import assert_greedy_matches, make_engine, rel_err

pytestmark = pytest.mark.gpu
SHAPE = O.ModelShape(1024, 256, 3, 4, 2, 64, 704, 512)
SEED, PS = 11, 16


def run(e: Engine, prompt, n_new, keep=False):
    slot = e.seq_add(prompt, InferenceParams(max_tokens=n_new + 1, temperature=0.0))
    toks, _ = e.prefill([slot])
    out, lg = [int(toks[0])], e.last_logits(1)[0].copy()
    for _ in range(n_new - 1):
        toks, _ = e.decode([slot])
        out.append(int(toks[0]))
    if not keep:
        e.seq_remove(slot)
    return out, lg, slot


@pytest.fixture()
def engines():
    cached = make_engine(SHAPE, seed=SEED, kv_pages=48, page_size=PS, flags=B.FLAG_PREFIX_CACHE)
    plain = make_engine(SHAPE, seed=SEED, kv_pages=48, page_size=PS)
    yield cached, plain
    cached.close(); plain.close()


def check_same(cached, plain, prompt, n_new, label):
    got, lg, _ = run(cached, prompt, n_new)
    want, lg0, _ = run(plain, prompt, n_new)
    assert rel_err(lg, lg0) <= 2e-2, (label, rel_err(lg, lg0))
    if got != want:  # allowed only where the fp32 oracle itself is within noise of a tie
        ref = O.Oracle(SHAPE, act_round=True); ref.fill_synthetic(SEED, True)
        assert_greedy_matches([got], ref, [prompt], n_new, label, max_flipped=1)
    return got


def test_shared_prefix_is_reused_not_recomputed(engines):
    cached, plain = engines
    rng = np.random.default_rng(0)
    a = rng.integers(0, SHAPE.vocab, 100).tolist()
    check_same(cached, plain, a, 6, "first request")
    s0 = cached.prefix_stats()
    assert s0["hit_tokens"] == 0 and s0["cached_pages"] == (100 + 5) // PS   # full pages of prompt + generated tokens
    b = a[:80] + rng.integers(0, SHAPE.vocab, 30).tolist()
    check_same(cached, plain, b, 6, "shared 80-token prefix")
    s1 = cached.prefix_stats()
    assert s1["hit_tokens"] - s0["hit_tokens"] == 80 and s1["hits"] == 1        # Property 9: five pages mapped, not computed
    # the identical prompt: everything but the last page comes from the cache (one token at least is computed)
    check_same(cached, plain, a, 6, "identical prompt")
    s2 = cached.prefix_stats()
    assert s2["hit_tokens"] - s1["hit_tokens"] == (100 - 1) // PS * PS
    # a follow-up turn quoting prompt + answer: the generated tokens' pages are reusable too
    ans, _, _ = run(plain, a, 30)
    run(cached, a, 30)
    turn2 = a + ans[:29] + rng.integers(0, SHAPE.vocab, 9).tolist()
    before = cached.prefix_stats()["hit_tokens"]
    got = check_same(cached, plain, turn2, 4, "second turn")
    assert cached.prefix_stats()["hit_tokens"] - before >= 100 // PS * PS     # at least the whole prompt
    assert len(got) == 4


def test_live_sequences_share_pages_and_decode_together(engines):
    cached, plain = engines
    rng = np.random.default_rng(1)
    common = rng.integers(0, SHAPE.vocab, 64).tolist()
    p1, p2 = common + [5, 6, 7], common + [9, 10, 11, 12, 13]
    P = InferenceParams(max_tokens=9, temperature=0.0)

    def both(e):
        s1 = e.seq_add(p1, P); t1, _ = e.prefill([s1])      # p1 is prefilled (and registered) while it stays live
        s2 = e.seq_add(p2, P); t2, _ = e.prefill([s2])
        out = [[int(t1[0])], [int(t2[0])]]
        for _ in range(7):
            t, _ = e.decode([s1, s2])
            out[0].append(int(t[0])); out[1].append(int(t[1]))
        st = e.status()
        e.seq_remove(s1); e.seq_remove(s2)
        return out, st

    got, st_c = both(cached)
    want, st_p = both(plain)
    assert cached.prefix_stats()["hit_tokens"] == 64
    assert st_p.kv_pages_free - st_c.kv_pages_free == -(64 // PS)   # four pages are held once instead of twice
    if got != want:
        ref = O.Oracle(SHAPE, act_round=True); ref.fill_synthetic(SEED, True)
        assert_greedy_matches(got, ref, [p1, p2], 8, "live sharing")


def test_lru_eviction_under_pressure_keeps_results_right():
    cached = make_engine(SHAPE, seed=SEED, kv_pages=12, page_size=PS, flags=B.FLAG_PREFIX_CACHE, max_batch=2)
    plain = make_engine(SHAPE, seed=SEED, kv_pages=12, page_size=PS, max_batch=2)
    rng = np.random.default_rng(2)
    prompts = [rng.integers(0, SHAPE.vocab, 70).tolist() for _ in range(6)]   # 5 pages each, pool of 12
    order = [0, 1, 0, 2, 3, 0, 4, 5, 1, 0]
    for k, i in enumerate(order):
        check_same(cached, plain, prompts[i], 3, f"request {k} (prompt {i})")
    s = cached.prefix_stats()
    assert s["evictions"] > 0 and s["hits"] >= 2                              # Property 10 at work, reuse still happening
    assert s["free_pages"] + s["evictable_pages"] == 12                      # nothing leaked
    assert cached.status().kv_pages_free == 12
    # a request that needs the whole pool is admitted by evicting everything that is not mapped
    big = rng.integers(0, SHAPE.vocab, 12 * PS - 8).tolist()
    check_same(cached, plain, big, 2, "pool-sized request")
    cached.close(); plain.close()
