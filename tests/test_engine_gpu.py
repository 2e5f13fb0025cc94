"""End-to-end parity of the CUDA engine, through the C ABI, against the oracle and the committed
transformers golden vectors.  Tolerances are the ones north_star states: logits max relative error
<= 2e-2 against the fp32 reference; greedy token ids identical (checked against the oracle run with
the GPU path's bf16 rounding points, and against the fp32 reference wherever its top-1/top-2 margin
is above bf16 noise -- any step where it is not is printed, not hidden)."""
import glob
import os

import numpy as np
import pytest

from distributed_inference_server_b200 import (BatchedRequest, Engine, FinishReason, InferenceBatch, InferenceParams,
                                               InferenceWorker, ModelShape, WorkerConfig, WorkerError)
from distributed_inference_server_b200 import binding as B
from oracle import oracle as O
from tests._gpu_util import shape_from_npz, to_engine_shape

pytestmark = pytest.mark.gpu
GOLDEN_DIR = os.path.join(os.path.dirname(__file__), "golden")
GPU_GOLDEN = ["mini_llama3", "mini_d128", "mini_g8", "mini_llama3_64"]  # head_dim 64/128 fixtures; the last one runs 64 greedy tokens
LOGIT_TOL = 2e-2


def make_engine(shape: O.ModelShape, oracle: O.Oracle = None, seed=None, **kw):
    cfg = WorkerConfig(to_engine_shape(shape), max_batch=kw.pop("max_batch", 8), kv_pages=kw.pop("kv_pages", 64),
                       page_size=kw.pop("page_size", 16), max_step_tokens=kw.pop("max_step_tokens", 256), **kw)
    e = Engine(cfg)
    if oracle is not None:
        for kind, layer, arr in oracle.tensors():
            e.load_tensor(kind, layer, arr)
    else:
        e.load_synthetic(seed)
    e.finalize()
    return e


def greedy(e: Engine, prompts, n_new, want_logits=False):
    P = InferenceParams(max_tokens=n_new + 1, temperature=0.0)
    slots = [e.seq_add(p, P) for p in prompts]
    toks, fin = e.prefill(slots)
    out = [[int(t)] for t in toks]
    logits = [e.last_logits(len(slots))] if want_logits else None
    for _ in range(n_new - 1):
        toks, fin = e.decode(slots)
        for i, t in enumerate(toks):
            out[i].append(int(t))
        if want_logits:
            logits.append(e.last_logits(len(slots)))
    for s in slots:
        e.seq_remove(s)
    return out, logits


def rel_err(a, b):
    return float(np.abs(a - b).max() / np.abs(b).max())


def assert_greedy_matches(got, oracle, prompts, n_new, label, max_flipped=None):
    """Greedy ids equal the oracle's wherever the oracle's own top-1/top-2 logit margin is above the
    noise left by a different fp32 summation order (and bf16 P in attention) once it has flipped a bf16
    rounding somewhere upstream: 5e-3 of max|logit|.  Random-init logits are nearly flat (SURVEY.md §7
    hard part 2); a sequence that diverges at a sub-noise margin is reported, not hidden."""
    want = oracle.generate(prompts, n_new, n_logit_steps=n_new)
    flipped = 0
    for b in range(len(prompts)):
        if got[b] == want["tokens"][b]:
            continue
        i = next(k for k in range(n_new) if got[b][k] != want["tokens"][b][k])
        margin = float(want["margins"][b, i])
        scale = float(np.abs(want["logits"][b, i]).max())
        flipped += 1
        print(f"[margin report] {label} seq {b}: first difference at step {i}/{n_new}, oracle top1-top2 margin "
              f"{margin:.3e} = {margin / scale:.1e} of max|logit| {scale:.2f}")
        assert margin <= 5e-3 * scale, (label, b, i, got[b], want["tokens"][b], margin, scale)
    limit = max(2, len(prompts) // 2) if max_flipped is None else max_flipped  # tiny models: near-flat logits
    assert flipped <= limit, (label, flipped)
    return want


@pytest.mark.parametrize("flags", [0, B.FLAG_NO_GRAPH | B.FLAG_NO_PDL], ids=["graph", "eager"])
def test_64_greedy_tokens_identical_to_transformers(flags):
    """north_star: greedy ids bit-exact for the first 64 tokens.  On mini_llama3_64 every step's fp32 top-1/top-2 margin
    is above bf16 noise (oracle/make_golden.py picked the seeds for that), so nothing may differ: the 64 ids equal the
    transformers continuation stored in the fixture, token for token."""
    z = np.load(os.path.join(GOLDEN_DIR, "mini_llama3_64.npz"))
    shape = shape_from_npz(z)
    fp32 = O.Oracle(shape, act_round=False); fp32.fill_synthetic(int(z["weight_seed"]), True)
    e = make_engine(shape, fp32, flags=flags)
    got, _ = greedy(e, [z["prompt"].tolist()], 64)
    e.close()
    assert len(z["greedy"]) == 64 and got[0] == z["greedy"].tolist()


@pytest.mark.parametrize("name", GPU_GOLDEN)
@pytest.mark.parametrize("flags", [0, B.FLAG_NO_GRAPH, B.FLAG_FORCE_GEMV | B.FLAG_NO_GRAPH, B.FLAG_NO_PDL],
                         ids=["graph", "eager", "gemv", "graph-nopdl"])
def test_golden_logits_and_greedy(name, flags):
    z = np.load(os.path.join(GOLDEN_DIR, name + ".npz"))
    shape = shape_from_npz(z)
    fp32 = O.Oracle(shape, act_round=False); fp32.fill_synthetic(int(z["weight_seed"]), True)
    mimic = O.Oracle(shape, act_round=True); mimic.fill_synthetic(int(z["weight_seed"]), True)
    e = make_engine(shape, fp32, flags=flags)
    prompt = z["prompt"].tolist()
    n_new = len(z["greedy"])
    got, logits = greedy(e, [prompt], n_new, want_logits=True)
    # (1) first-token logits against the transformers golden vector
    assert rel_err(logits[0][0], z["logits"][-1]) <= LOGIT_TOL
    # (2) every step's logits against the fp32 oracle run on the GPU's own token history
    sid = fp32.seq_new()
    hist = prompt + got[0]
    ref = fp32.forward(hist[:-1], [sid] * (len(hist) - 1), np.arange(len(hist) - 1))
    for i in range(n_new):
        assert rel_err(logits[i][0], ref[len(prompt) - 1 + i]) <= LOGIT_TOL, i
    # (3) greedy ids: identical to the oracle with the GPU's rounding points ...
    assert_greedy_matches(got, mimic, [prompt], n_new, name)
    # ... and to the fp32 reference / transformers wherever the margin is above bf16 noise
    ref_gen = fp32.generate([prompt], n_new)
    for i, (a, b) in enumerate(zip(got[0], ref_gen["tokens"][0])):
        if a != b:
            m = float(ref_gen["margins"][0, i])
            print(f"[margin report] {name}: step {i} gpu={a} fp32={b} top1-top2={m:.3e}")
            assert m < LOGIT_TOL * np.abs(ref[len(prompt) - 1 + i]).max()
            break
    e.close()


def test_synthetic_device_weights_match_oracle_generator():
    shape = O.ModelShape(1024, 256, 2, 4, 1, 128, 512, 256)
    e = make_engine(shape, seed=1234)
    for kind, layer in [(O.EMBED, 0), (O.LM_HEAD, 0), (O.WQ, 1), (O.WK, 0), (O.WV, 1), (O.WO, 1), (O.WGATE, 0),
                        (O.WUP, 1), (O.WDOWN, 0), (O.ATTN_NORM, 1), (O.MLP_NORM, 0), (O.FINAL_NORM, 0)]:
        got = O.from_bf16_bits(e.read_tensor(kind, layer))
        ref = O.synth_tensor(shape, 1234, kind, layer).reshape(-1)
        assert np.array_equal(got, ref), (kind, layer)
    e.close()


@pytest.mark.parametrize("flags", [0, B.FLAG_NO_GRAPH], ids=["graph", "eager"])
def test_ragged_batch_greedy_matches_oracle(flags):
    shape = O.ModelShape(2048, 512, 3, 8, 2, 64, 1408, 512)
    mimic = O.Oracle(shape, act_round=True); mimic.fill_synthetic(77, True)
    e = make_engine(shape, seed=77, flags=flags, max_batch=8, kv_pages=128)
    rng = np.random.default_rng(9)
    prompts = [rng.integers(0, shape.vocab, n).tolist() for n in (1, 5, 16, 17, 63, 100)]
    n_new = 24
    got, _ = greedy(e, prompts, n_new)
    assert_greedy_matches(got, mimic, prompts, n_new, "ragged")
    # the cache is reusable: a second round in the same engine, in another batch order (stream-K split points
    # depend on the tile a token lands in, so fp32 summation order -- not the result's validity -- may change)
    got2, _ = greedy(e, prompts[::-1], n_new)
    assert_greedy_matches(got2, mimic, prompts[::-1], n_new, "ragged-reversed")
    e.close()


def test_prefill_in_chunks_equals_one_pass():
    shape = O.ModelShape(1024, 256, 2, 4, 1, 128, 512, 512)
    rng = np.random.default_rng(4)
    prompts = [rng.integers(0, shape.vocab, n).tolist() for n in (70, 33, 150)]
    a = make_engine(shape, seed=5, max_step_tokens=512, kv_pages=128)
    b = make_engine(shape, seed=5, max_step_tokens=64, kv_pages=128)  # forces 4 chunks, sequences split mid-way
    ga, _ = greedy(a, prompts, 8)
    gb, _ = greedy(b, prompts, 8)
    assert ga == gb
    a.close(); b.close()


def test_llama3_8b_width_two_layers():
    """BASELINE configs[1] geometry (H=4096, 32/8 heads, d=128, I=14336, V=128256) at reduced depth so
    the CPU oracle finishes in seconds: batch-32 decode through the tcgen05 path."""
    shape = O.ModelShape(128256, 4096, 2, 32, 8, 128, 14336, 256)
    fp32 = O.Oracle(shape, act_round=False); fp32.fill_synthetic(3, True)
    mimic = O.Oracle(shape, act_round=True)
    for (k, l, dst), (_, _, src) in zip(mimic.tensors(), fp32.tensors()):
        dst[...] = src
    e = make_engine(shape, seed=3, max_batch=32, kv_pages=32 * 4, page_size=64, max_step_tokens=512)
    rng = np.random.default_rng(11)
    prompts = [rng.integers(0, shape.vocab, 12).tolist() for _ in range(32)]
    n_new = 6
    got, logits = greedy(e, prompts, n_new, want_logits=True)
    assert_greedy_matches(got, mimic, prompts, n_new, "8b-width", max_flipped=4)
    # logits of every step against the fp32 reference on the same history
    ref = fp32.generate(prompts, n_new, n_logit_steps=n_new)
    for b in range(32):
        if ref["tokens"][b] == got[b]:
            for i in range(n_new):
                assert rel_err(logits[i][b], ref["logits"][b, i]) <= LOGIT_TOL, (b, i)
    e.close()


def test_llama3_8b_full_size_properties():
    """BASELINE configs[1] at FULL size (32 layers, random-init): no CPU oracle finishes this in seconds, so the check is
    through properties that do not depend on size -- two copies of a prompt in one batch give bit-identical logits and
    tokens (row invariance of every kernel), a second run of the same batch reproduces the first bit for bit
    (determinism: no atomics, fixed summation orders), a prompt alone and inside a batch agree within the logit
    tolerance, and everything is finite and inside the vocabulary."""
    shape = O.LLAMA3_8B
    e = make_engine(shape, seed=0, max_batch=8, kv_pages=8 * 10, page_size=64, max_step_tokens=2048)
    rng = np.random.default_rng(5)
    p0, p2 = rng.integers(0, shape.vocab, 512).tolist(), rng.integers(0, shape.vocab, 300).tolist()
    batch = [p0, list(p0), p2, list(p2), rng.integers(0, shape.vocab, 17).tolist()]
    runs = []
    for _ in range(2):
        got, logits = greedy(e, batch, 16, want_logits=True)
        runs.append((got, logits))
    got, logits = runs[0]
    assert got[0] == got[1] and got[2] == got[3]
    for step in logits:
        assert np.isfinite(step).all()
        assert np.array_equal(step[0], step[1]) and np.array_equal(step[2], step[3])
    assert all(0 <= t < shape.vocab for seq in got for t in seq)
    assert runs[1][0] == got and all(np.array_equal(a, b) for a, b in zip(runs[1][1], logits))
    alone, lg = greedy(e, [p0], 2, want_logits=True)
    assert rel_err(lg[0][0], logits[0][0]) <= LOGIT_TOL
    e.close()


# ---- the worker interface: structural contracts of the spec -----------------------

def small_worker(**kw):
    shape = ModelShape(1024, 256, 2, 4, 1, 128, 512, 128)
    w = InferenceWorker(WorkerConfig(shape, max_batch=kw.pop("max_batch", 8), kv_pages=kw.pop("kv_pages", 64), page_size=16,
                                     max_step_tokens=256, synthetic_seed=8, **kw))
    w.initialize()
    return w


def test_infer_n_in_n_out_and_usage():
    w = small_worker()
    rng = np.random.default_rng(0)
    lens = [3, 9, 1, 20, 7]
    stride = max(lens)
    ids = [rng.integers(1, 1000, n).tolist() + [0] * (stride - n) for n in lens]
    mask = [[1] * n + [0] * (stride - n) for n in lens]
    reqs = [BatchedRequest(params=InferenceParams(max_tokens=4 + i, temperature=0.0)) for i in range(len(lens))]
    events = []
    res = w.infer(InferenceBatch(ids, mask, reqs), on_token=lambda i, t, k, f: events.append((i, t, k, f)) or True)
    assert len(res.results) == len(lens)                                   # Property 21 (design.md:806-810)
    for i, r in enumerate(res.results):
        assert r.request_id == reqs[i].id
        assert r.prompt_tokens == lens[i] and r.completion_tokens == 4 + i == len(r.tokens)
        assert r.finish_reason == FinishReason.Length and not r.failed
        assert [t for (j, t, _, _) in events if j == i] == r.tokens         # one TokenEvent per token, in order
        assert [k for (j, _, k, _) in events if j == i] == list(range(4 + i))
    assert res.tokens_generated == sum(4 + i for i in range(len(lens)))
    assert res.inference_time > 0 and res.time_to_first_token > 0
    st = w.status()
    assert st.is_healthy and st.active_batches == 0 and st.kv_pages_free == st.kv_pages_total
    assert st.kernel_launches > 0
    w.shutdown()


def test_finish_reasons_and_stop_sequences():
    w = small_worker()
    prompt = [5, 6, 7, 8]
    base = w.infer(InferenceBatch([prompt], requests=[BatchedRequest(params=InferenceParams(12, 0.0))])).results[0].tokens
    assert len(base) == 12
    # a stop sequence made of generated tokens 3..4 ends generation right after them
    stop = [base[3:5]]
    r = w.infer(InferenceBatch([prompt], requests=[BatchedRequest(params=InferenceParams(12, 0.0, 1.0, stop))])).results[0]
    first = next(i for i in range(1, len(base)) if base[i - 1:i + 1] == stop[0])
    assert r.finish_reason == FinishReason.StopSequence and r.tokens == base[:first + 1]
    w.shutdown()
    # EOS -> FinishReason::Stop
    w2 = small_worker(eos_token=base[2])
    r2 = w2.infer(InferenceBatch([prompt], requests=[BatchedRequest(params=InferenceParams(12, 0.0))])).results[0]
    k = base.index(base[2])
    assert r2.finish_reason == FinishReason.Stop and r2.tokens == base[:k + 1]
    w2.shutdown()


def test_failure_isolation_and_client_drop():
    w = small_worker()
    good = [1, 2, 3]
    too_long = list(range(1, 200))                       # does not fit max_ctx = 128
    out_of_vocab = [5, 999999]
    res = w.infer(InferenceBatch([good, too_long, out_of_vocab, good], requests=[
        BatchedRequest(params=InferenceParams(5, 0.0)) for _ in range(4)]))
    assert [r.failed for r in res.results] == [False, True, True, False]   # Property 22 (design.md:812-816)
    assert res.results[0].tokens == res.results[3].tokens and len(res.results[0].tokens) == 5
    assert res.results[1].tokens == [] and res.results[2].tokens == []
    # the host drops request 0 after two tokens (client disconnect, requirements.md:85)
    seen = []
    res = w.infer(InferenceBatch([good, good], requests=[BatchedRequest(params=InferenceParams(6, 0.0)) for _ in range(2)]),
                  on_token=lambda i, t, k, f: (seen.append((i, k)) or True) and not (i == 0 and k == 1))
    assert len(res.results[0].tokens) == 2 and len(res.results[1].tokens) == 6
    assert w.status().kv_pages_free == w.status().kv_pages_total
    w.shutdown()


def test_context_limit_and_page_exhaustion():
    w = small_worker(kv_pages=3)                         # 3 pages * 16 tokens
    r = w.infer(InferenceBatch([[1] * 10, [2] * 10], requests=[BatchedRequest(params=InferenceParams(60, 0.0)) for _ in range(2)]))
    # both fit at first; one of them runs out of pages later and fails alone, the other keeps its tokens
    assert sum(len(x.tokens) for x in r.results) > 0
    assert any(x.failed for x in r.results) or all(len(x.tokens) < 60 for x in r.results)
    w.shutdown()
    w = small_worker()
    r = w.infer(InferenceBatch([[3] * 100], requests=[BatchedRequest(params=InferenceParams(4096, 0.0))])).results[0]
    assert r.finish_reason == FinishReason.Length and r.prompt_tokens + r.completion_tokens <= 128
    w.shutdown()


def test_kv_pool_sized_from_free_memory():
    """kv_pages = 0: the pool takes kv_mem_fraction of the free device memory, but never more pages than max_batch
    sequences of max_ctx tokens can use."""
    w = small_worker(kv_pages=0, kv_mem_fraction=0.01, max_batch=4)         # context 128 = 8 pages of 16 per sequence
    st = w.status()
    assert st.kv_pages_total == 4 * 8 and st.kv_pages_free == st.kv_pages_total
    r = w.infer(InferenceBatch([[5] * 40, [6] * 3], requests=[BatchedRequest(params=InferenceParams(6, 0.0)) for _ in range(2)]))
    assert [len(x.tokens) for x in r.results] == [6, 6]
    assert w.status().kv_pages_free == st.kv_pages_total                      # everything handed back
    w.shutdown()


def test_replay_last_step_changes_nothing():
    """dis_engine_replay_last_step is a measurement aid: replaying the last decode graph must leave the sequence
    exactly where it was (same next tokens as an engine that never replayed)."""
    shape = ModelShape(1024, 256, 2, 4, 1, 128, 512, 128)
    outs = []
    for replay in (False, True):
        e = Engine(WorkerConfig(shape, max_batch=4, kv_pages=32, page_size=16, max_step_tokens=64))
        e.load_synthetic(3); e.finalize()
        slots = [e.seq_add(p, InferenceParams(max_tokens=12, temperature=0.0)) for p in ([1, 2, 3, 4, 5], [9] * 20)]
        toks = [list(map(int, e.prefill(slots)[0]))]
        for i in range(8):
            toks.append(list(map(int, e.decode(slots)[0])))
            if replay and i % 3 == 1:
                assert e.replay_last_step(5) > 0.0
        outs.append(toks)
        e.close()
    assert outs[0] == outs[1]


def test_error_codes():
    shape = ModelShape(1024, 256, 2, 4, 1, 128, 512, 128)
    e = Engine(WorkerConfig(shape, max_batch=2, kv_pages=8, page_size=16, max_step_tokens=64))
    with pytest.raises(WorkerError) as ei:
        e.seq_add([1, 2], InferenceParams(4, 0.0))       # before finalize
    assert ei.value.kind == WorkerError.ModelNotLoaded
    e.load_synthetic(1); e.finalize()
    s = e.seq_add([1, 2], InferenceParams(4, 0.0))
    with pytest.raises(WorkerError):
        e.decode([s])                                    # not prefilled yet
    e.close()
    w = InferenceWorker(WorkerConfig(shape, max_batch=2, kv_pages=8, page_size=16))
    with pytest.raises(WorkerError) as ei:
        w.infer(InferenceBatch([[1]]))
    assert ei.value.kind == WorkerError.ModelNotLoaded
    w.initialize(); w.shutdown()
    with pytest.raises(WorkerError) as ei:
        w.infer(InferenceBatch([[1]]))
    assert ei.value.kind == WorkerError.Shutdown
    with pytest.raises(WorkerError):
        Engine(WorkerConfig(ModelShape(1024, 256, 2, 4, 1, 96, 512, 128)))   # unsupported head_dim


def test_sampling_reproducible_and_close_to_oracle():
    shape = O.ModelShape(1024, 256, 2, 4, 1, 128, 512, 128)
    e = make_engine(shape, seed=8)
    P = InferenceParams(max_tokens=20, temperature=0.8, top_p=0.9, seed=4242)
    def run():
        s = e.seq_add([9, 8, 7], P)
        toks = [int(e.prefill([s])[0][0])]
        for _ in range(11):
            toks.append(int(e.decode([s])[0][0]))
        e.seq_remove(s)
        return toks
    a, b = run(), run()
    assert a == b                                        # counter-based RNG: same seed, same tokens
    mimic = O.Oracle(shape, act_round=True); mimic.fill_synthetic(8, True)
    want = mimic.generate([[9, 8, 7]], 12, temperature=0.8, top_p=0.9, seeds=[4242])["tokens"][0]
    same = 0
    for x, y in zip(a, want):
        if x != y:
            break
        same += 1
    print(f"[sampling report] identical prefix {same}/12")
    assert same >= 3  # one borderline draw changes the history; the step-by-step check below does not depend on it
    # step by step on the ENGINE's own history: at every step the oracle's sampler, given the engine's logits of that
    # step and the same (seed, step) counter, must draw the engine's token -- this pins the whole plumbing (temperature,
    # top_p, seed and step index reaching the kernel, the counter-based uniform) end to end, not only a prefix
    s = e.seq_add([9, 8, 7], P)
    toks, lg = [int(e.prefill([s])[0][0])], [e.last_logits(1)[0].copy()]
    lps = [float(e.last_logprobs([s])[0])]
    for _ in range(11):
        toks.append(int(e.decode([s])[0][0])); lg.append(e.last_logits(1)[0].copy()); lps.append(float(e.last_logprobs([s])[0]))
    e.seq_remove(s)
    assert toks == a
    off = 0
    for step, (t, z) in enumerate(zip(toks, lg)):
        ref, tau, mass = O.sample(z, 0.8, 0.9, 4242, step)
        if ref != t:                                     # only a draw within fp32 noise of a bucket edge may differ
            off += 1
            p = np.exp(z / 0.8 - (z / 0.8).max()); p /= p.sum()
            assert p[t] >= tau * (1 - 1e-5), (step, t, ref)
        zz = z.astype(np.float64) / 0.8
        want_lp = zz[t] - (np.log(np.exp(zz - zz.max()).sum()) + zz.max())
        assert abs(lps[step] - want_lp) <= 1e-3 * max(1.0, abs(want_lp)), (step, lps[step], want_lp)
    assert off <= 1, off
    e.close()


def test_decode_batch_padded_to_graph_bucket_changes_nothing():
    """Above 8 rows a decode step is padded to the next multiple of 8 with copies of its last row, so that the CUDA-graph
    cache holds a few sizes instead of one per batch size.  13 running sequences (padded to 16) must produce exactly the
    tokens and logits of the eager engine, which does not pad, and the oracle's tokens; afterwards 5 of them are dropped
    (8 rows: no padding) and the rest go on unchanged."""
    shape = O.ModelShape(2048, 512, 3, 8, 2, 64, 1408, 512)
    rng = np.random.default_rng(21)
    prompts = [rng.integers(0, shape.vocab, int(n)).tolist() for n in rng.integers(3, 70, 13)]
    outs = []
    for flags in (0, B.FLAG_NO_GRAPH):
        e = make_engine(shape, seed=77, flags=flags, max_batch=16, kv_pages=160)
        P = InferenceParams(max_tokens=40, temperature=0.0)
        slots = [e.seq_add(p, P) for p in prompts]
        toks = [list(map(int, e.prefill(slots)[0]))]
        logits = []
        for _ in range(6):
            toks.append(list(map(int, e.decode(slots)[0])))
            logits.append(e.last_logits(len(slots)))
        for s in slots[8:]:
            e.seq_remove(s)
        for _ in range(4):
            toks.append(list(map(int, e.decode(slots[:8])[0])) + [None] * 5)
        outs.append((toks, logits))
        e.close()
    assert outs[0][0] == outs[1][0]
    assert all(np.array_equal(a, b) for a, b in zip(outs[0][1], outs[1][1]))
    mimic = O.Oracle(shape, act_round=True); mimic.fill_synthetic(77, True)
    got = [[step[b] for step in outs[0][0][:7]] for b in range(13)]
    assert_greedy_matches(got, mimic, prompts, 7, "bucket-padded")
