"""Tensor parallel engine on 2, 4 and 8 GPUs (each case skipped when the box has fewer): one process per GPU,
the exchange after the O and down projections over peer memory (k_tp.cu) or NCCL, vocab-parallel lm_head.
Checked against the oracle and against the product's shard plan (each rank's weights are exactly the plan's
slice of the synthetic tensors).  World sizes above 2 exercise what a 2-rank run cannot: owner-block indexing
Hs = H/tp, the rank-order sum over several sources, all-gather pushes to several peers."""
import ctypes as C
import os

import numpy as np
import pytest
import torch
import torch.multiprocessing as mp

from oracle import oracle as O

pytestmark = pytest.mark.gpu
SHAPE = O.ModelShape(2048, 512, 3, 8, 2, 64, 1408, 256)
SHAPE8 = O.ModelShape(2048, 512, 3, 8, 8, 64, 2048, 256)   # 8 kv heads: shards over 4 and 8 ranks
SEED = 21


def needs(world):
    return pytest.mark.skipif(not torch.cuda.is_available() or torch.cuda.device_count() < world, reason=f"needs {world} GPUs")


WORLDS = [pytest.param(w, marks=needs(w), id=f"tp{w}") for w in (2, 4, 8)]


def shape_for(world):
    return SHAPE if world == 2 else SHAPE8


def _rank(rank, world, idq, outq, prompts, n_new, temps, hq, custom, SHAPE=None, SEED=None, step_tokens=256, env=None):
    SHAPE = SHAPE or globals()["SHAPE"]
    SEED = SEED or globals()["SEED"]
    os.environ.update(env or {})
    if custom:  # "oneshot" / "twoshot": which peer-memory exchange kernel the engine uses (k_tp.cu)
        os.environ["DIS_TP_EXCHANGE"] = custom
    from distributed_inference_server_b200 import Engine, InferenceParams, WorkerConfig, shard_plan as SP
    from distributed_inference_server_b200 import binding as B
    from tests._gpu_util import to_engine_shape
    torch.cuda.set_device(rank)
    if rank == 0:
        buf = (C.c_uint8 * 128)()
        assert B.load().dis_tp_unique_id(buf) == 0, B.load().dis_last_error()
        for _ in range(world - 1):
            idq.put(bytes(buf))
        nccl_id = bytes(buf)
    else:
        nccl_id = idq.get(timeout=120)
    e = Engine(WorkerConfig(to_engine_shape(SHAPE), device=rank, tp_rank=rank, tp_size=world, max_batch=8, kv_pages=64,
                            page_size=16, max_step_tokens=step_tokens, nccl_id=nccl_id))
    e.load_synthetic(SEED)
    e.finalize()
    if custom:  # fused all-reduce over peer memory: exchange the CUDA IPC handles of the symmetric buffers
        mine = e.tp_export()
        for r in range(world):
            if r != rank:
                hq[r].put((rank, mine))
        handles = {rank: mine}
        while len(handles) < world:
            r, hdl = hq[rank].get(timeout=120)
            handles[r] = hdl
        e.tp_import([handles[r] for r in range(world)])
    # each rank holds exactly the plan's slice
    for kind, layer in [(O.WQ, 1), (O.WK, 0), (O.WV, 2), (O.WO, 1), (O.WGATE, 0), (O.WUP, 2), (O.WDOWN, 1), (O.LM_HEAD, 0), (O.EMBED, 0)]:
        layer = min(layer, SHAPE.layers - 1)
        got = O.from_bf16_bits(e.read_tensor(kind, layer))
        want = O.synth_tensor(SHAPE, SEED, kind, layer)[SP.shard_slices(SHAPE, kind, rank, world)].reshape(-1)
        assert np.array_equal(got, want), (rank, kind, layer)
    slots = [e.seq_add(p, InferenceParams(max_tokens=n_new + 1, temperature=t, top_p=0.9, seed=5)) for p, t in zip(prompts, temps)]
    toks, _ = e.prefill(slots)
    out = [[int(t)] for t in toks]
    logits0 = e.last_logits(len(slots))
    lp0 = e.last_logprobs(slots)
    for _ in range(n_new - 1):
        toks, _ = e.decode(slots)
        for i, t in enumerate(toks):
            out[i].append(int(t))
    outq.put((rank, out, logits0, lp0))
    e.close()


def _run(prompts, n_new, temps, custom, world=2, shape=None, seed=None, step_tokens=256, env=None):
    ctx = mp.get_context("spawn")
    idq, outq = ctx.Queue(), ctx.Queue()
    hq = [ctx.Queue() for _ in range(world)]
    procs = [ctx.Process(target=_rank, args=(r, world, idq, outq, prompts, n_new, temps, hq, custom, shape, seed, step_tokens, env))
             for r in range(world)]
    for p in procs:
        p.start()
    res = {}
    for _ in range(world):
        r, out, lg, lp = outq.get(timeout=300)
        res[r] = (out, lg, lp)
    for p in procs:
        p.join(timeout=60)
        assert p.exitcode == 0
    return res



def _prompts():
    rng = np.random.default_rng(2)
    return [rng.integers(0, SHAPE.vocab, n).tolist() for n in (7, 33, 18)]


@pytest.mark.parametrize("world", WORLDS)
@pytest.mark.parametrize("custom", ["oneshot", "twoshot", "ll", "wide", None],
                         ids=["peer-memory", "peer-memory-twoshot", "peer-memory-ll", "peer-memory-wide", "nccl"])
@pytest.mark.parametrize("sampled", [False, True], ids=["greedy", "sampled"])
def test_tp_matches_oracle(sampled, custom, world):
    n_new = 12
    prompts = _prompts()
    temps = [0.7, 0.0, 1.1] if sampled else [0.0, 0.0, 0.0]
    SHAPE = shape_for(world)
    res = _run(prompts, n_new, temps, custom, world=world, shape=SHAPE)
    for r in range(1, world):
        assert res[0][0] == res[r][0]                   # every rank yields the same tokens
        assert np.array_equal(res[0][1], res[r][1])     # and the same gathered logits
    # TokenEvent.logprob: the vocabulary slices' (max, sum-exp) pairs merged across ranks = log softmax of the
    # gathered logits at the chosen token (sampled rows: of logits / temperature)
    for b, t in enumerate(temps):
        z = res[0][1][b].astype(np.float64) / (t if t > 0 else 1.0)
        want_lp = z[res[0][0][b][0]] - (np.log(np.exp(z - z.max()).sum()) + z.max())
        for r in range(world):
            assert abs(float(res[r][2][b]) - want_lp) <= 1e-3 * max(1.0, abs(want_lp)), (r, b, float(res[r][2][b]), want_lp)
    ref = O.Oracle(SHAPE, act_round=False); ref.fill_synthetic(SEED, True)
    want = ref.generate(prompts, 1, n_logit_steps=1)
    rel = np.abs(res[0][1] - want["logits"][:, 0]).max() / np.abs(want["logits"]).max()
    assert rel <= 2e-2, rel
    if not sampled:
        mimic = O.Oracle(SHAPE, act_round=True); mimic.fill_synthetic(SEED, True)
        w = mimic.generate(prompts, n_new, n_logit_steps=n_new)
        for b in range(len(prompts)):
            if res[0][0][b] != w["tokens"][b]:
                i = next(k for k in range(n_new) if res[0][0][b][k] != w["tokens"][b][k])
                assert w["margins"][b, i] <= 5e-3 * np.abs(w["logits"][b, i]).max(), (b, i)


@pytest.mark.parametrize("world", WORLDS)
def test_twoshot_exchange_is_bit_identical_to_oneshot(world):
    """All exchange kernels add the same contributions in the same (rank) order, then the residual: the logits
    and the greedy tokens must agree bit for bit -- at every world size."""
    prompts = _prompts()
    kw = dict(world=world, shape=shape_for(world))
    a = _run(prompts, 24, [0.0] * 3, "oneshot", **kw)
    a2 = _run(prompts, 24, [0.0] * 3, "oneshot", **kw)
    assert a[0][0] == a2[0][0] and np.array_equal(a[0][1], a2[0][1]), "the one-shot exchange is not run-to-run deterministic"
    b = _run(prompts, 24, [0.0] * 3, "twoshot", **kw)
    b2 = _run(prompts, 24, [0.0] * 3, "twoshot", **kw)
    assert b[0][0] == b2[0][0] and np.array_equal(b[0][1], b2[0][1]), "the two-shot exchange is not run-to-run deterministic"
    diff = np.abs(a[0][1] - b[0][1])
    assert np.array_equal(a[0][1], b[0][1]), (float(diff.max()), int((diff > 0).sum()), diff.size)
    assert a[0][0] == b[0][0]
    c = _run(prompts, 24, [0.0] * 3, "ll", **kw)      # flag-in-data variant of the two-shot schedule
    assert np.array_equal(a[0][1], c[0][1]) and a[0][0] == c[0][0]
    # the wide-grid variant associates the row's sum of squares per column block: deterministic and the same on every
    # rank, but not bit-identical to the three row kernels -- it is held to the logit tolerance against them
    d = _run(prompts, 24, [0.0] * 3, "wide", **kw)
    d2 = _run(prompts, 24, [0.0] * 3, "wide", **kw)
    assert d[0][0] == d2[0][0] and np.array_equal(d[0][1], d2[0][1]), "the wide-grid exchange is not run-to-run deterministic"
    for r in range(1, world):
        assert np.array_equal(d[0][1], d[r][1]) and d[0][0] == d[r][0]
    assert np.abs(d[0][1] - a[0][1]).max() <= 2e-2 * np.abs(a[0][1]).max()


WIDE = O.ModelShape(512, 8192, 1, 64, 8, 128, 1024, 128)   # hidden 8192 (Llama-3-70B's): two float4 per thread in the exchange kernels


@pytest.mark.parametrize("world", WORLDS[:2])
@pytest.mark.parametrize("custom", ["oneshot", "twoshot", "ll", "wide"])
def test_tp_wide_hidden_matches_oracle(custom, world):
    """The exchange kernels walk a row of more than 4096 columns in several passes (shared-memory row instead of
    registers): the path Llama-3-70B takes, checked here on a one-layer model of that width."""
    rng = np.random.default_rng(4)
    prompts = [rng.integers(0, WIDE.vocab, n).tolist() for n in (5, 21)]
    res = _run(prompts, 6, [0.0, 0.0], custom, shape=WIDE, world=world)
    for r in range(1, world):
        assert res[0][0] == res[r][0] and np.array_equal(res[0][1], res[r][1])
    ref = O.Oracle(WIDE, act_round=False); ref.fill_synthetic(SEED, True)
    want = ref.generate(prompts, 1, n_logit_steps=1)
    rel = np.abs(res[0][1] - want["logits"][:, 0]).max() / np.abs(want["logits"]).max()
    assert rel <= 2e-2, rel


@needs(2)
def test_tp2_64_greedy_tokens_identical_to_transformers():
    """The 64-token fixture of tests/golden (clear margin at every step) on two GPUs: both ranks produce the
    transformers continuation token for token."""
    z = np.load(os.path.join(os.path.dirname(__file__), "golden", "mini_llama3_64.npz"))
    shape = O.ModelShape(*[int(v) for v in z["shape"]], float(z["rope_theta"]), float(z["rms_eps"]))
    res = _run([z["prompt"].tolist()], 64, [0.0], "ll", shape=shape, seed=int(z["weight_seed"]))
    assert res[0][0] == res[1][0] == [z["greedy"].tolist()]



@pytest.mark.parametrize("world", WORLDS[:2])
def test_tp_prefill_pass_larger_than_the_decode_exchange_region(world):
    """DIS_TP_PREFILL_FUSED=1: a prefill pass of more rows than the decode-sized exchange region holds (128 here) takes the
    second region of the symmetric buffer and the fenced two-shot kernel instead of reduce + ncclAllReduce + norm (the
    default, which is faster: see Engine::init): 330 prompt tokens in one pass.  Checked against the fp32 oracle, and
    against the NCCL path (same additions up to the order in which the ranks' partial sums meet: held to the logit
    tolerance, greedy ids compared with the margin rule)."""
    shape = shape_for(world)
    rng = np.random.default_rng(6)
    prompts = [rng.integers(0, shape.vocab, n).tolist() for n in (160, 110, 60)]
    n_new = 6
    kw = dict(world=world, shape=shape, step_tokens=512)
    fused = _run(prompts, n_new, [0.0] * 3, "ll", env={"DIS_TP_PREFILL_FUSED": "1"}, **kw)
    nccl = _run(prompts, n_new, [0.0] * 3, "ll", **kw)
    for r in range(1, world):
        assert fused[0][0] == fused[r][0] and np.array_equal(fused[0][1], fused[r][1])
    ref = O.Oracle(shape, act_round=False); ref.fill_synthetic(SEED, True)
    want = ref.generate(prompts, 1, n_logit_steps=1)
    scale = np.abs(want["logits"]).max()
    assert np.abs(fused[0][1] - want["logits"][:, 0]).max() <= 2e-2 * scale
    assert np.abs(fused[0][1] - nccl[0][1]).max() <= 2e-2 * scale
    mimic = O.Oracle(shape, act_round=True); mimic.fill_synthetic(SEED, True)
    w = mimic.generate(prompts, n_new, n_logit_steps=n_new)
    for b in range(len(prompts)):
        if fused[0][0][b] != w["tokens"][b]:
            i = next(k for k in range(n_new) if fused[0][0][b][k] != w["tokens"][b][k])
            assert w["margins"][b, i] <= 5e-3 * np.abs(w["logits"][b, i]).max(), (b, i)


@pytest.mark.parametrize("world", WORLDS)
def test_tp_vocab_parallel_sampler_draws_what_the_gathered_sampler_draws(world):
    """Temperature / nucleus rows under tensor parallelism: the default keeps the logits sharded and exchanges scalars
    over peer memory (k_tp_sample.cu); DIS_TP_SAMPLE=gather all-gathers the logits through NCCL and runs the one-GPU
    sampler on every rank.  Same exchange kernel in front of both, so the logits are bit-identical and the draws must be
    the same tokens (a draw within fp32 noise of a bucket edge may differ: at most one of the six sequences), with the
    same log-probabilities, on every rank."""
    shape = shape_for(world)
    rng = np.random.default_rng(9)
    prompts = [rng.integers(0, shape.vocab, n).tolist() for n in (7, 33, 18, 2, 25, 11)]
    temps = [0.7, 1.1, 0.0, 1.6, 0.3, 0.9]
    kw = dict(world=world, shape=shape)
    fused = _run(prompts, 10, temps, "ll", **kw)
    gathered = _run(prompts, 10, temps, "ll", env={"DIS_TP_SAMPLE": "gather"}, **kw)
    for r in range(1, world):
        assert fused[0][0] == fused[r][0] and np.array_equal(fused[0][2], fused[r][2])
    assert np.array_equal(fused[0][1], gathered[0][1])             # the first pass's logits: same bits
    same = sum(a == b for a, b in zip(fused[0][0], gathered[0][0]))
    assert same >= len(prompts) - 1, (fused[0][0], gathered[0][0])
    assert fused[0][0][2] == gathered[0][0][2]                     # the greedy row among them is exact
    lp_f, lp_g = np.asarray(fused[0][2]), np.asarray(gathered[0][2])
    first_same = [fused[0][0][b][0] == gathered[0][0][b][0] for b in range(len(prompts))]
    assert np.abs(lp_f - lp_g)[first_same].max() <= 1e-4
