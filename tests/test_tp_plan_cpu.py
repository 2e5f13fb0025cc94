"""The N>1 path on CPU: world_size-2 gloo processes apply the product's tensor-parallel shard plan
(distributed_inference_server_b200/shard_plan.py, the same slicing csrc/engine.cu does) to the
oracle's weights, run the per-rank decoder with the oracle's primitives, exchange with the same two
all-reduces per layer + candidate all-gather the engine uses, and must reproduce the unsharded oracle."""
import os

import numpy as np
import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

from distributed_inference_server_b200 import shard_plan as SP
from oracle import oracle as O

SHAPE = O.ModelShape(512, 128, 2, 8, 2, 16, 352, 64)
SEED = 17


def _rank_main(rank, world, port, tokens, q):
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        s = SHAPE
        d, H = s.head_dim, s.hidden
        nh_l, nkv_l = s.n_heads // world, s.n_kv_heads // world
        def w(kind, layer=0):
            full = O.synth_tensor(s, SEED, kind, layer)
            return full[SP.shard_slices(s, kind, rank, world)] if full.ndim == 2 else full
        cs, sn = O.rope_table(s.max_ctx, d, s.rope_theta)
        T = len(tokens)
        pos = np.arange(T, dtype=np.int32)
        x = w(O.EMBED)[tokens].copy()

        def allreduce(a):
            t = torch.from_numpy(np.ascontiguousarray(a))
            dist.all_reduce(t)
            return t.numpy()

        for l in range(s.layers):
            n1 = O.rmsnorm(x, w(O.ATTN_NORM, l), s.rms_eps)
            qh = O.rope(O.linear(n1, w(O.WQ, l)), pos, nh_l, d, cs, sn)
            kh = O.rope(O.linear(n1, w(O.WK, l)), pos, nkv_l, d, cs, sn)
            vh = O.linear(n1, w(O.WV, l))
            att = np.stack([O.attention_one(qh[t], kh[:t + 1], vh[:t + 1], nh_l, nkv_l, d) for t in range(T)])
            x = x + allreduce(O.linear(att, w(O.WO, l)))                 # row-parallel -> all-reduce
            n2 = O.rmsnorm(x, w(O.MLP_NORM, l), s.rms_eps)
            act = O.silu_mul(O.linear(n2, w(O.WGATE, l)), O.linear(n2, w(O.WUP, l)))
            x = x + allreduce(O.linear(act, w(O.WDOWN, l)))              # row-parallel -> all-reduce
        logits_l = O.linear(O.rmsnorm(x, w(O.FINAL_NORM), s.rms_eps), w(O.LM_HEAD))
        val = logits_l.max(1).astype(np.float32)
        tok = (logits_l.argmax(1) + SP.vocab_offset(s, rank, world)).astype(np.int64)
        vals = [torch.zeros(T) for _ in range(world)]
        toks = [torch.zeros(T, dtype=torch.int64) for _ in range(world)]
        dist.all_gather(vals, torch.from_numpy(val))
        dist.all_gather(toks, torch.from_numpy(tok))
        full = [torch.zeros(logits_l.shape) for _ in range(world)]
        dist.all_gather(full, torch.from_numpy(logits_l))
        if rank == 0:
            q.put((np.concatenate([f.numpy() for f in full], axis=1),
                   SP.merge_greedy([v.tolist() for v in vals], [t.tolist() for t in toks])))
    finally:
        dist.destroy_process_group()


@pytest.mark.parametrize("world", [2])
def test_sharded_forward_equals_unsharded(world):
    tokens = np.random.default_rng(3).integers(0, SHAPE.vocab, 9).astype(np.int64)
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    port = 29500 + os.getpid() % 2000
    procs = [ctx.Process(target=_rank_main, args=(r, world, port, tokens, q)) for r in range(world)]
    for p in procs:
        p.start()
    logits, greedy = q.get(timeout=90)
    for p in procs:
        p.join(timeout=60)
        assert p.exitcode == 0
    ref = O.Oracle(SHAPE)
    ref.fill_synthetic(SEED, True)
    sid = ref.seq_new()
    want = ref.forward(tokens.astype(np.int32), [sid] * len(tokens), np.arange(len(tokens)))
    assert np.abs(logits - want).max() <= 1e-4 * np.abs(want).max()
    assert greedy == want.argmax(1).tolist()


def test_shard_slices_partition_every_tensor():
    s = O.LLAMA3_8B
    for tp in (1, 2, 4, 8):
        for kind in range(12):
            shp = O.tensor_shape(s, kind)
            if len(shp) == 1:
                continue
            cover = np.zeros(shp[:2], np.int8) if np.prod(shp) < 1e7 else None
            rows = cols = 0
            for r in range(tp):
                rs, cs_ = SP.shard_slices(s, kind, r, tp)
                nr = len(range(*rs.indices(shp[0]))); nc = len(range(*cs_.indices(shp[1])))
                rows += nr; cols += nc
                if cover is not None:
                    cover[rs, cs_] += 1
            replicated = kind in (O.EMBED,)
            if replicated:
                assert rows == tp * shp[0] and cols == tp * shp[1]
            else:
                assert (rows == shp[0] and cols == tp * shp[1]) or (rows == tp * shp[0] and cols == shp[1])
            if cover is not None and not replicated:
                assert (cover == 1).all()
    with pytest.raises(ValueError):
        SP.shard_slices(s, O.WQ, 0, 3)
