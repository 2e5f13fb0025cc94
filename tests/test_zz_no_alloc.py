"""dis_k_attn_paged_static -- split-KV attention for hosts that cannot allocate device memory (a server
whose KV-page pool already owns all free HBM, or a loop that must avoid cudaMalloc's device-wide stall).
Its partials live in an array inside the library, so this test drives it through the C ABI next to the
ordinary entry point (caller-provided scratch) and expects the same output, twice in a row with different
inputs: the second call must not be disturbed by what the first one left in the shared array.
The file name makes pytest collect it last."""
import numpy as np
import pytest
import torch

from distributed_inference_server_b200 import binding as B
from oracle import oracle as O
from tests import _gpu_util as G
from tests._gpu_util import DEV, check, dev_bf16, dev_i32, host, lib, ptr, rand_bf16, sync

pytestmark = pytest.mark.gpu
BF16_ULP = 2.0 ** -8
LIMIT = 2 << 20  # DIS_ATTN_STATIC_SCRATCH_FLOATS


def _case(seed, d, nh, nkv, page_size, ctx_lens, splits):
    rng = np.random.default_rng(seed)
    n_seq, n_pages = len(ctx_lens), 48
    max_pages = (max(ctx_lens) + page_size - 1) // page_size + 1
    perm = rng.permutation(n_pages)
    pt = np.zeros((n_seq, max_pages), np.int32)
    k = 0
    for s in range(n_seq):
        for p in range((ctx_lens[s] + page_size - 1) // page_size):
            pt[s, p] = perm[k]; k += 1
    pool = dev_bf16(rand_bf16(rng, (n_pages, 2, nkv, page_size, d)))
    q = dev_bf16(rand_bf16(rng, (n_seq, nh * d)))
    args = dict(q=q, pool=pool, n_pages=n_pages, pt=dev_i32(pt), max_pages=max_pages, row=dev_i32(np.arange(n_seq)),
                ctx=dev_i32(ctx_lens), T=n_seq, nh=nh, nkv=nkv, d=d, ps=page_size, splits=splits, host_pt=pt, ctx_lens=ctx_lens)
    return args


def _oracle_attention(a):
    """The CPU checker's single-query attention (oracle/llama_oracle.c orc_attention_one) over each sequence's pages."""
    pool, q = host(a["pool"]), host(a["q"])
    out = np.zeros((a["T"], a["nh"] * a["d"]), np.float32)
    for s in range(a["T"]):
        n = a["ctx_lens"][s]
        K = np.zeros((n, a["nkv"] * a["d"]), np.float32); V = np.zeros_like(K)
        for t in range(n):
            page, r = a["host_pt"][s, t // a["ps"]], t % a["ps"]
            K[t] = pool[page, 0, :, r, :].reshape(-1)
            V[t] = pool[page, 1, :, r, :].reshape(-1)
        out[s] = O.attention_one(q[s], K, V, a["nh"], a["nkv"], a["d"])
    return out


def _run_both(a):
    want = G.empty((a["T"], a["nh"] * a["d"]), dtype=torch.bfloat16, device=DEV)
    got = G.empty(want.shape, dtype=torch.bfloat16, device=DEV)
    got.fill_(float("nan"))
    n = lib().dis_k_attn_scratch_elems(a["T"], a["nh"], a["d"], a["splits"])
    scr = G.empty(max(n, 1), dtype=torch.float32, device=DEV)
    check(lib().dis_k_attn_paged(ptr(a["q"]), ptr(a["pool"]), a["n_pages"], ptr(a["pt"]), a["max_pages"], ptr(a["row"]),
                                 ptr(a["ctx"]), ptr(want), ptr(scr), a["splits"], a["T"], a["nh"], a["nkv"], a["d"], a["ps"], None))
    check(lib().dis_k_attn_paged_static(ptr(a["q"]), ptr(a["pool"]), a["n_pages"], ptr(a["pt"]), a["max_pages"], ptr(a["row"]),
                                        ptr(a["ctx"]), ptr(got), a["splits"], a["T"], a["nh"], a["nkv"], a["d"], a["ps"], None))
    sync()
    return host(got), host(want)


def test_static_scratch_matches_caller_scratch_twice_in_a_row():
    first = _case(1, 128, 8, 2, 16, [200, 37, 130, 1], splits=4)
    second = _case(2, 64, 8, 1, 64, [5, 500, 64], splits=7)       # other head size, GQA group and split count
    for case in (first, second):
        got, want = _run_both(case)
        assert np.isfinite(got).all()
        # same kernels, same summation order: identical; the bound is the one the attention parity tests use
        assert np.abs(got - want).max() <= 2 * BF16_ULP * np.abs(want).max() + 1e-5
        assert np.array_equal(got, want)
        # ... and, standing on its own, against the CPU oracle (P is rounded to bf16 for the P.V MMA: two bf16 ulps
        # of the output scale, the bound of the attention parity tests in test_kernels_gpu.py)
        ref = _oracle_attention(case)
        assert np.abs(got - ref).max() <= 2 * BF16_ULP * np.abs(ref).max() + 1e-5


def test_static_scratch_allocates_nothing_and_refuses_oversize():
    a = _case(3, 128, 8, 2, 16, [90, 300], splits=3)
    _run_both(a)                                                   # first call on the device: symbol lookup done
    out = G.empty((a["T"], a["nh"] * a["d"]), dtype=torch.bfloat16, device=DEV)
    sync()
    free_before, _ = torch.cuda.mem_get_info()
    check(lib().dis_k_attn_paged_static(ptr(a["q"]), ptr(a["pool"]), a["n_pages"], ptr(a["pt"]), a["max_pages"], ptr(a["row"]),
                                        ptr(a["ctx"]), ptr(out), a["splits"], a["T"], a["nh"], a["nkv"], a["d"], a["ps"], None))
    sync()
    free_after, _ = torch.cuda.mem_get_info()
    assert free_after == free_before                               # no device memory was taken
    # T * n_heads * n_splits * (2 + d) beyond the built-in array is refused, with both numbers in the message
    big_T, big_splits = 4096, 32
    assert lib().dis_k_attn_scratch_elems(big_T, a["nh"], a["d"], big_splits) > LIMIT
    rc = lib().dis_k_attn_paged_static(ptr(a["q"]), ptr(a["pool"]), a["n_pages"], ptr(a["pt"]), a["max_pages"], ptr(a["row"]),
                                       ptr(a["ctx"]), ptr(out), big_splits, big_T, a["nh"], a["nkv"], a["d"], a["ps"], None)
    assert rc == B.ERR_INVALID_ARGUMENT
    msg = lib().dis_last_error().decode()
    assert str(LIMIT) in msg and "do not fit" in msg
