"""Kernel-level parity: each hand-written CUDA kernel, driven through the C ABI, against the oracle's
primitive of the same name on the same seeded inputs.  Tolerances: bit-exact for integer/bf16-bit
work (synthetic weights, argmax, page writes of v); fp32-accumulation noise (1e-4 of the output scale)
for reductions whose summation order differs; one bf16 ulp where the kernel's output is bf16."""
import ctypes as C

import numpy as np
import pytest
import torch

from oracle import oracle as O
from tests import _gpu_util as G
from tests._gpu_util import (DEV, check, dev_bf16, dev_f32, dev_i32, host, lib, ptr, rand_bf16, sync)

pytestmark = pytest.mark.gpu
BF16_ULP = 2.0 ** -8  # relative spacing of bf16


def test_synth_bits_match_oracle():
    shape = O.ModelShape(512, 128, 3, 8, 2, 16, 352, 256)
    for kind, layer in [(O.WQ, 1), (O.ATTN_NORM, 0), (O.WDOWN, 2), (O.EMBED, 0)]:
        ref = O.synth_tensor(shape, 99, kind, layer).reshape(-1)
        amp, off = C.c_float(), C.c_float()
        cfg = shape.c()
        O.lib().orc_synth_params(C.byref(cfg), kind, C.byref(amp), C.byref(off))
        out = G.empty(ref.size, dtype=torch.bfloat16, device=DEV)
        check(lib().dis_k_synth_bf16(ptr(out), ref.size, 99, kind, layer if kind in O.PER_LAYER else 0, amp.value,
                                     off.value, None))
        sync()
        assert np.array_equal(host(out), ref)


@pytest.mark.parametrize("T,H", [(1, 128), (7, 512), (33, 4096), (4, 8192)])
def test_add_rmsnorm(T, H):
    rng = np.random.default_rng(T * 1000 + H)
    x = rng.standard_normal((T, H)).astype(np.float32)
    parts = rng.standard_normal((3, T, H)).astype(np.float32)
    gain = O.bf16_round(1 + 0.1 * rng.standard_normal(H).astype(np.float32))
    dx, dp, dg = dev_f32(x), dev_f32(parts), dev_f32(gain)
    xn = G.empty((T, H), dtype=torch.bfloat16, device=DEV)
    check(lib().dis_k_add_rmsnorm(ptr(dx), None, None, ptr(dp), 3, T * H, ptr(dg), ptr(xn), None, T, H, 1e-5, None))
    sync()
    xs = x + parts.sum(0)
    assert np.abs(host(dx) - xs).max() <= 1e-5 * np.abs(xs).max()
    ref = O.rmsnorm(xs, gain, 1e-5, round_out=True)
    assert np.abs(host(xn) - ref).max() <= BF16_ULP * np.abs(ref).max() * 1.01


def test_add_rmsnorm_embedding_and_rowmap():
    rng = np.random.default_rng(5)
    V, H, T = 300, 256, 9
    emb = rand_bf16(rng, (V, H), 0.02)
    toks = rng.integers(0, V, T).astype(np.uint32)
    gain = np.ones(H, np.float32)
    dx = G.zeros((T, H), dtype=torch.float32, device=DEV)
    xn = G.empty((T, H), dtype=torch.bfloat16, device=DEV)
    dt = torch.from_numpy(toks.astype(np.int64)).to(DEV).to(torch.int32)
    check(lib().dis_k_add_rmsnorm(ptr(dx), ptr(dt), ptr(dev_bf16(emb)), None, 0, 0, ptr(dev_f32(gain)), ptr(xn), None,
                                  T, H, 1e-5, None))
    sync()
    assert np.array_equal(host(dx), emb[toks])
    ref = O.rmsnorm(emb[toks], gain, 1e-5, round_out=True)
    assert np.abs(host(xn) - ref).max() <= BF16_ULP * np.abs(ref).max() * 1.01
    # row_map: compact rows, x untouched
    rows = np.array([8, 2, 5], np.int32)
    xn2 = G.empty((3, H), dtype=torch.bfloat16, device=DEV)
    before = host(dx).copy()
    check(lib().dis_k_add_rmsnorm(ptr(dx), None, None, None, 0, 0, ptr(dev_f32(gain)), ptr(xn2), ptr(dev_i32(rows)), 3,
                                  H, 1e-5, None))
    sync()
    assert np.array_equal(host(dx), before)
    assert np.array_equal(host(xn2), host(xn)[rows])


GEMM_SHAPES = [(32, 256, 128), (5, 200, 352), (64, 384, 512), (17, 1000, 1024), (128, 512, 256), (300, 640, 192),
               (32, 4096, 4096),
               # prefill-sized: the persistent, TMEM-double-buffered kernel (>= 2 tiles per SM), ragged T / N / K too
               (1024, 9600, 576), (1000, 9500, 320), (4096, 2500, 128)]


@pytest.mark.parametrize("T,N,K", GEMM_SHAPES)
def test_gemm_tcgen05(T, N, K):
    rng = np.random.default_rng(T + N + K)
    X, W = rand_bf16(rng, (T, K)), rand_bf16(rng, (N, K), 0.05)
    ref = O.linear(X, W)
    dX, dW = dev_bf16(X), dev_bf16(W)
    for split in sorted({1, lib().dis_k_gemm_pick_split(T, N, K)}):
        out = torch.full((split, T, N), float("nan"), dtype=torch.float32, device=DEV)
        check(lib().dis_k_gemm(ptr(dX), ptr(dW), ptr(out), T, N, K, split, None))
        sync()
        got = host(out).sum(0)
        assert np.isfinite(got).all()
        assert np.abs(got - ref).max() <= 1e-4 * np.abs(ref).max() + 1e-6, (split, np.abs(got - ref).max())


@pytest.mark.parametrize("T,N,K,silu", [(8192, 6144, 4096, False), (4096, 4096, 14336, False), (8192, 14336, 4096, True),
                                        (32, 28672 // 2, 4096, True), (512, 6144, 4096, False)])
def test_gemm_full_size_against_torch(T, N, K, silu):
    """BASELINE-sized projections (Llama-3-8B qkv / down / gate-up at an 8192-token prefill chunk, a decode batch, a
    single 512-token request): too large for the CPU oracle in seconds, so the reference is a plain torch fp32 matmul
    of the same bf16 operands on the device (tolerance: fp32 accumulation order only, 2e-4 of max|y|; the SiLU-gate
    output is bf16, one ulp)."""
    g = torch.Generator(device=DEV); g.manual_seed(T + N + K)
    X = (torch.randn((T, K), device=DEV, generator=g)).to(torch.bfloat16)
    W = (torch.randn(((2 if silu else 1) * N, K), device=DEV, generator=g) * 0.02).to(torch.bfloat16)
    if silu:
        act = G.empty((T, N), dtype=torch.bfloat16, device=DEV)
        check(lib().dis_k_gemm_silu(ptr(X), ptr(W), ptr(act), T, N, K, None))
        sync()
        gate, up = X.float() @ W[:N].float().T, X.float() @ W[N:].float().T
        ref = torch.nn.functional.silu(gate) * up
        err = (act.float() - ref).abs().max().item()
        assert err <= BF16_ULP * ref.abs().max().item() * 1.05 + 1e-6, err
    else:
        out = torch.full((1, T, N), float("nan"), dtype=torch.float32, device=DEV)
        check(lib().dis_k_gemm(ptr(X), ptr(W), ptr(out), T, N, K, 1, None))
        sync()
        ref = X.float() @ W.float().T
        assert torch.isfinite(out).all()
        err = (out[0] - ref).abs().max().item()
        assert err <= 2e-4 * ref.abs().max().item(), err


def test_gemm_rejects_empty_split():
    X, W = dev_bf16(np.zeros((8, 576), np.float32)), dev_bf16(np.zeros((128, 576), np.float32))
    out = G.zeros((4, 8, 128), dtype=torch.float32, device=DEV)
    assert lib().dis_k_gemm(ptr(X), ptr(W), ptr(out), 8, 128, 576, 4, None) != 0  # 9 k-blocks / 4 -> empty slice


@pytest.mark.parametrize("T,I,K", [(32, 256, 128), (9, 200, 320), (64, 512, 256), (128, 384, 128), (32, 14336, 4096),
                                   (1024, 4800, 192), (1000, 4792, 320), (2048, 2432, 64)])  # persistent dual kernel
def test_gemm_silu(T, I, K):
    rng = np.random.default_rng(T + I + K)
    X, Wgu = rand_bf16(rng, (T, K)), rand_bf16(rng, (2 * I, K), 0.05)
    ref = O.silu_mul(O.linear(X, Wgu[:I]), O.linear(X, Wgu[I:]), round_out=False)
    act = G.empty((T, I), dtype=torch.bfloat16, device=DEV)
    check(lib().dis_k_gemm_silu(ptr(dev_bf16(X)), ptr(dev_bf16(Wgu)), ptr(act), T, I, K, None))
    sync()
    assert np.abs(host(act) - ref).max() <= BF16_ULP * np.abs(ref).max() * 1.05 + 1e-6


@pytest.mark.parametrize("T,N,K", [(1, 256, 128), (2, 201, 352), (4, 512, 4096), (7, 64, 256),
                                   # several row groups per persistent CTA, ragged N and K tails
                                   (1, 4096, 4096), (1, 3001, 6152), (2, 1500, 14336), (3, 2501, 520)])
def test_gemv(T, N, K):
    rng = np.random.default_rng(T + N + K)
    X, W = rand_bf16(rng, (T, K)), rand_bf16(rng, (N, K), 0.05)
    ref = O.linear(X, W)
    out = G.empty((T, N), dtype=torch.float32, device=DEV)
    check(lib().dis_k_gemv(ptr(dev_bf16(X)), ptr(dev_bf16(W)), ptr(out), T, N, K, None))
    sync()
    assert np.abs(host(out) - ref).max() <= 1e-4 * np.abs(ref).max() + 1e-6
    Wgu = rand_bf16(rng, (2 * N, K), 0.05)
    refs = O.silu_mul(O.linear(X, Wgu[:N]), O.linear(X, Wgu[N:]))
    act = G.empty((T, N), dtype=torch.bfloat16, device=DEV)
    check(lib().dis_k_gemv_silu(ptr(dev_bf16(X)), ptr(dev_bf16(Wgu)), ptr(act), T, N, K, None))
    sync()
    assert np.abs(host(act) - refs).max() <= BF16_ULP * np.abs(refs).max() * 1.05 + 1e-6


def _paged_setup(rng, n_seq, ctx_lens, n_kv, d, page_size, n_pages):
    max_pages = (max(ctx_lens) + page_size - 1) // page_size + 1
    perm = rng.permutation(n_pages)
    pt = np.zeros((n_seq, max_pages), np.int32)
    k = 0
    for s in range(n_seq):
        for p in range((ctx_lens[s] + page_size - 1) // page_size):
            pt[s, p] = perm[k]; k += 1
    return pt, max_pages


@pytest.mark.parametrize("d,nh,nkv", [(128, 8, 2), (64, 8, 2), (128, 4, 4), (64, 4, 2), (128, 8, 1), (64, 16, 2)])
def test_rope_kv_write_and_attention(d, nh, nkv):
    rng = np.random.default_rng(d + nh + nkv)
    page_size, n_pages, max_ctx = 16, 64, 256
    ctx_lens = [1, 37, 130, 200]
    n_seq = len(ctx_lens)
    pt, max_pages = _paged_setup(rng, n_seq, ctx_lens, nkv, d, page_size, n_pages)
    cs, sn = O.rope_table(max_ctx, d, 500000.0)
    nq, nk = nh * d, nkv * d
    pool = G.zeros((n_pages, 2, nkv, page_size, d), dtype=torch.bfloat16, device=DEV)
    dpt = dev_i32(pt)
    K_hist = [np.zeros((c, nk), np.float32) for c in ctx_lens]
    V_hist = [np.zeros((c, nk), np.float32) for c in ctx_lens]
    q_last = np.zeros((n_seq, nq), np.float32)
    # write every position of every sequence through the kernel, a ragged batch at a time, with 2 split-K slices
    toks = [(s, p) for s in range(n_seq) for p in range(ctx_lens[s])]
    for b0 in range(0, len(toks), 50):
        chunk = toks[b0:b0 + 50]
        T = len(chunk)
        full = (rng.standard_normal((T, nq + 2 * nk))).astype(np.float32)
        part = rng.standard_normal((2, T, nq + 2 * nk)).astype(np.float32)
        part[1] = full - part[0]
        summed = part[0] + part[1]
        pos = np.array([p for _, p in chunk], np.int32)
        row = np.array([s for s, _ in chunk], np.int32)
        dq = G.empty((T, nq), dtype=torch.bfloat16, device=DEV)
        check(lib().dis_k_rope_kv_write(ptr(dev_f32(part)), 2, T * (nq + 2 * nk), ptr(dq), ptr(pool), ptr(dpt), max_pages,
                                        ptr(dev_i32(row)), ptr(dev_i32(pos)), ptr(dev_f32(cs)), ptr(dev_f32(sn)), T, nh,
                                        nkv, d, page_size, None))
        sync()
        q_ref = O.bf16_round(O.rope(summed[:, :nq], pos, nh, d, cs, sn))
        k_ref = O.bf16_round(O.rope(summed[:, nq:nq + nk], pos, nkv, d, cs, sn))
        v_ref = O.bf16_round(summed[:, nq + nk:])
        got_q = host(dq)
        assert np.abs(got_q - q_ref).max() <= BF16_ULP * np.abs(q_ref).max() * 1.01
        for i, (s, p) in enumerate(chunk):
            K_hist[s][p], V_hist[s][p] = k_ref[i], v_ref[i]
            if p == ctx_lens[s] - 1:
                q_last[s] = got_q[i]
    hp = host(pool)
    for s in range(n_seq):
        for p in range(ctx_lens[s]):
            page, slot = pt[s, p // page_size], p % page_size
            got_k = hp[page, 0, :, slot, :].reshape(-1)
            got_v = hp[page, 1, :, slot, :].reshape(-1)
            assert np.abs(got_k - K_hist[s][p]).max() <= BF16_ULP * max(np.abs(K_hist[s][p]).max(), 1e-3) * 1.01
            assert np.array_equal(got_v, V_hist[s][p])
            K_hist[s][p], V_hist[s][p] = got_k, got_v  # attention below is checked on what is in the pages
    # attention of the last token of every sequence, 1 and several splits
    dq = dev_bf16(q_last)
    ref = np.stack([O.attention_one(q_last[s], K_hist[s], V_hist[s], nh, nkv, d) for s in range(n_seq)])
    for splits in (1, 3, 8):
        out = G.empty((n_seq, nq), dtype=torch.bfloat16, device=DEV)
        n_scr = lib().dis_k_attn_scratch_elems(n_seq, nh, d, splits)
        scr = G.empty(max(n_scr, 1), dtype=torch.float32, device=DEV)
        check(lib().dis_k_attn_paged(ptr(dq), ptr(pool), n_pages, ptr(dpt), max_pages, ptr(dev_i32(np.arange(n_seq))),
                                     ptr(dev_i32(ctx_lens)), ptr(out), ptr(scr), splits, n_seq, nh, nkv, d, page_size, None))
        sync()
        got = host(out)
        # P is rounded to bf16 for the P.V tensor-core MMA: allow two bf16 ulps of the output scale
        assert np.abs(got - ref).max() <= 2 * BF16_ULP * np.abs(ref).max() + 1e-5, (splits, np.abs(got - ref).max())


def test_attention_full_size_against_torch():
    """Llama-3-8B geometry (32 query heads, 8 kv heads, d = 128, 64-token pages), batch 32 with contexts from 1 to 8000
    tokens on shuffled pages: decode attention with 1 and 4 history splits against softmax(q K^T / sqrt(d)) V in fp32
    torch on the same bf16 K/V (tolerance: P is rounded to bf16 for the P.V MMA -- two bf16 ulps of the output scale)."""
    g = torch.Generator(device=DEV); g.manual_seed(9)
    d, nh, nkv, ps, B = 128, 32, 8, 64, 32
    rng = np.random.default_rng(9)
    ctx = np.concatenate([[1, 63, 64, 65, 8000], rng.integers(100, 4000, B - 5)]).astype(np.int32)
    pages_per = [(int(c) + ps - 1) // ps for c in ctx]
    n_pages, max_pages = sum(pages_per) + 3, max(pages_per)
    perm = rng.permutation(n_pages)
    pt = np.zeros((B, max_pages), np.int32)
    k0 = 0
    for b in range(B):
        pt[b, :pages_per[b]] = perm[k0:k0 + pages_per[b]]; k0 += pages_per[b]
    pool = torch.randn((n_pages, 2, nkv, ps, d), device=DEV, generator=g).to(torch.bfloat16)
    q = torch.randn((B, nh * d), device=DEV, generator=g).to(torch.bfloat16)
    dpt, dctx, drow = dev_i32(pt), dev_i32(ctx), dev_i32(np.arange(B))
    ref = G.empty((B, nh, d), dtype=torch.float32, device=DEV)
    for b in range(B):
        pages = torch.as_tensor(pt[b, :pages_per[b]].astype(np.int64), device=DEV)
        kv = pool[pages].float()                                   # [pages, 2, nkv, ps, d]
        K = kv[:, 0].permute(1, 0, 2, 3).reshape(nkv, -1, d)[:, :int(ctx[b])]
        V = kv[:, 1].permute(1, 0, 2, 3).reshape(nkv, -1, d)[:, :int(ctx[b])]
        qb = q[b].float().reshape(nkv, nh // nkv, d)
        p = torch.softmax(qb @ K.transpose(1, 2) / d ** 0.5, dim=-1)
        ref[b] = (p @ V).reshape(nh, d)
    for splits in (1, 4):
        out = G.empty((B, nh * d), dtype=torch.bfloat16, device=DEV)
        scr = G.empty(max(lib().dis_k_attn_scratch_elems(B, nh, d, splits), 1), dtype=torch.float32, device=DEV)
        check(lib().dis_k_attn_paged(ptr(q), ptr(pool), n_pages, ptr(dpt), max_pages, ptr(drow), ptr(dctx), ptr(out), ptr(scr),
                                     splits, B, nh, nkv, d, ps, None))
        sync()
        err = (out.float().reshape(B, nh, d) - ref).abs().max().item()
        assert err <= 2 * BF16_ULP * ref.abs().max().item() + 1e-5, (splits, err)


def test_attention_long_context_tiles():
    rng = np.random.default_rng(3)
    d, nh, nkv, page_size = 128, 4, 1, 64
    ctx = 1500  # several 256-token tiles per split
    n_pages = 32
    pt, max_pages = _paged_setup(rng, 1, [ctx], nkv, d, page_size, n_pages)
    K = rand_bf16(rng, (ctx, d), 1.0)
    V = rand_bf16(rng, (ctx, d), 1.0)
    pool = np.zeros((n_pages, 2, nkv, page_size, d), np.float32)
    for p in range(ctx):
        pool[pt[0, p // page_size], 0, 0, p % page_size] = K[p]
        pool[pt[0, p // page_size], 1, 0, p % page_size] = V[p]
    q = rand_bf16(rng, (1, nh * d), 1.0)
    ref = O.attention_one(q[0], K, V, nh, nkv, d)
    dpool = dev_bf16(pool)
    for splits in (1, 2, 5):
        out = G.empty((1, nh * d), dtype=torch.bfloat16, device=DEV)
        scr = G.empty(max(lib().dis_k_attn_scratch_elems(1, nh, d, splits), 1), dtype=torch.float32, device=DEV)
        check(lib().dis_k_attn_paged(ptr(dev_bf16(q)), ptr(dpool), n_pages, ptr(dev_i32(pt)), max_pages, ptr(dev_i32([0])),
                                     ptr(dev_i32([ctx])), ptr(out), ptr(scr), splits, 1, nh, nkv, d, page_size, None))
        sync()
        assert np.abs(host(out)[0] - ref).max() <= 2 * BF16_ULP * np.abs(ref).max() + 1e-5


@pytest.mark.parametrize("d,nh,nkv,page_size", [(128, 8, 2, 64), (64, 8, 2, 16), (128, 8, 1, 32), (64, 2, 2, 64), (128, 4, 1, 16),
                                                 (128, 16, 2, 32)])
@pytest.mark.parametrize("short_blocks", [False, True], ids=["full-blocks", "short-blocks"])
def test_attention_prefill_blocks(d, nh, nkv, page_size, short_blocks):
    """Causal attention of query blocks over the paged cache vs the oracle, ragged sequences.  Blocks hold up to
    dis_k_attn_prefill_block_tokens tokens: 32 / 16 on the tcgen05 kernel (head_dim 128 with 4 / 8 query heads per kv
    head), 16 on the mma.sync kernel; short-blocks cuts them at 5 tokens (mostly padding rows, as the decode rows of a
    mixed step are)."""
    BT = lib().dis_k_attn_prefill_block_tokens(nh, nkv, d)
    assert BT == (128 // (nh // nkv) if d == 128 and nh // nkv in (4, 8) else 16)
    if short_blocks:
        BT = 5
    rng = np.random.default_rng(d + nh + nkv + page_size)
    seq_lens = [1, 15, 16, 17, 70, 200]
    n_pages = 32
    pt, max_pages = _paged_setup(rng, len(seq_lens), seq_lens, nkv, d, page_size, n_pages)
    pool = np.zeros((n_pages, 2, nkv, page_size, d), np.float32)
    K = [rand_bf16(rng, (n, nkv * d)) for n in seq_lens]
    V = [rand_bf16(rng, (n, nkv * d)) for n in seq_lens]
    Q = [rand_bf16(rng, (n, nh * d)) for n in seq_lens]
    for s, n in enumerate(seq_lens):
        for p in range(n):
            pool[pt[s, p // page_size], 0, :, p % page_size] = K[s][p].reshape(nkv, d)
            pool[pt[s, p // page_size], 1, :, p % page_size] = V[s][p].reshape(nkv, d)
    # the second half of the longest sequence only: blocks need not start at position 0 (chunked prefill)
    starts = [0, 0, 0, 0, 0, 96]
    q_flat, tok0, qlen, pos0, ptrow, ref = [], [], [], [], [], []
    t = 0
    for s, n in enumerate(seq_lens):
        for p0 in range(starts[s], n, BT):
            ln = min(BT, n - p0)
            tok0.append(t); qlen.append(ln); pos0.append(p0); ptrow.append(s)
            for i in range(ln):
                q_flat.append(Q[s][p0 + i])
                ref.append(O.attention_one(Q[s][p0 + i], K[s][:p0 + i + 1], V[s][:p0 + i + 1], nh, nkv, d))
            t += ln
    q_flat, ref = np.stack(q_flat), np.stack(ref)
    out = G.empty((t, nh * d), dtype=torch.bfloat16, device=DEV)
    check(lib().dis_k_attn_prefill(ptr(dev_bf16(q_flat)), ptr(dev_bf16(pool)), n_pages, ptr(dev_i32(pt)), max_pages,
                                   ptr(dev_i32(tok0)), ptr(dev_i32(qlen)), ptr(dev_i32(pos0)), ptr(dev_i32(ptrow)), ptr(out),
                                   len(tok0), nh, nkv, d, page_size, None))
    sync()
    got = host(out)
    assert np.isfinite(got).all()
    assert np.abs(got - ref).max() <= 2 * BF16_ULP * np.abs(ref).max() + 1e-5, np.abs(got - ref).max()


def test_attention_prefill_full_size_against_torch():
    """Llama-3-8B geometry, a 3000-token prompt prefilled from position 0 and the last 1000 tokens of a 4000-token one
    (a later chunk of a chunked prefill), 64-token pages in shuffled order: causal attention against fp32 torch."""
    g = torch.Generator(device=DEV); g.manual_seed(13)
    d, nh, nkv, ps = 128, 32, 8, 64
    lens, starts = [3000, 4000], [0, 3000]
    rng = np.random.default_rng(13)
    pages_per = [(n + ps - 1) // ps for n in lens]
    n_pages, max_pages = sum(pages_per) + 2, max(pages_per)
    perm = rng.permutation(n_pages)
    pt = np.zeros((2, max_pages), np.int32)
    pt[0, :pages_per[0]] = perm[:pages_per[0]]; pt[1, :pages_per[1]] = perm[pages_per[0]:pages_per[0] + pages_per[1]]
    pool = torch.randn((n_pages, 2, nkv, ps, d), device=DEV, generator=g).to(torch.bfloat16)
    tok0, qlen, pos0, ptrow, t = [], [], [], [], 0
    BT = lib().dis_k_attn_prefill_block_tokens(nh, nkv, d)      # 32: the tcgen05 kernel's 128 query rows = 32 tokens x 4 heads
    assert BT == 32
    for s in range(2):
        for p0 in range(starts[s], lens[s], BT):
            ln = min(BT, lens[s] - p0)
            tok0.append(t); qlen.append(ln); pos0.append(p0); ptrow.append(s); t += ln
    q = torch.randn((t, nh * d), device=DEV, generator=g).to(torch.bfloat16)
    out = G.empty((t, nh * d), dtype=torch.bfloat16, device=DEV)
    check(lib().dis_k_attn_prefill(ptr(q), ptr(pool), n_pages, ptr(dev_i32(pt)), max_pages, ptr(dev_i32(tok0)), ptr(dev_i32(qlen)),
                                   ptr(dev_i32(pos0)), ptr(dev_i32(ptrow)), ptr(out), len(tok0), nh, nkv, d, ps, None))
    sync()
    row, worst, scale = 0, 0.0, 0.0
    for s in range(2):
        pages = torch.as_tensor(pt[s, :pages_per[s]].astype(np.int64), device=DEV)
        kv = pool[pages].float()
        K = kv[:, 0].permute(1, 0, 2, 3).reshape(nkv, -1, d)[:, :lens[s]]
        V = kv[:, 1].permute(1, 0, 2, 3).reshape(nkv, -1, d)[:, :lens[s]]
        n_q = lens[s] - starts[s]
        qs = q[row:row + n_q].float().reshape(n_q, nkv, nh // nkv, d).permute(1, 2, 0, 3)       # [nkv, G, n_q, d]
        sc = qs @ K.transpose(1, 2).unsqueeze(1) / d ** 0.5                                     # [nkv, G, n_q, keys]
        keys = torch.arange(lens[s], device=DEV)[None, :]
        qpos = torch.arange(starts[s], lens[s], device=DEV)[:, None]
        sc = sc.masked_fill(keys > qpos, float("-inf"))
        ref = (torch.softmax(sc, dim=-1) @ V.unsqueeze(1)).permute(2, 0, 1, 3).reshape(n_q, nh * d)
        worst = max(worst, (out[row:row + n_q].float() - ref).abs().max().item())
        scale = max(scale, ref.abs().max().item())
        row += n_q
    assert worst <= 2 * BF16_ULP * scale + 1e-5, (worst, scale)


def test_sampler_argmax_and_ties():
    rng = np.random.default_rng(0)
    T, V = 6, 128256
    lg = rng.standard_normal((T, V)).astype(np.float32)
    lg[1, [500, 90000]] = 9.0      # tie -> lowest index
    lg[2, V - 1] = 11.0            # last element
    lg[3, 0] = 11.0                # first element
    out = G.zeros(T, dtype=torch.int32, device=DEV)
    lp = G.zeros(T, dtype=torch.float32, device=DEV)
    check(lib().dis_k_sample(ptr(dev_f32(lg)), V, T, V, None, None, None, None, ptr(out), ptr(lp), None))
    sync()
    ref = [O.argmax(lg[t])[0] for t in range(T)]
    assert out.cpu().tolist() == ref
    assert ref[1] == 500
    # TokenEvent.logprob (models.rs:276): log softmax(logits)[token], against float64 numpy
    z = lg.astype(np.float64)
    want = z[np.arange(T), ref] - (np.log(np.exp(z - z.max(1, keepdims=True)).sum(1)) + z.max(1))
    assert np.abs(lp.cpu().numpy() - want).max() <= 1e-4


@pytest.mark.parametrize("T,V", [(48, 4099), (24, 128256)], ids=["ragged-vocab", "llama3-vocab"])
def test_sampler_top_p_matches_oracle(T, V):
    rng = np.random.default_rng(1)
    lg = (rng.standard_normal((T, V)) * 2.5).astype(np.float32)
    temps = rng.choice([0.0, 0.5, 1.0, 1.7], T).astype(np.float32)
    tps = rng.choice([0.1, 0.5, 0.9, 0.95, 1.0], T).astype(np.float32)
    seeds = rng.integers(0, 2 ** 40, T).astype(np.uint64)
    steps = rng.integers(0, 100, T).astype(np.uint64)
    out = G.zeros(T, dtype=torch.int32, device=DEV)
    ds = torch.from_numpy(seeds.view(np.int64)).to(DEV)
    dst = torch.from_numpy(steps.view(np.int64)).to(DEV)
    lp = G.zeros(T, dtype=torch.float32, device=DEV)
    check(lib().dis_k_sample(ptr(dev_f32(lg)), V, T, V, ptr(dev_f32(temps)), ptr(dev_f32(tps)), ptr(ds), ptr(dst), ptr(out), ptr(lp), None))
    sync()
    got = out.cpu().tolist()
    # log-probability of the drawn token under softmax(logits / temperature) (temperature 0: the plain softmax)
    for t in range(T):
        z = lg[t].astype(np.float64) / (float(temps[t]) if temps[t] > 0 else 1.0)
        want = z[got[t]] - (np.log(np.exp(z - z.max()).sum()) + z.max())
        assert abs(float(lp[t]) - want) <= 1e-3 * max(1.0, abs(want)), (t, float(lp[t]), want)
    mism = 0
    for t in range(T):
        ref, tau, mass = O.sample(lg[t], float(temps[t]), float(tps[t]), int(seeds[t]), int(steps[t]))
        if got[t] != ref:
            mism += 1
            # only a draw that lands within fp32 noise of a bucket edge may differ
            z = lg[t] / temps[t]; p = np.exp(z - z.max()); p /= p.sum()
            kept = p >= tau * (1 - 1e-5)
            assert kept[got[t]], (t, got[t], ref)
    assert mism <= 2, mism
    # greedy rows are exact
    for t in range(T):
        if temps[t] == 0:
            assert got[t] == int(np.argmax(lg[t]))


@pytest.mark.parametrize("tp,T,V", [(1, 48, 4099), (1, 32, 128256), (2, 48, 4100), (3, 20, 4098), (4, 24, 128256), (8, 16, 128256)],
                         ids=["single-rank-ragged-vocab", "single-rank-llama3-vocab", "tp2-small-vocab", "tp3-small-vocab",
                              "tp4-llama3-vocab", "tp8-llama3-vocab"])
def test_vocab_parallel_sampler_matches_the_one_gpu_sampler_and_the_oracle(tp, T, V):
    """The tensor-parallel engines' sampler (k_tp_sample.cu) with its ranks emulated on one GPU: every rank keeps V/tp
    logit columns and exchanges scalars (candidates, 16 partial masses per round of the threshold search, the drawn
    token).  All ranks must write the same token and log-probability; greedy rows are exact; nucleus rows must draw what
    dis_k_sample and the oracle draw, except where the draw lands within fp32 noise of a bucket edge (the partial sums
    are associated per rank).  Two calls on the same exchange area: the second must not see the first one's words."""
    rng = np.random.default_rng(3 + tp)
    rows = T + 3                                                   # area rows > T: the row stride is the area's, not the call's
    n_area = tp * lib().dis_k_sample_vocab_parallel_area_bytes(tp, rows) // 4
    area = G.zeros(n_area, dtype=torch.int32, device=DEV)
    for call_no in (1, 2):
        lg = (rng.standard_normal((T, V)) * 2.5).astype(np.float32)
        lg[0, [V // tp - 1, (V // tp) % V]] = 40.0                 # a tie across a slice boundary: the lower id wins
        temps = rng.choice([0.0, 0.5, 1.0, 1.7], T).astype(np.float32)
        temps[0] = 0.0
        tps = rng.choice([0.1, 0.5, 0.9, 0.95, 1.0], T).astype(np.float32)
        seeds = rng.integers(0, 2 ** 40, T).astype(np.uint64)
        steps = rng.integers(0, 100, T).astype(np.uint64)
        dl, dt, dp = dev_f32(lg), dev_f32(temps), dev_f32(tps)
        ds = torch.from_numpy(seeds.view(np.int64)).to(DEV)
        dst = torch.from_numpy(steps.view(np.int64)).to(DEV)
        out = G.zeros((tp, T), dtype=torch.int32, device=DEV)
        lp = G.zeros((tp, T), dtype=torch.float32, device=DEV)
        check(lib().dis_k_sample_vocab_parallel(ptr(dl), T, V, tp, ptr(dt), ptr(dp), ptr(ds), ptr(dst), call_no, ptr(area), rows,
                                                ptr(out), ptr(lp), None))
        one = G.zeros(T, dtype=torch.int32, device=DEV)
        one_lp = G.zeros(T, dtype=torch.float32, device=DEV)
        check(lib().dis_k_sample(ptr(dl), V, T, V, ptr(dt), ptr(dp), ptr(ds), ptr(dst), ptr(one), ptr(one_lp), None))
        sync()
        got_all, lp_all = out.cpu().numpy(), lp.cpu().numpy()
        for r in range(1, tp):                                     # every rank computes the same bits
            assert np.array_equal(got_all[0], got_all[r]) and np.array_equal(lp_all[0], lp_all[r]), r
        got, single = got_all[0].tolist(), one.cpu().tolist()
        assert got[0] == (V // tp - 1 if tp > 1 else 0)
        mism_single = mism_oracle = 0
        for t in range(T):
            z = lg[t].astype(np.float64) / (float(temps[t]) if temps[t] > 0 else 1.0)
            want_lp = z[got[t]] - (np.log(np.exp(z - z.max()).sum()) + z.max())
            assert abs(float(lp_all[0, t]) - want_lp) <= 1e-3 * max(1.0, abs(want_lp)), (t, float(lp_all[0, t]), want_lp)
            if temps[t] == 0:
                assert got[t] == int(np.argmax(lg[t])) == single[t]
                continue
            ref, tau, mass = O.sample(lg[t], float(temps[t]), float(tps[t]), int(seeds[t]), int(steps[t]))
            if got[t] != ref or got[t] != single[t]:
                mism_oracle += got[t] != ref
                mism_single += got[t] != single[t]
                p = np.exp(z - z.max()); p /= p.sum()
                assert (p >= tau * (1 - 1e-5))[got[t]], (t, got[t], ref, single[t])   # still inside the nucleus
        assert mism_oracle <= 2 and mism_single <= 2, (mism_oracle, mism_single)
