"""Megatron tensor-parallel shard plan of the decoder, as the engine applies it (csrc/engine.cu,
Engine::tensor_slot).  The reference specifies whole-model replicas only (requirements.md:100-110);
tensor parallelism is north_star's addition, so this plan has no reference counterpart.

  Wq/Wk/Wv  column-parallel: heads split (n_heads/tp query heads, n_kv_heads/tp kv heads per rank)
  Wo        row-parallel   : input columns of the local heads   -> all-reduce(sum) of [T, H]
  Wgate/Wup column-parallel: inter/tp rows
  Wdown     row-parallel   : inter/tp input columns              -> all-reduce(sum) of [T, H]
  lm_head   vocab-parallel : vocab/tp rows; greedy = local (max, id) -> all-gather -> merge
  embed, norms             : replicated
"""
from __future__ import annotations

from . import binding as B


def check(shape, tp_size: int) -> None:
    if shape.n_kv_heads % tp_size or shape.n_heads % tp_size or shape.inter % tp_size or shape.vocab % tp_size:
        raise ValueError("n_heads, n_kv_heads, inter and vocab must be divisible by tp_size")


def shard_slices(shape, kind: int, tp_rank: int, tp_size: int):
    """(row slice, column slice) of the full [out, in] tensor that rank `tp_rank` keeps."""
    check(shape, tp_size)
    d = shape.head_dim
    qd_l, kd_l = shape.n_heads // tp_size * d, shape.n_kv_heads // tp_size * d
    i_l, v_l = shape.inter // tp_size, shape.vocab // tp_size
    full = slice(None)
    r = tp_rank
    if kind == B.T_WQ:
        return slice(r * qd_l, (r + 1) * qd_l), full
    if kind in (B.T_WK, B.T_WV):
        return slice(r * kd_l, (r + 1) * kd_l), full
    if kind == B.T_WO:
        return full, slice(r * qd_l, (r + 1) * qd_l)
    if kind in (B.T_WGATE, B.T_WUP):
        return slice(r * i_l, (r + 1) * i_l), full
    if kind == B.T_WDOWN:
        return full, slice(r * i_l, (r + 1) * i_l)
    if kind == B.T_LM_HEAD:
        return slice(r * v_l, (r + 1) * v_l), full
    if kind in (B.T_EMBED, B.T_ATTN_NORM, B.T_MLP_NORM, B.T_FINAL_NORM):
        return full, full
    raise ValueError(f"unknown tensor kind {kind}")


def vocab_offset(shape, tp_rank: int, tp_size: int) -> int:
    return tp_rank * (shape.vocab // tp_size)


def merge_greedy(cand_val, cand_tok):
    """Final step of the vocab-parallel argmax: [tp][T] local maxima and their global ids ->
    [T] token ids; the larger value wins, the lower id on a tie (same rule as the single-GPU sampler)."""
    tp, T = len(cand_val), len(cand_val[0])
    out = []
    for t in range(T):
        bv, bt = cand_val[0][t], cand_tok[0][t]
        for r in range(1, tp):
            v, k = cand_val[r][t], cand_tok[r][t]
            if v > bv or (v == bv and k < bt):
                bv, bt = v, k
        out.append(int(bt))
    return out
