// kernels.h -- host-side launchers of the sm_100a kernels (internal; the C ABI is include/dis_engine.h).
#pragma once
#include "common.cuh"

namespace dis {

// A 2-D bf16 row-major [rows][cols] tensor map with a {64, box_rows} box and 128-byte swizzle.
struct TmaMap { alignas(64) unsigned char bytes[128]; };

// k_elementwise.cu
cudaError_t launch_add_rmsnorm(float *x, const uint32_t *tokens, const uint16_t *embed, const float *partial,
                               int n_partial, long long partial_stride, const float *gain, uint16_t *xn,
                               const int *row_map, int T, int H, float eps, cudaStream_t st);
cudaError_t launch_reduce_partials(const float *partial, int n_partial, long long stride, float *out, long long n,
                                   cudaStream_t st);
cudaError_t launch_silu_mul_reduce(const float *partial, int n_partial, long long stride, uint16_t *act, int T, int I,
                                   cudaStream_t st);
cudaError_t launch_rope_kv_write(const float *qkv, int n_partial, long long partial_stride, uint16_t *q,
                                 uint16_t *kv_pool, const int *page_table, int max_pages, const int *pt_row,
                                 const int *pos, const float *rope_cos, const float *rope_sin, int T,
                                 int n_heads, int n_kv, int d, int page_size, cudaStream_t st);
cudaError_t launch_sample(const float *logits, long long ld, int T, int V, const float *temperature,
                          const float *top_p, const unsigned long long *seeds, const unsigned long long *steps,
                          uint32_t *out_tokens, int vocab_offset, float *out_val, float *out_sumexp, float *out_logprob,
                          cudaStream_t st);
cudaError_t launch_argmax_merge(const float *cand_val, const uint32_t *cand_tok, const float *cand_sum, int tp, int T,
                                uint32_t *out, float *out_logprob, cudaStream_t st);
cudaError_t launch_synth_bf16(uint16_t *dst, size_t n, unsigned long long seed, int kind, int layer, float amp,
                              float offset, size_t idx0, size_t row_len, size_t src_ld, size_t src_col0,
                              cudaStream_t st);
cudaError_t launch_bf16_to_f32(const uint16_t *src, float *dst, size_t n, cudaStream_t st);

// k_tp.cu : exchange over peer memory.  Every rank owns one symmetric allocation:
//   data [2 parities][tp sources][cap_rows][H] fp32 | flags [2 hops][2][tp][cap_rows] u32 | cand [2][tp][cap_rows][4] u32 |
//   cand_flags [2][tp] u32     (peers PUSH into the slot of their own rank; the owner polls and reads locally)
// two_shot: the data area is used as a reduce-scatter area + an all-gather area instead (see k_tp.cu); the
// flag-in-data variant needs 8 * cap_rows * H floats of it.
constexpr int kTpMaxRanks = 8;
struct TpPeers {
  float *data[kTpMaxRanks];
  uint32_t *flags[kTpMaxRanks];
  uint32_t *cand[kTpMaxRanks];
  uint32_t *cand_flags[kTpMaxRanks];
  int rank, tp, cap_rows;
  int two_shot;  // 0: one-shot push all-reduce, 1: reduce-scatter + all-gather over column blocks, 2: the same with flag-in-data stores,
                 // 3: flag-in-data on a (row, column block) grid (falls back to 2 when the grid would not be co-resident),
                 // 4: streaming two-shot for prefill-sized passes (resident CTAs, two flag flights per CTA, fenced layout)
};
cudaError_t launch_tp_add_rmsnorm(const TpPeers &peers, float *x, const float *partial, int n_partial,
                                  long long partial_stride, const float *gain, uint16_t *xn, const uint32_t *step_ptr,
                                  uint32_t idx, int parity, int T, int H, float eps, cudaStream_t st);
cudaError_t launch_tp_argmax_merge(const TpPeers &peers, const float *val, const uint32_t *tok, const float *sumexp,
                                   const uint32_t *step_ptr, int T, uint32_t *out, float *out_logprob, cudaStream_t st);

// k_tp_sample.cu : vocab-parallel sampling (greedy, temperature, nucleus) over peer memory.  buf[r] = rank r's exchange area,
// tp_sample_area_bytes() of zero-initialised memory laid out [round][source rank][cap_rows][kTpSampleVals] x {bits, epoch}.
// rank < 0: the tp ranks are emulated on one GPU (grid (rows, tp); logits is the full [T][tp * Vl] matrix, out_tokens /
// out_logprob are [tp][T]).  step_ptr (device) or step_imm numbers the call: it must differ from the previous call's on
// the same area and be the same on every rank.
constexpr int kTpSampleRounds = 12, kTpSampleVals = 16;
struct TpSampPeers {
  uint32_t *buf[kTpMaxRanks];
  int rank, tp, cap_rows;
};
size_t tp_sample_area_bytes(int tp, int cap_rows);
cudaError_t launch_tp_sample(const TpSampPeers &sp, const float *logits, long long ld, int T, int Vl,
                             const float *temperature, const float *top_p, const unsigned long long *seeds,
                             const unsigned long long *steps, const uint32_t *step_ptr, uint32_t step_imm,
                             uint32_t *out_tokens, float *out_logprob, cudaStream_t st);

// k_gemv.cu : out[t][n] = sum_k X[t][k] W[n][k]  (T <= kGemvMaxT per launch; launcher loops)
constexpr int kGemvMaxT = 4;
cudaError_t launch_gemv(const uint16_t *X, const uint16_t *W, float *out, int T, int N, int K, cudaStream_t st);
cudaError_t launch_gemv_silu(const uint16_t *X, const uint16_t *Wgu, uint16_t *act, int T, int I, int K,
                             cudaStream_t st);

// k_attn_paged.cu
size_t attn_scratch_elems(int T, int n_heads, int d, int n_splits);
// kv_map: make_kv_map over a pool of d-element rows [..][page][2][n_kv][page_size]; row0 = first row of the
// layer to read (the engine keeps one map over all layers).
cudaError_t make_kv_map(TmaMap *out, const uint16_t *pool, long long total_rows, int d, int page_size);
cudaError_t launch_attn_paged(const TmaMap &kv_map, long long row0, const uint16_t *q, const int *page_table,
                              int max_pages, const int *pt_row, const int *ctx, uint16_t *out, float *scratch,
                              int n_splits, int T, int n_heads, int n_kv, int d, int page_size, cudaStream_t st);

cudaError_t launch_attn_decode_fused(const TmaMap &kv_map, long long row0, uint16_t *kv_layer_base, const float *qkv,
                                     int n_partial, long long partial_stride, const int *pos, const float *rope_cos,
                                     const float *rope_sin, const int *page_table, int max_pages, const int *pt_row,
                                     const int *ctx, uint16_t *out, float *scratch, int n_splits, int T, int n_heads,
                                     int n_kv, int d, int page_size, cudaStream_t st);
cudaError_t launch_attn_prefill(const TmaMap &kv_map, long long row0, const uint16_t *q, const int *page_table,
                                int max_pages, const int *qb_tok0, const int *qb_len, const int *qb_pos0,
                                const int *qb_ptrow, uint16_t *out, int n_blocks, int n_heads, int n_kv, int d,
                                int page_size, cudaStream_t st);

// k_attn_prefill_tc.cu : the same contract on tcgen05 / TMEM for head_dim 128 and GQA groups of 4 or 8; query blocks
// hold up to attn_prefill_block_tokens() tokens (128 / G there, 16 on the mma.sync kernel)
bool attn_prefill_uses_tc(int n_heads, int n_kv, int d);
int attn_prefill_block_tokens(int n_heads, int n_kv, int d);
cudaError_t launch_attn_prefill_tc(const TmaMap &kv_map, long long row0, const uint16_t *q, const int *page_table,
                                   int max_pages, const int *qb_tok0, const int *qb_len, const int *qb_pos0,
                                   const int *qb_ptrow, uint16_t *out, int n_blocks, int n_heads, int n_kv, int d,
                                   int page_size, cudaStream_t st);

// k_gemm_tcgen05.cu
cudaError_t make_tma_map_2d(TmaMap *out, const void *base, uint64_t rows, uint64_t cols, uint32_t box_rows);
int gemm_token_tile(int T);  // 32 / 64 / 128: the UMMA N used for T tokens
// out[S][T][N] fp32 partials.  w_row0: first row of W's map to use (stacked per-layer weights).
cudaError_t launch_gemm_partial(const TmaMap &w_map, const TmaMap &x_map, float *out, int T, int N, int K,
                                int w_row0, int split_k, cudaStream_t st);
// act[T][I] = bf16(silu(gate) * up); gate rows at w_row0, up rows at w_row0 + I.  w_map's box height must be
// tile_rows = gemm_pick_tile_rows(I).
int gemm_pick_tile_rows(int N);
// k_gemm_persist.cu : persistent, TMEM-double-buffered variant for prefill-sized T (no split-K)
bool gemm_use_persist(int T, int N, bool dual);
cudaError_t launch_gemm_persist_partial(const TmaMap &w_map128, const TmaMap &x_map256, float *out, int T, int N, int K,
                                        int w_row0, cudaStream_t st);
cudaError_t launch_gemm_persist_silu(const TmaMap &w_map64, const TmaMap &x_map256, uint16_t *act, int T, int I, int K,
                                     int w_row0, cudaStream_t st);
cudaError_t launch_gemm_silu(const TmaMap &w_map, const TmaMap &x_map, uint16_t *act, int T, int I, int K,
                             int w_row0, int tile_rows, cudaStream_t st);
int gemm_pick_split_k(int T, int N, int K);

}  // namespace dis
