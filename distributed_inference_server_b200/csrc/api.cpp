// api.cpp -- extern "C" surface declared in include/dis_engine.h.
#include <cstring>
#include <new>

#include "engine.h"

using dis::Engine;

struct dis_engine {
  Engine eng;
};

static int guard_null(dis_engine *e) {
  if (!e) { dis::set_last_error("null engine handle"); return DIS_ERR_INVALID_ARGUMENT; }
  return DIS_OK;
}
#define LOCKED(e) std::lock_guard<std::mutex> _lk((e)->eng.mu)

extern "C" {

int dis_abi_version(void) { return DIS_ABI_VERSION; }
const char *dis_last_error(void) { return dis::get_last_error(); }
const char *dis_engine_last_error(dis_engine *e) { return e ? e->eng.last_error.c_str() : dis::get_last_error(); }

int dis_engine_create(const dis_engine_config *cfg, dis_engine **out) {
  if (!cfg || !out) { dis::set_last_error("dis_engine_create: null argument"); return DIS_ERR_INVALID_ARGUMENT; }
  *out = nullptr;
  dis_engine *e = new (std::nothrow) dis_engine();
  if (!e) { dis::set_last_error("host allocation failed"); return DIS_ERR_OUT_OF_MEMORY; }
  const int rc = e->eng.init(cfg);
  if (rc != DIS_OK) { std::string msg = e->eng.last_error; delete e; dis::set_last_error(msg); return rc; }
  *out = e;
  return DIS_OK;
}

int dis_engine_destroy(dis_engine *e) {
  if (int rc = guard_null(e)) return rc;
  { LOCKED(e); e->eng.shut = true; }
  delete e;
  return DIS_OK;
}

int dis_engine_load_tensor(dis_engine *e, int kind, int layer, const void *host, int dtype) {
  if (int rc = guard_null(e)) return rc;
  LOCKED(e);
  return e->eng.load_tensor(kind, layer, host, dtype);
}
int dis_engine_load_synthetic(dis_engine *e, uint64_t seed) {
  if (int rc = guard_null(e)) return rc;
  LOCKED(e);
  return e->eng.load_synthetic(seed);
}
int dis_engine_read_tensor(dis_engine *e, int kind, int layer, uint16_t *host, size_t *n_elems) {
  if (int rc = guard_null(e)) return rc;
  LOCKED(e);
  return e->eng.read_tensor(kind, layer, host, n_elems);
}
int dis_engine_finalize(dis_engine *e) {
  if (int rc = guard_null(e)) return rc;
  LOCKED(e);
  return e->eng.finalize();
}
int dis_engine_seq_add(dis_engine *e, const uint32_t *prompt, int32_t n, const dis_sampling_params *p, int32_t *slot) {
  if (int rc = guard_null(e)) return rc;
  LOCKED(e);
  return e->eng.seq_add(prompt, n, p, slot);
}
int dis_engine_seq_remove(dis_engine *e, int32_t slot) {
  if (int rc = guard_null(e)) return rc;
  LOCKED(e);
  return e->eng.seq_remove(slot);
}
int dis_engine_prefill(dis_engine *e, int32_t n, const int32_t *slots, uint32_t *out_tokens, int32_t *out_finish) {
  if (int rc = guard_null(e)) return rc;
  LOCKED(e);
  return e->eng.prefill(n, slots, out_tokens, out_finish);
}
int dis_engine_decode(dis_engine *e, int32_t n, const int32_t *slots, uint32_t *out_tokens, int32_t *out_finish) {
  if (int rc = guard_null(e)) return rc;
  LOCKED(e);
  return e->eng.decode(n, slots, out_tokens, out_finish);
}
int dis_engine_step(dis_engine *e, int32_t n, const int32_t *slots, int32_t max_prefill_tokens, uint32_t *out_tokens,
                    int32_t *out_finish, int32_t *out_produced) {
  if (int rc = guard_null(e)) return rc;
  LOCKED(e);
  return e->eng.step_mixed(n, slots, max_prefill_tokens, out_tokens, out_finish, out_produced);
}
int dis_serving_plan_step(int32_t n, const int32_t *remaining_prompt, int32_t budget, int32_t cap, int32_t *take) {
  if (n < 0 || (n > 0 && (!remaining_prompt || !take))) return DIS_ERR_INVALID_ARGUMENT;
  dis::plan_mixed_step(n, remaining_prompt, budget, cap, take);
  return DIS_OK;
}
int dis_engine_profile_decode(dis_engine *e, int32_t n, const int32_t *slots, uint32_t *out_tokens, int32_t *out_finish,
                              float *class_ms, int32_t *class_launches) {
  if (int rc = guard_null(e)) return rc;
  LOCKED(e);
  return e->eng.profile_decode(n, slots, out_tokens, out_finish, class_ms, class_launches);
}
int dis_engine_replay_last_step(dis_engine *e, int32_t n, float *avg_ms) {
  if (int rc = guard_null(e)) return rc;
  LOCKED(e);
  return e->eng.replay_last_step(n, avg_ms);
}
int dis_engine_last_logits(dis_engine *e, float *host, int32_t n_rows) {
  if (int rc = guard_null(e)) return rc;
  LOCKED(e);
  return e->eng.last_logits_host(host, n_rows);
}
int dis_engine_last_logprobs(dis_engine *e, int32_t n, const int32_t *slots, float *out) {
  if (int rc = guard_null(e)) return rc;
  LOCKED(e);
  return e->eng.last_logprobs(n, slots, out);
}
int dis_engine_status(dis_engine *e, dis_worker_status *out) {
  if (int rc = guard_null(e)) return rc;
  LOCKED(e);
  return e->eng.status(out);
}
int dis_engine_prefix_stats(dis_engine *e, dis_prefix_stats *out) {
  if (int rc = guard_null(e)) return rc;
  LOCKED(e);
  return e->eng.prefix_stats(out);
}
int dis_engine_model_info(dis_engine *e, dis_model_config *out) {
  if (int rc = guard_null(e)) return rc;
  if (!out) return DIS_ERR_INVALID_ARGUMENT;
  *out = e->eng.m;
  return DIS_OK;
}

int dis_tp_unique_id(uint8_t out[128]) {
  std::string err;
  dis::NcclApi *api = dis::nccl_api(&err);
  if (!api) { dis::set_last_error("tensor parallel needs NCCL: " + err); return DIS_ERR_INFERENCE_FAILED; }
  dis::NcclId id;
  memset(&id, 0, sizeof(id));
  const int r = api->GetUniqueId(&id);
  if (r != 0) { dis::set_last_error(std::string("ncclGetUniqueId: ") + api->GetErrorString(r)); return DIS_ERR_INFERENCE_FAILED; }
  memcpy(out, id.bytes, 128);
  return DIS_OK;
}

int dis_engine_tp_export(dis_engine *e, uint8_t out[64]) {
  if (int rc = guard_null(e)) return rc;
  LOCKED(e);
  return e->eng.tp_export(out);
}
int dis_engine_tp_import(dis_engine *e, const uint8_t *handles) {
  if (int rc = guard_null(e)) return rc;
  LOCKED(e);
  return e->eng.tp_import(handles);
}

// ---- kernel-level entry points -------------------------------------------------------------------
static int k_ret(cudaError_t e) { return e == cudaSuccess ? DIS_OK : DIS_ERR_INFERENCE_FAILED; }
static int need_gpu() {
  int n = 0;
  if (cudaGetDeviceCount(&n) != cudaSuccess || n == 0) { dis::set_last_error("no CUDA device: there is no CPU fallback"); return DIS_ERR_INFERENCE_FAILED; }
  return DIS_OK;
}

int dis_k_add_rmsnorm(float *x, const uint32_t *tokens, const uint16_t *embed, const float *partial, int32_t n_partial,
                      int64_t partial_stride, const float *gain, uint16_t *xn, const int32_t *row_map, int32_t T,
                      int32_t H, float eps, void *stream) {
  if (int rc = need_gpu()) return rc;
  return k_ret(dis::launch_add_rmsnorm(x, tokens, embed, partial, n_partial, partial_stride, gain, xn, row_map, T, H, eps,
                                       (cudaStream_t)stream));
}

static int gemm_common(const uint16_t *X, const uint16_t *W, float *out, uint16_t *act, int T, int N, int K, int split,
                       bool silu, cudaStream_t st) {
  if (int rc = need_gpu()) return rc;
  dis::TmaMap wm, xm;
  const bool persist = split <= 1 && dis::gemm_use_persist(T, N, silu);  // the engine's choice for this shape
  const int bn = persist ? 256 : dis::gemm_token_tile(T);
  const int rows = silu ? (persist ? 64 : dis::gemm_pick_tile_rows(N)) : 128;
  cudaError_t e = dis::make_tma_map_2d(&wm, W, (uint64_t)(silu ? 2 * N : N), (uint64_t)K, (uint32_t)rows);
  if (e == cudaSuccess) e = dis::make_tma_map_2d(&xm, X, (uint64_t)T, (uint64_t)K, (uint32_t)bn);
  if (e != cudaSuccess) return k_ret(e);
  if (persist && silu) return k_ret(dis::launch_gemm_persist_silu(wm, xm, act, T, N, K, 0, st));
  if (persist) return k_ret(dis::launch_gemm_persist_partial(wm, xm, out, T, N, K, 0, st));
  if (silu) return k_ret(dis::launch_gemm_silu(wm, xm, act, T, N, K, 0, rows, st));
  return k_ret(dis::launch_gemm_partial(wm, xm, out, T, N, K, 0, split, st));
}
int dis_k_gemm(const uint16_t *X, const uint16_t *W, float *out, int32_t T, int32_t N, int32_t K, int32_t split_k,
               void *stream) {
  return gemm_common(X, W, out, nullptr, T, N, K, split_k, false, (cudaStream_t)stream);
}
int dis_k_gemm_pick_split(int32_t T, int32_t N, int32_t K) { return dis::gemm_pick_split_k(T, N, K); }
int dis_k_gemm_silu(const uint16_t *X, const uint16_t *Wgu, uint16_t *act, int32_t T, int32_t I, int32_t K, void *stream) {
  return gemm_common(X, Wgu, nullptr, act, T, I, K, 1, true, (cudaStream_t)stream);
}
int dis_k_gemv(const uint16_t *X, const uint16_t *W, float *out, int32_t T, int32_t N, int32_t K, void *stream) {
  if (int rc = need_gpu()) return rc;
  return k_ret(dis::launch_gemv(X, W, out, T, N, K, (cudaStream_t)stream));
}
int dis_k_gemv_silu(const uint16_t *X, const uint16_t *Wgu, uint16_t *act, int32_t T, int32_t I, int32_t K, void *stream) {
  if (int rc = need_gpu()) return rc;
  return k_ret(dis::launch_gemv_silu(X, Wgu, act, T, I, K, (cudaStream_t)stream));
}
int dis_k_rope_kv_write(const float *qkv, int32_t n_partial, int64_t partial_stride, uint16_t *q, uint16_t *kv_pool,
                        const int32_t *page_table, int32_t max_pages, const int32_t *pt_row, const int32_t *pos,
                        const float *rope_cos, const float *rope_sin, int32_t T, int32_t n_heads, int32_t n_kv, int32_t d,
                        int32_t page_size, void *stream) {
  if (int rc = need_gpu()) return rc;
  return k_ret(dis::launch_rope_kv_write(qkv, n_partial, partial_stride, q, kv_pool, page_table, max_pages, pt_row, pos,
                                         rope_cos, rope_sin, T, n_heads, n_kv, d, page_size, (cudaStream_t)stream));
}
size_t dis_k_attn_scratch_elems(int32_t T, int32_t n_heads, int32_t d, int32_t n_splits) {
  return dis::attn_scratch_elems(T, n_heads, d, n_splits);
}
int dis_k_attn_paged(const uint16_t *q, const uint16_t *kv_pool, int64_t n_pages, const int32_t *page_table,
                     int32_t max_pages, const int32_t *pt_row, const int32_t *ctx, uint16_t *out, float *scratch,
                     int32_t n_splits, int32_t T, int32_t n_heads, int32_t n_kv, int32_t d, int32_t page_size, void *stream) {
  if (int rc = need_gpu()) return rc;
  dis::TmaMap map;
  cudaError_t e = dis::make_kv_map(&map, kv_pool, (long long)n_pages * 2 * n_kv * page_size, d, page_size);
  if (e != cudaSuccess) return k_ret(e);
  return k_ret(dis::launch_attn_paged(map, 0, q, page_table, max_pages, pt_row, ctx, out, scratch, n_splits, T, n_heads, n_kv,
                                      d, page_size, (cudaStream_t)stream));
}
int dis_k_attn_prefill(const uint16_t *q, const uint16_t *kv_pool, int64_t n_pages, const int32_t *page_table,
                       int32_t max_pages, const int32_t *qb_tok0, const int32_t *qb_len, const int32_t *qb_pos0,
                       const int32_t *qb_ptrow, uint16_t *out, int32_t n_blocks, int32_t n_heads, int32_t n_kv, int32_t d,
                       int32_t page_size, void *stream) {
  if (int rc = need_gpu()) return rc;
  dis::TmaMap map;
  cudaError_t e = dis::make_kv_map(&map, kv_pool, (long long)n_pages * 2 * n_kv * page_size, d, page_size);
  if (e != cudaSuccess) return k_ret(e);
  auto fn = dis::attn_prefill_uses_tc(n_heads, n_kv, d) ? dis::launch_attn_prefill_tc : dis::launch_attn_prefill;
  return k_ret(fn(map, 0, q, page_table, max_pages, qb_tok0, qb_len, qb_pos0, qb_ptrow, out, n_blocks, n_heads, n_kv, d, page_size,
                  (cudaStream_t)stream));
}
int dis_k_attn_prefill_block_tokens(int32_t n_heads, int32_t n_kv, int32_t d) {
  return dis::attn_prefill_block_tokens(n_heads, n_kv, d);
}
int dis_k_sample(const float *logits, int64_t ld, int32_t T, int32_t V, const float *temperature, const float *top_p,
                 const uint64_t *seeds, const uint64_t *steps, uint32_t *out_tokens, float *out_logprob, void *stream) {
  if (int rc = need_gpu()) return rc;
  return k_ret(dis::launch_sample(logits, ld, T, V, temperature, top_p, reinterpret_cast<const unsigned long long *>(seeds),
                                  reinterpret_cast<const unsigned long long *>(steps), out_tokens, 0, nullptr, nullptr,
                                  out_logprob, (cudaStream_t)stream));
}
size_t dis_k_sample_vocab_parallel_area_bytes(int32_t tp, int32_t rows) {
  return tp > 0 && rows > 0 ? dis::tp_sample_area_bytes(tp, rows) : 0;
}
int dis_k_sample_vocab_parallel(const float *logits, int32_t T, int32_t V, int32_t tp, const float *temperature,
                                const float *top_p, const uint64_t *seeds, const uint64_t *steps, uint32_t call_no,
                                void *area, int32_t area_rows, uint32_t *out_tokens, float *out_logprob, void *stream) {
  if (int rc = need_gpu()) return rc;
  if (!logits || !area || !out_tokens || tp < 1 || tp > dis::kTpMaxRanks || V <= 0 || V % tp != 0) {
    dis::set_last_error("sample_vocab_parallel: 1 <= tp <= 8, the vocabulary a multiple of tp, logits / area / out_tokens set");
    return DIS_ERR_INVALID_ARGUMENT;
  }
  dis::TpSampPeers sp{};
  for (int r = 0; r < tp; r++) sp.buf[r] = static_cast<uint32_t *>(area) + (size_t)r * (dis::tp_sample_area_bytes(tp, area_rows) / 4);
  sp.rank = -1;  // every rank is a slice of this grid
  sp.tp = tp;
  sp.cap_rows = area_rows;
  return k_ret(dis::launch_tp_sample(sp, logits, V, T, V / tp, temperature, top_p, reinterpret_cast<const unsigned long long *>(seeds),
                                     reinterpret_cast<const unsigned long long *>(steps), nullptr, call_no, out_tokens, out_logprob,
                                     (cudaStream_t)stream));
}
int dis_k_synth_bf16(uint16_t *dst, size_t n, uint64_t seed, int32_t kind, int32_t layer, float amp, float offset,
                     void *stream) {
  if (int rc = need_gpu()) return rc;
  return k_ret(dis::launch_synth_bf16(dst, n, seed, kind, layer, amp, offset, 0, n, n, 0, (cudaStream_t)stream));
}

}  // extern "C"
