// engine.h -- the decoder engine behind include/dis_engine.h (internal C++).
#pragma once

#include <map>
#include <mutex>
#include <string>
#include <vector>

#include "../../include/dis_engine.h"
#include "kernels.h"
#include "prefix_index.hpp"
#include "ptx.cuh"

namespace dis {

// tp.cpp: NCCL entry points resolved with dlopen (no link-time dependency; reuses the copy a host
// process such as PyTorch has already loaded).
struct NcclId { char bytes[128]; };
struct NcclApi {
  void *lib = nullptr;
  int (*GetUniqueId)(NcclId *) = nullptr;
  int (*CommInitRank)(void **, int, NcclId, int) = nullptr;
  int (*CommDestroy)(void *) = nullptr;
  int (*AllReduce)(const void *, void *, size_t, int, int, void *, cudaStream_t) = nullptr;
  int (*AllGather)(const void *, void *, size_t, int, void *, cudaStream_t) = nullptr;
  const char *(*GetErrorString)(int) = nullptr;
};
NcclApi *nccl_api(std::string *err);

struct Sequence {
  bool used = false;
  bool prefilled = false;
  std::vector<uint32_t> prompt;
  std::vector<uint32_t> gen;  // sampled tokens (kept only while the prefix cache is on)
  int cached = 0;          // leading prompt tokens whose K/V came from the prefix cache
  int prefill_pos = 0;     // prompt tokens whose K/V are in the cache already (prefix hit + chunks run so far)
  int len = 0;             // tokens whose K/V are in the cache
  uint32_t last_token = 0;  // sampled, not yet in the cache
  float last_logprob = 0.f; // log-probability of last_token (TokenEvent.logprob)
  int generated = 0;
  int finish = DIS_FINISH_NONE;
  int32_t max_tokens = 0;
  float temperature = 0.f, top_p = 1.f;
  uint64_t seed = 0;
  std::vector<std::vector<uint32_t>> stop;  // stop sequences
  std::vector<uint32_t> tail;               // last generated tokens, for stop matching
  std::vector<int> pages;
};

// Host-side image of the per-step device metadata (one pinned block, one H2D copy per step).
struct StepMeta {
  int cap_tokens = 0, cap_rows = 0;
  size_t bytes = 0;
  uint8_t *host = nullptr;  // pinned
  uint8_t *dev = nullptr;
  // sections (offsets in bytes)
  size_t o_tokens, o_pos, o_ptrow, o_ctx, o_rowmap, o_temp, o_topp, o_seed, o_step;
  size_t o_qb_tok0, o_qb_len, o_qb_pos0, o_qb_ptrow;  // prefill query blocks
  size_t o_arstep;                                     // forward-pass counter (epoch of the peer-memory exchange)
  template <class T_> T_ *h(size_t off) { return reinterpret_cast<T_ *>(host + off); }
  template <class T_> T_ *d(size_t off) { return reinterpret_cast<T_ *>(dev + off); }
};

void plan_mixed_step(int n, const int *remaining_prompt, int budget, int cap, int *take);

struct Engine {
  dis_engine_config cfg{};
  dis_model_config m{};
  int tp_rank = 0, tp_size = 1;
  // local (per tensor-parallel rank) dimensions
  int nh_l = 0, nkv_l = 0, qd_l = 0, kvd_l = 0, nqkv_l = 0, inter_l = 0, vocab_l = 0, vocab_off = 0;
  int max_batch = 0, cap_tokens = 0, page_size = 0, max_pages = 0;
  int64_t n_pages = 0;
  cudaStream_t stream = nullptr;
  bool own_stream = false;
  bool loaded = false, shut = false, healthy = true;
  std::mutex mu;
  std::string last_error;
  uint64_t bytes_allocated = 0, tokens_generated = 0, kernel_launches = 0;
  double last_step_ms = 0.0;
  cudaEvent_t ev0 = nullptr, ev1 = nullptr;

  // weights
  uint16_t *embed = nullptr, *wqkv = nullptr, *wo = nullptr, *wgu = nullptr, *wdown = nullptr, *lm_head = nullptr;
  float *attn_norm = nullptr, *mlp_norm = nullptr, *final_norm = nullptr;
  TmaMap wqkv_map, wo_map, wgu_map, wgu_map128, wgu_map64, wdown_map, lm_map;  // wgu_map: balanced tile height, wgu_map128 / 64: 128- / 64-row boxes
  // activations
  float *x = nullptr, *partial = nullptr, *logits = nullptr, *ar_buf = nullptr, *attn_scratch = nullptr;
  float *full_logits = nullptr, *gather_buf = nullptr, *cand_val = nullptr, *cand_val_all = nullptr;
  uint32_t *cand_tok = nullptr, *cand_tok_all = nullptr;
  float *cand_sum = nullptr, *cand_sum_all = nullptr;  // sum exp(z - max) over the rank's vocabulary slice (logprob)
  uint16_t *xn = nullptr, *q = nullptr, *attn = nullptr, *act = nullptr;
  size_t partial_elems = 0, attn_scratch_elems = 0;
  TmaMap xn_map[4], attn_map[4], act_map[4];  // box rows 32 / 64 / 128 / 256
  uint32_t *d_out_tokens = nullptr, *h_out_tokens = nullptr;
  // KV
  uint16_t *kv_pool = nullptr;
  size_t kv_layer_stride = 0;  // elements between layers in the pool
  long long kv_rows_per_layer = 0;
  TmaMap kv_map;  // the whole pool as [rows][head_dim]
  bool use_pdl = true;  // programmatic dependent launch between the kernels of a step
  bool use_persist = true;  // prefill-sized projections through the persistent GEMM (DIS_GEMM_PERSIST=0: A/B runs)
  int max_attn_splits = 0;  // DIS_ATTN_MAX_SPLITS: cap on the split-KV factor of the decode attention (0 = none)
  unsigned skip_mask = 0;  // DIS_DEBUG_SKIP (measurement only): kernel classes left out of the step, bit = ProfClass
  int gemv_max_t = 0;   // steps of at most this many tokens take the GEMV path.  0: the tcgen05 path is faster at every T (T = 1: 3.06 vs 3.18 ms per step, T = 2: 3.03 vs 4.09); DIS_GEMV_MAX_T / DIS_FLAG_FORCE_GEMV select the GEMV
  PrefixIndex pages_ix;  // page allocator; with DIS_FLAG_PREFIX_CACHE also the prefix index (row f2)
  int *d_page_table = nullptr, *h_page_table = nullptr;
  bool pt_dirty = true;
  float *rope_cos = nullptr, *rope_sin = nullptr;
  StepMeta meta_dec, meta_pre;      // decode steps / prefill chunks
  StepMeta *cur_meta = &meta_dec;
  int n_qblocks = 0;                // > 0 while a prefill chunk is being enqueued
  int pf_block = 16;                // tokens per prefill query block (attn_prefill_block_tokens)
  bool pf_tc = false;               // prefill attention on the tcgen05 kernel (head_dim 128, GQA group 4 or 8)
  std::vector<Sequence> seqs;
  std::vector<uint64_t> slot_stamp;  // duplicate-slot detection in O(n): stamp of the last call that listed the slot
  uint64_t call_stamp = 0;
  int last_rows = 0;  // logits rows of the last call
  // decode graphs keyed by (T, n_splits, all_greedy)
  std::map<uint64_t, cudaGraphExec_t> graphs;
  std::map<uint64_t, uint64_t> graph_launches;  // kernels per replay of each graph
  // projection output still to be added to the residual when enqueue_layers returns
  const float *pend_ptr = nullptr;
  int pend_count = 0;
  long long pend_str = 0;
  // per-kernel-class timing of an eager step (dis_engine_profile_decode)
  enum ProfClass { P_NORM = 0, P_QKV, P_ROPE, P_ATTN, P_OPROJ, P_GATEUP, P_DOWN, P_ALLREDUCE, P_HEAD_NORM, P_LM_HEAD, P_SAMPLE, P_COUNT };
  bool profiling = false;
  std::vector<cudaEvent_t> prof_events;  // pool
  std::vector<std::pair<int, int>> prof_marks;  // (class, index of the event recorded BEFORE that class)
  void prof_mark(int cls);
  cudaGraphExec_t last_graph = nullptr;
  int replay_last_step(int n, float *avg_ms);
  int profile_decode(int n, const int *slots, uint32_t *out_tokens, int *out_finish, float *class_ms, int *class_launches);
  // tensor parallel
  NcclApi *nccl = nullptr;
  void *comm = nullptr;
  // peer-memory exchange (k_tp.cu): one symmetric allocation per rank, mapped into every peer through CUDA IPC
  uint8_t *sym = nullptr;
  size_t sym_bytes = 0, sym_big_off = 0, sym_samp_off = 0;
  void *peer_base[kTpMaxRanks] = {};
  TpPeers peers{};
  TpSampPeers samp_peers{};  // vocab-parallel sampler's exchange area (k_tp_sample.cu)
  bool sampler_buckets = true;  // one GPU, sampled rows: bucket search (k_tp_sample.cu as a single rank); DIS_SAMPLER=bisect: sample_kernel
  bool tp_sample_fused = true;  // DIS_TP_SAMPLE=gather: all-gather the logits through NCCL and sample on every rank instead
  TpPeers peers_big{};      // second region of the symmetric buffer: fenced two-shot exchange for prefill-sized passes
  int ar_cap_big = 0;       // rows that region holds (0: absent, such passes go through NCCL)
  bool tp_custom = false;   // peers imported: all-reduce fused into the norm kernels
  int ar_cap_rows = 0;
  uint32_t ar_step = 0;
  bool pend_needs_ar = false;
  int ar_idx = 0;           // index of the next all-reduce inside the current forward pass
  int last_T = 0;           // tokens of the forward pass being enqueued
  bool rowmap_identity = true;  // decode: logits row i is token i
  int tp_export(uint8_t out[64]);
  int tp_import(const uint8_t *handles);
  cudaError_t consume_pending(const float *gain, const int *row_map, int rows, int T);

  ~Engine();
  int fail(int code, const std::string &msg);
  int init(const dis_engine_config *c);
  int load_tensor(int kind, int layer, const void *host, int dtype);
  int load_synthetic(uint64_t seed);
  int read_tensor(int kind, int layer, uint16_t *host, size_t *n);
  int finalize();
  int seq_add(const uint32_t *prompt, int n, const dis_sampling_params *p, int *slot);
  int seq_remove(int slot);
  int prefill(int n, const int *slots, uint32_t *out_tokens, int *out_finish);
  int decode(int n, const int *slots, uint32_t *out_tokens, int *out_finish);
  int step_mixed(int n, const int *slots, int max_prefill_tokens, uint32_t *out_tokens, int *out_finish, int *out_produced);
  int last_logits_host(float *host, int rows);
  int last_logprobs(int n, const int *slots, float *out);
  int status(dis_worker_status *out);
  int prefix_stats(dis_prefix_stats *out);

  // internals
  bool ensure_pages(Sequence &s, int slot, int n_tokens);
  void release_pages(Sequence &s);
  cudaError_t upload_step(int T, int R);
  cudaError_t enqueue_layers(int T, int n_splits);
  cudaError_t enqueue_head(int R, bool all_greedy);
  cudaError_t linear_partial(const TmaMap *xmaps, const uint16_t *xptr, const TmaMap &wmap, const uint16_t *wptr,
                             int w_row0, int T, int N, int K, int *split_out);
  cudaError_t allreduce_partials(int split, int T);
  cudaError_t gather_full_logits(int R);
  int pick_attn_splits(int T, int max_ctx_in_batch) const;
  void finish_check(Sequence &s, uint32_t tok);
  struct TensorSlot { void *ptr; size_t rows, cols, ld, row0, col0, full_rows, full_cols; bool f32; };
  bool tensor_slot(int kind, int layer, TensorSlot *out);
};

}  // namespace dis
