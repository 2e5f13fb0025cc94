/*
 * dis_engine.h -- C ABI of the B200-native decoder forward path for
 * its-me-ojas/distributed-inference-server.
 *
 * The reference has no FFI boundary today (no extern "C", no build.rs).  Its only seam
 * for this path is the spec'd `InferenceWorker` trait (.kiro/specs/distributed-inference-server/
 * design.md:334-342), whose implementing file is empty (crates/inference/src/worker.rs:1).  Every
 * entry point below is what a Rust `impl InferenceWorker` in crates/inference would bind through a
 * `-sys` shim; each cites the reference item it serves.  INTEGRATION.md shows that binding.
 *
 * Conventions
 *  - Plain C: pointers, sizes, PODs.  No torch / C++ types.  Handles are opaque.
 *  - The caller owns every host buffer passed in or out (Vec<u32> in, Vec<u32> out:
 *    design.md:244,355); the engine owns all device memory.
 *  - Every function returns a dis_status code.  The codes mirror `WorkerError`
 *    (crates/core/src/error.rs:115-128); the `String` payload of `InferenceFailed` is
 *    dis_engine_last_error().
 *  - The trait is `Send + Sync` with `&self` methods (design.md:335-338): calls may come from any
 *    host thread; the engine serialises them per handle.  Blocking; never calls back.
 *  - There is no CPU fallback: every compute entry point needs a CUDA device of compute
 *    capability 10.0 and fails with DIS_ERR_INFERENCE_FAILED otherwise.
 */
#ifndef DIS_ENGINE_H_
#define DIS_ENGINE_H_

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define DIS_API __attribute__((visibility("default")))
#define DIS_ABI_VERSION 2 /* 2: 16-byte request ids, logprob on token events / callback, dis_serving_cancel */

/* ---- status codes: WorkerError, crates/core/src/error.rs:115-128 ------------------------------ */
typedef enum dis_status {
  DIS_OK = 0,
  DIS_ERR_MODEL_NOT_LOADED = 1, /* WorkerError::ModelNotLoaded  error.rs:118 */
  DIS_ERR_INFERENCE_FAILED = 2, /* WorkerError::InferenceFailed error.rs:121 (message: last_error) */
  DIS_ERR_SHUTDOWN = 3,         /* WorkerError::Shutdown        error.rs:124 */
  DIS_ERR_OUT_OF_MEMORY = 4,    /* WorkerError::OutOfMemory     error.rs:127 */
  DIS_ERR_INVALID_ARGUMENT = 5, /* caller bug; a Rust shim maps it to InferenceFailed */
  DIS_ERR_QUEUE_FULL = 6        /* QueueError::Full (error.rs:79-89) -> ApiError::QueueFull, HTTP 503 (error.rs:42) */
} dis_status;

/* ---- FinishReason: crates/core/src/models.rs:26-32 (plus "still running") ---------------------- */
typedef enum dis_finish_reason {
  DIS_FINISH_NONE = 0,          /* not finished */
  DIS_FINISH_STOP = 1,          /* FinishReason::Stop         (EOS token) */
  DIS_FINISH_LENGTH = 2,        /* FinishReason::Length       (max_tokens or context limit) */
  DIS_FINISH_STOP_SEQUENCE = 3, /* FinishReason::StopSequence */
  DIS_FINISH_FAILED = 4         /* per-sequence failure (design.md:812-816, Property 22) */
} dis_finish_reason;

/* ---- tensors of the model (for weight loading) -------------------------------------------------- */
typedef enum dis_tensor_kind {
  DIS_T_EMBED = 0, DIS_T_ATTN_NORM = 1, DIS_T_WQ = 2, DIS_T_WK = 3, DIS_T_WV = 4, DIS_T_WO = 5,
  DIS_T_MLP_NORM = 6, DIS_T_WGATE = 7, DIS_T_WUP = 8, DIS_T_WDOWN = 9, DIS_T_FINAL_NORM = 10,
  DIS_T_LM_HEAD = 11
} dis_tensor_kind;

typedef enum dis_dtype { DIS_F32 = 0, DIS_BF16 = 1 } dis_dtype;

/* Llama decoder dimensions.  The reference names the model family only ("Llama-3 8B or compatible",
 * requirements.md:5); ModelInfo is named but never defined (design.md:341). */
typedef struct dis_model_config {
  int32_t vocab, hidden, layers, n_heads, n_kv_heads, head_dim, inter, max_ctx;
  float rope_theta, rms_eps;
} dis_model_config;

/* Engine configuration: WorkerConfig (design.md:316-322) plus what it lacks for a GPU path.
 * `Quantization` (design.md:326-332) has no bf16 variant; this engine is bf16 weights / fp32
 * accumulate only. */
#define DIS_FLAG_NO_GRAPH 1u   /* do not capture the decode step in a CUDA graph */
#define DIS_FLAG_FORCE_GEMV 2u /* debugging: run every projection through the GEMV kernel */
#define DIS_FLAG_NO_PDL 4u     /* launch the kernels of a step without programmatic dependent launch */
#define DIS_FLAG_PREFIX_CACHE 8u /* keep the full KV pages of finished prompts and map them into later requests that start
                                   with the same tokens (KVCacheManager::get_prefix, design.md:396; Property 9) */
typedef struct dis_engine_config {
  dis_model_config model;
  int32_t device;              /* CUDA ordinal (WorkerConfig.gpu_layers has no meaning here) */
  int32_t tp_rank, tp_size;    /* Megatron tensor parallel position; 0,1 = single GPU */
  int32_t max_batch;           /* concurrent sequences (requirements.md:46 default 32) */
  int32_t max_step_tokens;     /* tokens per forward pass (prefill chunk); 0 = default */
  int32_t page_size;           /* tokens per KV page; 0 = 64 */
  int64_t kv_pages;            /* KV pages to allocate; 0 = derive from kv_mem_fraction */
  float kv_mem_fraction;       /* of free device memory after weights; 0 = 0.85 */
  uint32_t flags;
  int32_t eos_token;           /* -1 = none */
  void *stream;                /* cudaStream_t to run on; NULL = engine-owned stream */
  uint8_t nccl_id[128];        /* ncclUniqueId from dis_tp_unique_id when tp_size > 1 */
} dis_engine_config;

/* InferenceParams: design.md:666-671, fed from GenerateRequest.{max_tokens,temperature,top_p,
 * stop_sequences} (models.rs:60-74).  temperature 0 = greedy (0.0 is valid: validator.rs:22).
 * There is no top_k in the reference's request model. */
typedef struct dis_sampling_params {
  int32_t max_tokens;
  float temperature;
  float top_p;
  uint64_t seed;
  const uint32_t *stop_tokens; /* concatenated stop sequences (Vec<Vec<u32>>) */
  const int32_t *stop_lens;    /* length of each */
  int32_t n_stop;
} dis_sampling_params;

/* WorkerStatus: design.md:290-297. */
typedef struct dis_worker_status {
  int32_t active_sequences;    /* WorkerStatus.active_batches, at sequence granularity */
  int32_t is_healthy;
  uint64_t memory_used;        /* bytes of device memory held by this engine */
  uint64_t memory_available;   /* bytes still free on the device */
  int64_t kv_pages_total, kv_pages_free;
  uint64_t tokens_generated;   /* BatchResult.tokens_generated, cumulative (design.md:349) */
  double last_step_ms;         /* device time of the last prefill/decode call (CUDA events) */
  uint64_t kernel_launches;    /* kernels launched by this engine so far (incl. graph replays) */
} dis_worker_status;

typedef struct dis_engine dis_engine;

/* ---- lifecycle: InferenceWorker::initialize / shutdown (design.md:337,339) --------------------- */
DIS_API int dis_abi_version(void);
/* sha256 prefix over the sources, headers and compiler flags this binary was built from (distributed_inference_server_b200/
 * build.py source_id()); lets a host or a test refuse a stale library.  No reference counterpart. */
DIS_API const char *dis_build_id(void);
DIS_API int dis_engine_create(const dis_engine_config *cfg, dis_engine **out);
DIS_API int dis_engine_destroy(dis_engine *e);
/* Global message for failures that have no handle yet (create). */
DIS_API const char *dis_last_error(void);
DIS_API const char *dis_engine_last_error(dis_engine *e);

/* Weights.  WorkerConfig.model_path (design.md:317) is resolved by the host: it hands every tensor
 * over as a full, unsharded [out,in] row-major matrix and the engine keeps its tensor-parallel
 * slice.  `layer` is ignored for EMBED / FINAL_NORM / LM_HEAD. */
DIS_API int dis_engine_load_tensor(dis_engine *e, int kind, int layer, const void *host, int dtype);
/* Random-init weights generated on the device from a counter-based hash (benchmarks, tests). */
DIS_API int dis_engine_load_synthetic(dis_engine *e, uint64_t seed);
/* Copy the engine's (sharded) copy of a tensor back as bf16 bit patterns; *n_elems in = capacity. */
DIS_API int dis_engine_read_tensor(dis_engine *e, int kind, int layer, uint16_t *host, size_t *n_elems);
/* After this the model counts as loaded; compute calls before it return MODEL_NOT_LOADED. */
DIS_API int dis_engine_finalize(dis_engine *e);

/* ---- InferenceWorker::infer, step-wise (design.md:338; InferenceBatch design.md:241-248) -------- *
 * Step-wise so that the host can emit TokenEvent::Token per token (models.rs:272-277, design.md:265),
 * drop a sequence when its client disconnects (requirements.md:85) and admit new sequences between
 * decode steps. */
/* Register one request: prompt tokens + params.  Reserves KV pages for the prompt. */
DIS_API int dis_engine_seq_add(dis_engine *e, const uint32_t *prompt, int32_t n_prompt,
                               const dis_sampling_params *params, int32_t *out_slot);
DIS_API int dis_engine_seq_remove(dis_engine *e, int32_t slot);
/* Run the prompts of `n` registered sequences and sample their first token.
 * out_tokens[n], out_finish[n] (dis_finish_reason). */
DIS_API int dis_engine_prefill(dis_engine *e, int32_t n, const int32_t *slots, uint32_t *out_tokens,
                               int32_t *out_finish);
/* One decode step: each listed sequence consumes its last sampled token and yields the next. */
DIS_API int dis_engine_decode(dis_engine *e, int32_t n, const int32_t *slots, uint32_t *out_tokens,
                              int32_t *out_finish);
/* One forward pass shared by running sequences and prompts that are still being read (chunked prefill; SURVEY.md §8
 * row f1; the reference's batcher is a 50 ms window batcher in its spec only, design.md:227-267).  A listed sequence
 * that has been prefilled decodes one token exactly as in dis_engine_decode; one that has not advances through its
 * prompt by a chunk -- all unfinished prompts together get at most max_prefill_tokens tokens (0 = the engine's
 * max_step_tokens), handed out in list order -- and samples its first token in the pass that reaches the prompt's
 * end.  out_produced[i] = 1 where out_tokens[i] / out_finish[i] carry a new token.  dis_engine_prefill and
 * dis_engine_decode remain the one-kind-per-call entry points; the three can be mixed on one engine. */
DIS_API int dis_engine_step(dis_engine *e, int32_t n, const int32_t *slots, int32_t max_prefill_tokens,
                            uint32_t *out_tokens, int32_t *out_finish, int32_t *out_produced);
/* dis_engine_decode run eagerly with a CUDA event before every kernel class (no reference counterpart;
 * it is how bench.py times the dominant kernel on the stream it runs on).  class_ms / class_launches
 * have DIS_PROF_COUNT entries: device milliseconds and launches per class for this one step. */
enum { DIS_PROF_NORM = 0, DIS_PROF_QKV, DIS_PROF_ROPE, DIS_PROF_ATTN, DIS_PROF_OPROJ, DIS_PROF_GATEUP, DIS_PROF_DOWN,
       DIS_PROF_ALLREDUCE, DIS_PROF_HEAD_NORM, DIS_PROF_LM_HEAD, DIS_PROF_SAMPLE, DIS_PROF_COUNT };
DIS_API int dis_engine_profile_decode(dis_engine *e, int32_t n, const int32_t *slots, uint32_t *out_tokens,
                                      int32_t *out_finish, float *class_ms, int32_t *class_launches);
/* Measurement only (no reference counterpart): replays the CUDA graph of the most recent dis_engine_decode call
 * `n` times back to back -- same inputs, same outputs, nothing advances -- and returns the average device time of
 * one replay.  The difference to the time of a real step is what the host adds between steps (metadata upload,
 * launch, token read-back, the caller's own work). */
DIS_API int dis_engine_replay_last_step(dis_engine *e, int32_t n, float *avg_ms);
/* Logits of the last prefill/decode call, row i = i-th listed sequence, [n][vocab] fp32 (gathered
 * across tensor-parallel ranks). For parity tests. */
DIS_API int dis_engine_last_logits(dis_engine *e, float *host, int32_t n_rows);
/* TokenEvent.logprob (crates/core/src/models.rs:272-277): the log-probability of the token each listed sequence
 * received from its most recent prefill/decode call, computed by the sampler kernel and copied back with the token
 * ids (no logits leave the device).  Greedy (temperature 0): log softmax(logits)[token]; sampled: the same under
 * softmax(logits / temperature), before the nucleus cut.  Under tensor parallelism the vocabulary slices' (max,
 * sum-exp) pairs are merged with the candidates. */
DIS_API int dis_engine_last_logprobs(dis_engine *e, int32_t n, const int32_t *slots, float *out_logprobs);

/* ---- InferenceWorker::status / model_info (design.md:340-341) ---------------------------------- */
DIS_API int dis_engine_status(dis_engine *e, dis_worker_status *out);
DIS_API int dis_engine_model_info(dis_engine *e, dis_model_config *out);

/* ---- tensor parallel bootstrap (no reference counterpart: the reference specifies replicas only,
 * requirements.md:100-110; north_star asks for Megatron TP) ------------------------------------- */
DIS_API int dis_tp_unique_id(uint8_t out[128]);
/* Peer-memory exchange: every rank exports a 64-byte CUDA IPC handle of its symmetric buffer, the host
 * all-gathers them (rank order) and hands all tp_size handles to every rank.  After the import the all-reduce
 * after the O and down projections is fused into the residual + RMSNorm kernel (P2P loads over NVLink, flag
 * signalling), the greedy candidates are merged the same way, and temperature / nucleus rows are sampled with the
 * logits left sharded (a few scalars per row exchanged over peer memory, csrc/k_tp_sample.cu); without it the engine
 * uses NCCL (all-reduce, and an all-gather of the logits for sampled rows). */
DIS_API int dis_engine_tp_export(dis_engine *e, uint8_t out[64]);
DIS_API int dis_engine_tp_import(dis_engine *e, const uint8_t *handles /* [tp_size][64] */);

/* ---- InferenceWorker::infer as one blocking call (design.md:338,345-360) ------------------------ *
 * The C++ InferenceWorker mirror (csrc/worker.hpp) exported for non-C++ hosts.  input_ids is the
 * padded [n][stride] batch of InferenceBatch (design.md:244), lens[b] the unpadded length
 * (BatchedRequest.original_length, design.md:261).  out_tokens is [n][max_new_cap].
 * on_token (optional) is called once per generated token, in step order, from the calling thread
 * (TokenEvent::Token{token, index, logprob}, models.rs:272-277); returning non-zero drops that sequence. */
typedef int (*dis_token_callback)(void *user, int32_t request_index, uint32_t token, int32_t index,
                                  int32_t finish_reason, float logprob);
typedef struct dis_batch_result {
  double inference_ms;          /* BatchResult.inference_time  design.md:348 */
  double ttft_ms;               /* time to first token of the batch (prefill) */
  uint64_t tokens_generated;    /* BatchResult.tokens_generated design.md:349 */
} dis_batch_result;
DIS_API int dis_worker_infer(dis_engine *e, int32_t n, const uint32_t *input_ids, int32_t stride,
                             const int32_t *lens, const dis_sampling_params *params /*[n]*/,
                             uint32_t *out_tokens, int32_t max_new_cap, int32_t *out_lens,
                             int32_t *out_finish, dis_batch_result *result, dis_token_callback on_token,
                             void *user);

/* ---- admission queue + continuous batching ------------------------------- *
 * Host-side callers of the step-wise entry points above.  The queue restates the behaviour of the reference's
 * PriorityQueueManager (crates/core/src/queue.rs:75-250): three FIFO lanes served High -> Normal -> Low
 * (Priority, types.rs:17-28), enqueue refused (DIS_ERR_QUEUE_FULL) at max_queue_size or while back-pressure
 * is on, back-pressure switched on above high_watermark and off below low_watermark, requests older than
 * request_timeout_s expired; defaults (cfg == NULL) are queue.rs:24-33.  The batcher replaces the spec's
 * 50 ms window batcher (design.md:227-267): every dis_serving_step admits queued requests into free engine
 * slots (KV pages permitting) and runs one forward pass in which the running sequences decode a token and the admitted
 * prompts advance by a bounded chunk (dis_engine_step), and
 * reports one event per generated token (TokenEvent, models.rs:268-288).  engine == NULL gives a queue-only
 * object (admission logic without a GPU).  The *_at variants take the clock as an argument. */
typedef enum dis_priority { DIS_PRIORITY_LOW = 0, DIS_PRIORITY_NORMAL = 1, DIS_PRIORITY_HIGH = 2 } dis_priority;
typedef struct dis_queue_config {
  int32_t high_watermark, low_watermark, max_queue_size;
  double request_timeout_s;
} dis_queue_config;
typedef struct dis_queue_depth { int32_t high, normal, low, total, accepting, active; } dis_queue_depth;
typedef enum dis_event_kind { DIS_EVENT_TOKEN = 0, DIS_EVENT_DONE = 1, DIS_EVENT_EXPIRED = 2 } dis_event_kind;
typedef struct dis_serving_event {
  uint8_t request_id[16]; /* RequestID = Uuid (crates/core/src/types.rs:7): 16 opaque bytes, as submitted */
  int32_t kind;           /* dis_event_kind */
  uint32_t token;         /* TOKEN */
  int32_t index;          /* TOKEN: position in the completion; DONE: completion_tokens (Usage, models.rs:9-23) */
  int32_t finish_reason;  /* dis_finish_reason; DIS_FINISH_NONE on a token that does not end the request */
  int32_t prompt_tokens;
  double waited_ms;       /* first TOKEN: time from submit to first token (TTFT); EXPIRED: time spent queued */
  float logprob;          /* TOKEN: TokenEvent.logprob (models.rs:276), see dis_engine_last_logprobs */
} dis_serving_event;
typedef struct dis_serving dis_serving;
DIS_API int dis_serving_create(dis_engine *engine, const dis_queue_config *cfg, int32_t max_active, dis_serving **out);
DIS_API int dis_serving_destroy(dis_serving *s);
DIS_API const char *dis_serving_last_error(dis_serving *s);
DIS_API int dis_serving_submit(dis_serving *s, const uint8_t request_id[16], const uint32_t *prompt, int32_t n_prompt,
                               const dis_sampling_params *params, int32_t priority);
DIS_API int dis_serving_submit_at(dis_serving *s, const uint8_t request_id[16], const uint32_t *prompt, int32_t n_prompt,
                                  const dis_sampling_params *params, int32_t priority, double now_s);
/* Client disconnect (requirements.md:85): a queued request leaves the queue, a running one frees its slot and KV
 * pages; one DONE event with DIS_FINISH_STOP follows.  DIS_ERR_INVALID_ARGUMENT if the id is neither queued nor running. */
DIS_API int dis_serving_cancel(dis_serving *s, const uint8_t request_id[16]);
DIS_API int dis_serving_step(dis_serving *s, dis_serving_event *events, int32_t cap, int32_t *n_events);
DIS_API int dis_serving_step_at(dis_serving *s, double now_s, dis_serving_event *events, int32_t cap, int32_t *n_events);
DIS_API int dis_serving_depth(dis_serving *s, dis_queue_depth *out);
/* Prompt tokens read per scheduler step beside the running sequences' decode rows (default 4096): the knob between
 * time-to-first-token of waiting requests and inter-token latency of running ones. */
DIS_API int dis_serving_set_prefill_budget(dis_serving *s, int32_t tokens);
/* Which admitted prompt is read first inside a pass.  The admission queue keeps the reference's order (priority lane,
 * then arrival: crates/core/src/queue.rs:130-158); among requests that already hold a slot -- an order the reference
 * does not define -- mode 1 (default) reads the shortest prompt first inside a priority class, a waiting request
 * gaining aging_tokens_per_s (default 4096) tokens of precedence per second, ties in arrival order; mode 0 keeps
 * admission order.  dis_serving_order_prompts is that rule as a pure function (no engine, no GPU): order[] receives the
 * indices 0..n-1 in reading order. */
DIS_API int dis_serving_set_prefill_order(dis_serving *s, int32_t mode, double aging_tokens_per_s);
DIS_API int dis_serving_order_prompts(int32_t n, const int32_t *priority, const int32_t *prompt_tokens,
                                      const double *waited_s, double aging_tokens_per_s, int32_t *order);
/* The rule dis_engine_step uses to share a pass, as a pure function (no engine, no GPU) so that it can be property-
 * tested: remaining_prompt[i] <= 0 marks a decode row (take[i] = 1); the other entries receive, in list order, as
 * much of their remaining prompt as still fits into min(budget, cap - decode rows) tokens. */
DIS_API int dis_serving_plan_step(int32_t n, const int32_t *remaining_prompt, int32_t budget, int32_t cap, int32_t *take);
DIS_API int dis_serving_debug_pop(dis_serving *s, int32_t max, uint8_t *ids /* [max][16] */, int32_t *n);

/* ---- prefix KV reuse -------------------------------------------------------- *
 * The spec's KVCacheManager (design.md:363-413) caches K/V per token sequence as host Vec<f32> and hands the
 * longest cached prefix back through get_prefix (Property 9, design.md:734-738), evicting least-recently-used
 * entries under memory pressure (Property 10).  Here the cache is the device page pool itself: with
 * DIS_FLAG_PREFIX_CACHE the full pages of a prefilled prompt (and, when the sequence is removed, of what it
 * generated) stay registered under a hash of the tokens up to and including the page; dis_engine_seq_add maps
 * the longest registered prefix into the new sequence and dis_engine_prefill computes only the remaining
 * tokens.  Pages nobody maps are handed back to the allocator in LRU order when it runs dry.
 * dis_prefix_stats mirrors CacheStats (design.md:404-411). */
typedef struct dis_prefix_stats {
  int64_t cached_pages;    /* CacheStats.entries: registered pages (mapped or not) */
  int64_t evictable_pages; /* registered, mapped by no live sequence */
  int64_t free_pages;      /* never registered or already evicted */
  uint64_t lookups;        /* requests looked up */
  uint64_t hits;           /* CacheStats.hit_count: requests that reused at least one page (miss_count = lookups - hits) */
  uint64_t lookup_tokens;  /* prompt tokens looked up */
  uint64_t hit_tokens;     /* prompt tokens whose K/V were reused instead of computed */
  uint64_t evictions;      /* CacheStats.eviction_count */
} dis_prefix_stats;
DIS_API int dis_engine_prefix_stats(dis_engine *e, dis_prefix_stats *out);

/* The same page index without an engine, so that the admission / reuse / eviction rules can be tested
 * without a GPU: admit = longest-prefix match + allocation of the remaining pages for n_tokens + 1 positions
 * (all or nothing), commit = register the full pages, release = the sequence is gone. */
typedef struct dis_prefix_index dis_prefix_index;
DIS_API int dis_prefix_index_create(int32_t page_size, int64_t n_pages, dis_prefix_index **out);
DIS_API int dis_prefix_index_destroy(dis_prefix_index *ix);
DIS_API int dis_prefix_index_admit(dis_prefix_index *ix, const uint32_t *tokens, int32_t n_tokens, int32_t *pages,
                                   int32_t pages_cap, int32_t *n_pages, int32_t *matched_tokens);
DIS_API int dis_prefix_index_commit(dis_prefix_index *ix, const uint32_t *tokens, int32_t n_tokens,
                                    const int32_t *pages, int32_t n_pages);
DIS_API int dis_prefix_index_release(dis_prefix_index *ix, const int32_t *pages, int32_t n_pages);
DIS_API int dis_prefix_index_stats(dis_prefix_index *ix, dis_prefix_stats *out);

/* ---- kernel-level entry points ------------------------------------------------------------------ *
 * One per hand-written kernel, taking DEVICE pointers and a cudaStream_t (NULL = default stream).
 * They exist so that parity tests can drive each kernel through the C ABI; the reference has no
 * counterpart for any of them ("None").  bf16 buffers are uint16_t bit patterns. */

/* x[T][H] fp32 residual.  If tokens != NULL: v = embed[tokens[r]] (bf16 table); else
 * v = x[r] + sum_s partial[s*partial_stride + r*H ...] (n_partial may be 0), with r = row_map ? row_map[t] : t.
 * Then xn[t] = bf16(v * rsqrt(mean v^2 + eps) * gain).  Without row_map, x[t] = v is written back;
 * with row_map (the rows that need logits) x is left untouched and xn rows are compact. */
DIS_API int dis_k_add_rmsnorm(float *x, const uint32_t *tokens, const uint16_t *embed, const float *partial,
                              int32_t n_partial, int64_t partial_stride, const float *gain, uint16_t *xn,
                              const int32_t *row_map, int32_t T, int32_t H, float eps, void *stream);
/* Y = X[T][K] * W[N][K]^T, bf16 in, fp32 out as split-K partials out[S][T][N] (S chosen by the
 * caller, 1 = plain result).  tcgen05/TMEM/TMA kernel. */
DIS_API int dis_k_gemm(const uint16_t *X, const uint16_t *W, float *out, int32_t T, int32_t N, int32_t K,
                       int32_t split_k, void *stream);
/* The split-K factor the engine uses for a [T][K] x [N][K]^T projection (never leaves a slice empty). */
DIS_API int dis_k_gemm_pick_split(int32_t T, int32_t N, int32_t K);
/* act[T][I] = bf16(silu(X*Wg^T) * (X*Wu^T)); Wgu is [2I][K]: gate rows then up rows. */
DIS_API int dis_k_gemm_silu(const uint16_t *X, const uint16_t *Wgu, uint16_t *act, int32_t T, int32_t I,
                            int32_t K, void *stream);
/* Same contracts on the bandwidth-bound CUDA-core GEMV path (batch-1 decode). */
DIS_API int dis_k_gemv(const uint16_t *X, const uint16_t *W, float *out, int32_t T, int32_t N, int32_t K,
                       void *stream);
DIS_API int dis_k_gemv_silu(const uint16_t *X, const uint16_t *Wgu, uint16_t *act, int32_t T, int32_t I,
                            int32_t K, void *stream);
/* qkv partials [S][T][(nh+2nkv)*d] fp32 -> RoPE(q) bf16 [T][nh*d]; RoPE(k), v -> KV pages.
 * Page pool layout (per layer): [page][2 (K,V)][nkv][page_size][d] bf16.  page_table is
 * [rows][max_pages]; token t uses row pt_row[t] and position pos[t]. */
DIS_API int dis_k_rope_kv_write(const float *qkv, int32_t n_partial, int64_t partial_stride, uint16_t *q,
                                uint16_t *kv_pool, const int32_t *page_table, int32_t max_pages,
                                const int32_t *pt_row, const int32_t *pos, const float *rope_cos,
                                const float *rope_sin, int32_t T, int32_t n_heads, int32_t n_kv,
                                int32_t d, int32_t page_size, void *stream);
/* Attention of each query token over its paged history (ctx[t] tokens, causal by construction).
 * out bf16 [T][nh*d].  n_splits > 1 uses `scratch` (fp32, dis_k_attn_scratch_elems()) for the
 * split-KV partials and a combine pass.  kv_pool holds n_pages pages (the TMA map is bounded by it);
 * rows of a page that lie past ctx[t] must hold finite values (zero-fill the pool once). */
DIS_API size_t dis_k_attn_scratch_elems(int32_t T, int32_t n_heads, int32_t d, int32_t n_splits);
DIS_API int dis_k_attn_paged(const uint16_t *q, const uint16_t *kv_pool, int64_t n_pages, const int32_t *page_table,
                             int32_t max_pages, const int32_t *pt_row, const int32_t *ctx, uint16_t *out,
                             float *scratch, int32_t n_splits, int32_t T, int32_t n_heads, int32_t n_kv,
                             int32_t d, int32_t page_size, void *stream);
/* dis_k_attn_paged without a scratch argument, for callers that cannot allocate device memory: a server
 * that has already given all free HBM to its KV pages, or a request loop that must never run into the
 * device-wide synchronisation of cudaMalloc / cudaFree.  It has no counterpart in the reference.
 * The split-KV partials go to an array built into the library (DIS_ATTN_STATIC_SCRATCH_FLOATS floats =
 * 8 MiB, present on every GPU that loads the library until the process exits).  A call needing more than
 * that -- dis_k_attn_scratch_elems(T, n_heads, d, n_splits) -- is refused with DIS_ERR_INVALID_ARGUMENT and a
 * message naming both numbers.  The entry point itself performs no cudaMalloc, cudaFree or other allocation.
 * The first call on each device looks up that device's copy of the array (a one-time runtime query made
 * with the caller's current device); make that first call before any stream capture begins.
 * All calls share the one array: two calls whose kernels can overlap on the GPU (different streams, or
 * different host threads without ordering) race on it and corrupt each other's result -- order them.
 * Nothing in it needs clearing between calls: each call overwrites every slot it later reads. */
#define DIS_ATTN_STATIC_SCRATCH_FLOATS (2u << 20)
DIS_API int dis_k_attn_paged_static(const uint16_t *q, const uint16_t *kv_pool, int64_t n_pages,
                                    const int32_t *page_table, int32_t max_pages, const int32_t *pt_row,
                                    const int32_t *ctx, uint16_t *out, int32_t n_splits, int32_t T, int32_t n_heads,
                                    int32_t n_kv, int32_t d, int32_t page_size, void *stream);
/* Prefill attention: query block b is the qb_len[b] consecutive tokens [qb_tok0[b], ..) of the flattened
 * batch, at positions qb_pos0[b].. of the sequence with page-table row qb_ptrow[b]; causal over keys [0, pos].
 * The block's own K/V must already be in the pages.  q / out are [tokens][nh*d] bf16.
 * qb_len[b] <= dis_k_attn_prefill_block_tokens(n_heads, n_kv, d): 128 / (n_heads / n_kv) where the tcgen05 / TMEM
 * kernel applies (head_dim 128, 4 or 8 query heads per kv head: S = Q K^T and P V as tcgen05.mma over 128 query rows,
 * k_attn_prefill_tc.cu), 16 on the mma.sync kernel that serves every other geometry. */
DIS_API int dis_k_attn_prefill_block_tokens(int32_t n_heads, int32_t n_kv, int32_t d);
DIS_API int dis_k_attn_prefill(const uint16_t *q, const uint16_t *kv_pool, int64_t n_pages, const int32_t *page_table,
                               int32_t max_pages, const int32_t *qb_tok0, const int32_t *qb_len,
                               const int32_t *qb_pos0, const int32_t *qb_ptrow, uint16_t *out, int32_t n_blocks,
                               int32_t n_heads, int32_t n_kv, int32_t d, int32_t page_size, void *stream);
/* Sampler over fp32 logits [T][V] (row stride ld): temperature 0 -> argmax, lowest index wins ties;
 * else nucleus sampling as oracle/llama_oracle.c states it.  out_tokens[T]; out_logprob[T] (may be NULL) receives
 * the chosen token's log-probability (see dis_engine_last_logprobs). */
DIS_API int dis_k_sample(const float *logits, int64_t ld, int32_t T, int32_t V, const float *temperature,
                         const float *top_p, const uint64_t *seeds, const uint64_t *steps,
                         uint32_t *out_tokens, float *out_logprob, void *stream);
/* The vocab-parallel sampler of tensor-parallel engines (csrc/k_tp_sample.cu: each rank keeps its V/tp logit columns
 * and exchanges a few scalars per row over peer memory -- candidates for greedy rows, 16 partial masses per round of
 * a 16-ary threshold search for nucleus rows, then the drawn token) with its tp ranks emulated on ONE GPU, so that it
 * can be checked against the oracle and against dis_k_sample without a second GPU.  logits [T][V] fp32, V a multiple
 * of tp, T * tp <= 148 (the emulated ranks wait for each other inside one grid).  area: tp *
 * dis_k_sample_vocab_parallel_area_bytes(tp, area_rows) bytes of device memory, zero-filled once by the caller;
 * call_no must differ between consecutive calls on the same area.  out_tokens / out_logprob are [tp][T]: what each
 * emulated rank wrote -- all slices must be equal. */
DIS_API size_t dis_k_sample_vocab_parallel_area_bytes(int32_t tp, int32_t rows);
DIS_API int dis_k_sample_vocab_parallel(const float *logits, int32_t T, int32_t V, int32_t tp, const float *temperature,
                                        const float *top_p, const uint64_t *seeds, const uint64_t *steps,
                                        uint32_t call_no, void *area, int32_t area_rows, uint32_t *out_tokens,
                                        float *out_logprob, void *stream);
/* Synthetic weights: dst[i] = bf16(offset + (2u-1)*amp), u from the counter hash (seed, kind, layer, i). */
DIS_API int dis_k_synth_bf16(uint16_t *dst, size_t n, uint64_t seed, int32_t kind, int32_t layer, float amp,
                             float offset, void *stream);

#ifdef __cplusplus
}
#endif
#endif /* DIS_ENGINE_H_ */
