// engine.cu -- decoder engine: weights, paged KV cache, prefill / decode steps, CUDA-graph replay of
// the decode step, Megatron tensor parallelism.  This is the body of the spec'd InferenceWorker
// (.kiro/specs/distributed-inference-server/design.md:334-342); the reference's own worker file is
// empty (crates/inference/src/worker.rs:1), so the step structure follows SURVEY.md §3.3.
#include "engine.h"

#include <algorithm>
#include <cmath>
#include <cstdlib>
#include <cstring>

namespace dis {

// ------------------------------------------------------------------------------------ error state
static thread_local std::string g_last_error;
void set_last_error(const std::string &msg) { g_last_error = msg; }
const char *get_last_error() { return g_last_error.c_str(); }

int Engine::fail(int code, const std::string &msg) {
  last_error = msg;
  g_last_error = msg;
  return code;
}

#define ENG_CUDA(expr)                                                                                   \
  do {                                                                                                   \
    cudaError_t _e = (expr);                                                                             \
    if (_e != cudaSuccess) {                                                                             \
      healthy = false;                                                                                   \
      std::string _m = g_last_error.empty() ? std::string(cudaGetErrorString(_e)) : g_last_error;        \
      return fail(_e == cudaErrorMemoryAllocation ? DIS_ERR_OUT_OF_MEMORY : DIS_ERR_INFERENCE_FAILED,    \
                  std::string(#expr) + ": " + _m);                                                       \
    }                                                                                                    \
  } while (0)

#define RET_CUDA(expr)                      \
  do {                                      \
    cudaError_t _e = (expr);                \
    if (_e != cudaSuccess) return _e;       \
  } while (0)

template <class T_>
static cudaError_t dev_alloc(T_ **p, size_t n, uint64_t *acct) {
  *p = nullptr;
  if (n == 0) return cudaSuccess;
  cudaError_t e = cudaMalloc(reinterpret_cast<void **>(p), n * sizeof(T_));
  if (e == cudaSuccess) *acct += n * sizeof(T_);
  else set_last_error("cudaMalloc of " + std::to_string(n * sizeof(T_)) + " bytes failed");
  return e;
}

static size_t align_up(size_t v, size_t a) { return (v + a - 1) / a * a; }

// ------------------------------------------------------------------------------------ init
int Engine::init(const dis_engine_config *c) {
  cfg = *c;
  m = c->model;
  tp_rank = c->tp_rank;
  tp_size = c->tp_size <= 0 ? 1 : c->tp_size;
  if (m.vocab <= 0 || m.hidden <= 0 || m.layers <= 0 || m.n_heads <= 0 || m.n_kv_heads <= 0 || m.head_dim <= 0 ||
      m.inter <= 0 || m.max_ctx <= 0)
    return fail(DIS_ERR_INVALID_ARGUMENT, "model config: every dimension must be positive");
  if (m.head_dim != 64 && m.head_dim != 128) return fail(DIS_ERR_INVALID_ARGUMENT, "head_dim must be 64 or 128");
  if (m.n_heads % m.n_kv_heads != 0) return fail(DIS_ERR_INVALID_ARGUMENT, "n_heads must be a multiple of n_kv_heads");
  if (tp_rank < 0 || tp_rank >= tp_size) return fail(DIS_ERR_INVALID_ARGUMENT, "tp_rank out of range");
  if (m.n_kv_heads % tp_size || m.inter % tp_size || m.vocab % tp_size)
    return fail(DIS_ERR_INVALID_ARGUMENT, "n_kv_heads, inter and vocab must be divisible by tp_size");
  nh_l = m.n_heads / tp_size;
  nkv_l = m.n_kv_heads / tp_size;
  qd_l = nh_l * m.head_dim;
  kvd_l = nkv_l * m.head_dim;
  nqkv_l = qd_l + 2 * kvd_l;
  inter_l = m.inter / tp_size;
  vocab_l = m.vocab / tp_size;
  vocab_off = tp_rank * vocab_l;
  if (m.hidden % 8 || inter_l % 8 || qd_l % 8) return fail(DIS_ERR_INVALID_ARGUMENT, "hidden, inter/tp and n_heads*head_dim/tp must be multiples of 8");
  max_batch = c->max_batch > 0 ? c->max_batch : 32;
  cap_tokens = c->max_step_tokens > 0 ? c->max_step_tokens : 8192;
  cap_tokens = std::max(cap_tokens, max_batch);
  page_size = c->page_size > 0 ? c->page_size : 64;
  if ((page_size & (page_size - 1)) || page_size < 8) return fail(DIS_ERR_INVALID_ARGUMENT, "page_size must be a power of two >= 8");
  max_pages = (m.max_ctx + page_size - 1) / page_size;

  int ndev = 0;
  if (cudaGetDeviceCount(&ndev) != cudaSuccess || ndev == 0)
    return fail(DIS_ERR_INFERENCE_FAILED, "no CUDA device: this engine has no CPU fallback");
  if (c->device < 0 || c->device >= ndev) return fail(DIS_ERR_INVALID_ARGUMENT, "device ordinal out of range");
  ENG_CUDA(cudaSetDevice(c->device));
  cudaDeviceProp prop;
  ENG_CUDA(cudaGetDeviceProperties(&prop, c->device));
  if (prop.major != 10)
    return fail(DIS_ERR_INFERENCE_FAILED, std::string("device is sm_") + std::to_string(prop.major * 10 + prop.minor) +
                                              "; kernels are built for sm_100a only");
  if (c->stream) stream = (cudaStream_t)c->stream;
  else { ENG_CUDA(cudaStreamCreateWithFlags(&stream, cudaStreamNonBlocking)); own_stream = true; }
  ENG_CUDA(cudaEventCreate(&ev0));
  ENG_CUDA(cudaEventCreate(&ev1));

  const size_t L = m.layers, H = m.hidden;
  // ---- weights
  ENG_CUDA(dev_alloc(&embed, (size_t)m.vocab * H, &bytes_allocated));
  ENG_CUDA(dev_alloc(&wqkv, L * nqkv_l * H, &bytes_allocated));
  ENG_CUDA(dev_alloc(&wo, L * H * qd_l, &bytes_allocated));
  ENG_CUDA(dev_alloc(&wgu, L * 2 * inter_l * H, &bytes_allocated));
  ENG_CUDA(dev_alloc(&wdown, L * H * inter_l, &bytes_allocated));
  ENG_CUDA(dev_alloc(&lm_head, (size_t)vocab_l * H, &bytes_allocated));
  ENG_CUDA(dev_alloc(&attn_norm, L * H, &bytes_allocated));
  ENG_CUDA(dev_alloc(&mlp_norm, L * H, &bytes_allocated));
  ENG_CUDA(dev_alloc(&final_norm, H, &bytes_allocated));
  ENG_CUDA(make_tma_map_2d(&wqkv_map, wqkv, L * nqkv_l, H, 128));
  ENG_CUDA(make_tma_map_2d(&wo_map, wo, L * H, qd_l, 128));
  ENG_CUDA(make_tma_map_2d(&wgu_map, wgu, L * 2 * inter_l, H, (uint32_t)gemm_pick_tile_rows(inter_l)));
  ENG_CUDA(make_tma_map_2d(&wgu_map128, wgu, L * 2 * inter_l, H, 128));
  ENG_CUDA(make_tma_map_2d(&wgu_map64, wgu, L * 2 * inter_l, H, 64));
  ENG_CUDA(make_tma_map_2d(&wdown_map, wdown, L * H, inter_l, 128));
  ENG_CUDA(make_tma_map_2d(&lm_map, lm_head, vocab_l, H, 128));
  // ---- activations
  const size_t cap = cap_tokens;
  const size_t widest = std::max<size_t>(std::max<size_t>(nqkv_l, H), inter_l <= 4096 ? (size_t)2 * inter_l : 0);
  partial_elems = std::max(cap * widest, (size_t)16 * 128 * widest);
  ENG_CUDA(dev_alloc(&x, cap * H, &bytes_allocated));
  ENG_CUDA(dev_alloc(&xn, cap * H, &bytes_allocated));
  ENG_CUDA(dev_alloc(&q, cap * qd_l, &bytes_allocated));
  ENG_CUDA(dev_alloc(&attn, cap * qd_l, &bytes_allocated));
  ENG_CUDA(dev_alloc(&act, cap * inter_l, &bytes_allocated));
  ENG_CUDA(dev_alloc(&partial, partial_elems, &bytes_allocated));
  ENG_CUDA(dev_alloc(&logits, (size_t)max_batch * vocab_l, &bytes_allocated));
  attn_scratch_elems = (size_t)(max_batch + 2 * kNumSMs + 64) * nh_l * (2 + m.head_dim);
  ENG_CUDA(dev_alloc(&attn_scratch, attn_scratch_elems, &bytes_allocated));
  // sampler output: [max_batch] token ids, then [max_batch] log-probabilities (fp32) -- one D2H copy per step
  ENG_CUDA(dev_alloc(&d_out_tokens, (size_t)2 * max_batch, &bytes_allocated));
  ENG_CUDA(cudaMallocHost(reinterpret_cast<void **>(&h_out_tokens), (size_t)2 * max_batch * sizeof(uint32_t)));
  if (tp_size > 1) {
    ENG_CUDA(dev_alloc(&ar_buf, cap * H, &bytes_allocated));
    ENG_CUDA(dev_alloc(&full_logits, (size_t)max_batch * m.vocab, &bytes_allocated));
    ENG_CUDA(dev_alloc(&gather_buf, (size_t)max_batch * m.vocab, &bytes_allocated));
    ENG_CUDA(dev_alloc(&cand_val, (size_t)max_batch, &bytes_allocated));
    ENG_CUDA(dev_alloc(&cand_tok, (size_t)max_batch, &bytes_allocated));
    ENG_CUDA(dev_alloc(&cand_val_all, (size_t)max_batch * tp_size, &bytes_allocated));
    ENG_CUDA(dev_alloc(&cand_tok_all, (size_t)max_batch * tp_size, &bytes_allocated));
    ENG_CUDA(dev_alloc(&cand_sum, (size_t)max_batch, &bytes_allocated));
    ENG_CUDA(dev_alloc(&cand_sum_all, (size_t)max_batch * tp_size, &bytes_allocated));
  }
  const uint32_t boxes[4] = {32, 64, 128, 256};
  for (int i = 0; i < 4; i++) {
    ENG_CUDA(make_tma_map_2d(&xn_map[i], xn, cap, H, boxes[i]));
    ENG_CUDA(make_tma_map_2d(&attn_map[i], attn, cap, qd_l, boxes[i]));
    ENG_CUDA(make_tma_map_2d(&act_map[i], act, cap, inter_l, boxes[i]));
  }
  // ---- RoPE tables (computed in double on the host, like the oracle; values only depend on the math)
  {
    const int half = m.head_dim / 2;
    std::vector<float> cs((size_t)m.max_ctx * half), sn((size_t)m.max_ctx * half);
    for (int p = 0; p < m.max_ctx; p++)
      for (int i = 0; i < half; i++) {
        const double inv_freq = std::pow((double)m.rope_theta, -2.0 * (double)i / (double)m.head_dim);
        const double a = (double)p * inv_freq;
        cs[(size_t)p * half + i] = (float)std::cos(a);
        sn[(size_t)p * half + i] = (float)std::sin(a);
      }
    ENG_CUDA(dev_alloc(&rope_cos, cs.size(), &bytes_allocated));
    ENG_CUDA(dev_alloc(&rope_sin, sn.size(), &bytes_allocated));
    ENG_CUDA(cudaMemcpy(rope_cos, cs.data(), cs.size() * 4, cudaMemcpyHostToDevice));
    ENG_CUDA(cudaMemcpy(rope_sin, sn.data(), sn.size() * 4, cudaMemcpyHostToDevice));
  }
  // ---- step metadata: a small block for decode steps (one H2D copy per step) and a large one for prefill chunks
  auto make_meta = [&](StepMeta &s, int cap_t) -> cudaError_t {
    s.cap_tokens = cap_t;
    s.cap_rows = max_batch;
    size_t o = 0;
    auto sec = [&](size_t bytes) { size_t at = o; o = align_up(o + bytes, 16); return at; };
    const size_t ct = (size_t)cap_t;
    s.o_tokens = sec(ct * 4); s.o_pos = sec(ct * 4); s.o_ptrow = sec(ct * 4); s.o_ctx = sec(ct * 4);
    s.o_rowmap = sec((size_t)max_batch * 4); s.o_temp = sec((size_t)max_batch * 4); s.o_topp = sec((size_t)max_batch * 4);
    s.o_seed = sec((size_t)max_batch * 8); s.o_step = sec((size_t)max_batch * 8);
    s.o_qb_tok0 = sec(ct * 4); s.o_qb_len = sec(ct * 4); s.o_qb_pos0 = sec(ct * 4); s.o_qb_ptrow = sec(ct * 4);
    s.o_arstep = sec(16);
    s.bytes = o;
    cudaError_t e1 = cudaMallocHost(reinterpret_cast<void **>(&s.host), s.bytes);
    if (e1 != cudaSuccess) return e1;
    memset(s.host, 0, s.bytes);
    return dev_alloc(&s.dev, s.bytes, &bytes_allocated);
  };
  ENG_CUDA(make_meta(meta_dec, max_batch));
  ENG_CUDA(make_meta(meta_pre, cap_tokens));
  cur_meta = &meta_dec;
  // ---- KV pages: [layer][page][2][nkv_l][page_size][d]
  {
    const size_t page_elems = (size_t)2 * nkv_l * page_size * m.head_dim;
    const size_t page_bytes_all_layers = page_elems * 2 * L;
    int64_t want = c->kv_pages;
    const int64_t useful = (int64_t)max_batch * max_pages;
    if (want <= 0) {
      size_t free_b = 0, total_b = 0;
      ENG_CUDA(cudaMemGetInfo(&free_b, &total_b));
      const float frac = c->kv_mem_fraction > 0.f ? c->kv_mem_fraction : 0.85f;
      want = (int64_t)((double)free_b * frac / (double)page_bytes_all_layers);
      want = std::min(want, useful);
    }
    if (want < 1) return fail(DIS_ERR_OUT_OF_MEMORY, "no device memory left for KV pages");
    n_pages = want;
    kv_layer_stride = (size_t)n_pages * page_elems;
    ENG_CUDA(dev_alloc(&kv_pool, kv_layer_stride * L, &bytes_allocated));
    // stale rows inside a TMA box get probability 0 in attention, so they must never be NaN
    ENG_CUDA(cudaMemset(kv_pool, 0, kv_layer_stride * L * sizeof(uint16_t)));
    kv_rows_per_layer = (long long)n_pages * 2 * nkv_l * page_size;
    ENG_CUDA(make_kv_map(&kv_map, kv_pool, kv_rows_per_layer * (long long)L, m.head_dim, page_size));
    pages_ix.init(page_size, n_pages, (c->flags & DIS_FLAG_PREFIX_CACHE) != 0);
    ENG_CUDA(dev_alloc(&d_page_table, (size_t)max_batch * max_pages, &bytes_allocated));
    ENG_CUDA(cudaMallocHost(reinterpret_cast<void **>(&h_page_table), (size_t)max_batch * max_pages * sizeof(int)));
    memset(h_page_table, 0, (size_t)max_batch * max_pages * sizeof(int));
  }
  seqs.resize((size_t)max_batch);
  slot_stamp.assign((size_t)max_batch, 0);
  use_pdl = !(c->flags & DIS_FLAG_NO_PDL);
  pf_block = attn_prefill_block_tokens(nh_l, nkv_l, m.head_dim);
  pf_tc = attn_prefill_uses_tc(nh_l, nkv_l, m.head_dim);
  if (const char *g = getenv("DIS_GEMM_PERSIST")) use_persist = atoi(g) != 0;
  if (const char *g = getenv("DIS_GEMV_MAX_T")) gemv_max_t = std::max(0, std::min(kGemvMaxT, atoi(g)));
  if (const char *g = getenv("DIS_ATTN_MAX_SPLITS")) max_attn_splits = atoi(g);
  if (const char *g = getenv("DIS_SAMPLER")) sampler_buckets = strcmp(g, "bisect") != 0;
  // Measurement aid, never set in production: DIS_DEBUG_SKIP=norm,attn,... leaves those kernel classes out of the
  // decode step (its results are then garbage) so that step time with and without a class gives the class's marginal
  // cost INSIDE the graph, where programmatic dependent launch overlaps neighbours and ncu's isolated times mislead.
  if (const char *g = getenv("DIS_DEBUG_SKIP")) {
    const char *names[] = {"norm", "qkv", "rope", "attn", "oproj", "gateup", "down"};
    for (int i = 0; i < 7; i++) if (strstr(g, names[i])) skip_mask |= 1u << i;
  }
  // ---- tensor parallel communicator
  if (tp_size > 1) {
    std::string err;
    nccl = nccl_api(&err);
    if (!nccl) return fail(DIS_ERR_INFERENCE_FAILED, "tensor parallel needs NCCL: " + err);
    NcclId id;
    memcpy(id.bytes, c->nccl_id, 128);
    const int r = nccl->CommInitRank(&comm, tp_size, id, tp_rank);
    if (r != 0) return fail(DIS_ERR_INFERENCE_FAILED, std::string("ncclCommInitRank: ") + nccl->GetErrorString(r));
  }
  if (tp_size > 1) {
    if (tp_size > kTpMaxRanks) return fail(DIS_ERR_INVALID_ARGUMENT, "tp_size > 8");
    ar_cap_rows = std::min(std::max(max_batch, 128), 256);
    const size_t data_b = (size_t)std::max(2 * tp_size, 8) * ar_cap_rows * m.hidden * sizeof(float);  // 8: flag-in-data two-shot areas
    const size_t flags_b = align_up((size_t)2 * 2 * tp_size * ar_cap_rows * 4, 256);  // two hops
    const size_t cand_b = align_up((size_t)2 * tp_size * ar_cap_rows * 4 * 4, 256);
    sym_samp_off = align_up(data_b + flags_b + cand_b + 256, 256);
    sym_bytes = sym_samp_off + tp_sample_area_bytes(tp_size, ar_cap_rows);
    if (const char *e = getenv("DIS_TP_SAMPLE")) tp_sample_fused = strcmp(e, "gather") != 0;
    // Prefill-sized passes (more rows than the decode region holds) go through ncclAllReduce (reading the projection
    // output in place when it is one slice) + the norm kernel.  DIS_TP_PREFILL_FUSED=1 gives them their own region of the
    // symmetric buffer and a fused peer-memory exchange instead (one kernel, 2 (tp-1)/tp * T*H*4 bytes on the wire),
    // parity-green (tests/test_tp_gpu.py, and bench.py's gate at TP=2 and 8).  Measured, TTFT of 64 x 512 tokens:
    //   row kernel (one 1024-thread CTA per row, two release fences each)   TP=2 451 ms            NCCL 323
    //   streaming kernel (one resident CTA per SM, two flag flights in all) TP=2 309 ms  TP=8 189  NCCL 319 / 183
    // The streaming kernel is at the wire's bandwidth and NCCL (in-switch reduction at 8 ranks) is too: a tie at TP=2, NCCL
    // ahead at TP=8, so NCCL stays the default for these passes; DIS_TP_PREFILL_ROWKERNEL=1 selects the row kernel.
    if (cap_tokens > ar_cap_rows && m.hidden % (4 * tp_size) == 0 && getenv("DIS_TP_PREFILL_FUSED") && atoi(getenv("DIS_TP_PREFILL_FUSED"))) {
      ar_cap_big = cap_tokens;
      sym_big_off = align_up(sym_bytes, 256);
      sym_bytes = sym_big_off + (size_t)4 * ar_cap_big * m.hidden * sizeof(float) + align_up((size_t)2 * 2 * tp_size * ar_cap_big * 4, 256);
    }
    ENG_CUDA(dev_alloc(&sym, sym_bytes, &bytes_allocated));
    ENG_CUDA(cudaMemset(sym, 0, sym_bytes));
  }
  ENG_CUDA(cudaDeviceSynchronize());
  return DIS_OK;
}

// ---- peer-memory exchange set-up: export this rank's symmetric buffer, import everybody's
int Engine::tp_export(uint8_t out[64]) {
  if (tp_size <= 1 || !sym) return fail(DIS_ERR_INVALID_ARGUMENT, "tp_export: engine is not tensor parallel");
  ENG_CUDA(cudaSetDevice(cfg.device));
  cudaIpcMemHandle_t hdl;
  ENG_CUDA(cudaIpcGetMemHandle(&hdl, sym));
  static_assert(sizeof(hdl) == 64, "cudaIpcMemHandle_t is 64 bytes");
  memcpy(out, &hdl, 64);
  return DIS_OK;
}

int Engine::tp_import(const uint8_t *handles) {
  if (tp_size <= 1 || !sym || !handles) return fail(DIS_ERR_INVALID_ARGUMENT, "tp_import: engine is not tensor parallel");
  // exchange epochs are step * 256 + index of the exchange inside the forward pass (2 per layer, 255 = candidate merge)
  if (2 * m.layers >= 255) return fail(DIS_ERR_INVALID_ARGUMENT, "tp_import: the peer-memory exchange supports at most 127 layers (NCCL path remains)");
  ENG_CUDA(cudaSetDevice(cfg.device));
  const size_t data_b = (size_t)std::max(2 * tp_size, 8) * ar_cap_rows * m.hidden * sizeof(float);  // 8: flag-in-data two-shot areas
  const size_t flags_b = align_up((size_t)2 * 2 * tp_size * ar_cap_rows * 4, 256);  // two hops
  const size_t cand_b = align_up((size_t)2 * tp_size * ar_cap_rows * 4 * 4, 256);
  for (int r = 0; r < tp_size; r++) {
    uint8_t *base = sym;
    if (r != tp_rank) {
      cudaIpcMemHandle_t hdl;
      memcpy(&hdl, handles + (size_t)r * 64, 64);
      void *p = nullptr;
      ENG_CUDA(cudaIpcOpenMemHandle(&p, hdl, cudaIpcMemLazyEnablePeerAccess));
      peer_base[r] = p;
      base = static_cast<uint8_t *>(p);
    }
    peers.data[r] = reinterpret_cast<float *>(base);
    peers.flags[r] = reinterpret_cast<uint32_t *>(base + data_b);
    peers.cand[r] = reinterpret_cast<uint32_t *>(base + data_b + flags_b);
    peers.cand_flags[r] = reinterpret_cast<uint32_t *>(base + data_b + flags_b + cand_b);
    samp_peers.buf[r] = reinterpret_cast<uint32_t *>(base + sym_samp_off);
  }
  samp_peers.rank = tp_rank;
  samp_peers.tp = tp_size;
  samp_peers.cap_rows = ar_cap_rows;
  peers.rank = tp_rank;
  peers.tp = tp_size;
  peers.cap_rows = ar_cap_rows;
  if (ar_cap_big > 0) {
    peers_big = peers;
    for (int r = 0; r < tp_size; r++) {
      uint8_t *base = reinterpret_cast<uint8_t *>(peers.data[r]) + sym_big_off;
      peers_big.data[r] = reinterpret_cast<float *>(base);
      peers_big.flags[r] = reinterpret_cast<uint32_t *>(base + (size_t)4 * ar_cap_big * m.hidden * sizeof(float));
    }
    peers_big.cap_rows = ar_cap_big;
    peers_big.two_shot = getenv("DIS_TP_PREFILL_ROWKERNEL") && atoi(getenv("DIS_TP_PREFILL_ROWKERNEL")) ? 1 : 4;
  }
  // one-shot (tp-1)*T*H*4 bytes and one flag hop, or two-shot 2*(tp-1)/tp*T*H*4 bytes and two hops
  // measured at batch 64, H = 4096, ms per step (one-shot / fenced two-shot / flag-in-data two-shot):
  //   TP=2  3.44 / 3.76 / 3.28      TP=4  2.80 / 3.00 / 2.64      TP=8  2.88 / 2.67 / 2.42
  // round 2, flag-in-data row kernel (ll) against the wide-grid kernel:
  //   H = 4096 (8B, batch 64)   TP=2 2.942 / 2.941   TP=4 2.424 / 2.464   TP=8 2.390 / 2.509   -> ll
  //   H = 8192 (70B, batch 32)  TP=8 7.986 / 7.667                                             -> wide: a row that needs two
  //   float4 per thread in the row kernel is where spreading the local work over more CTAs starts to pay
  // The choice is fixed per engine (the kernels lay the symmetric buffer out differently).
  const bool blocks_ok = m.hidden % (4 * tp_size) == 0;
  const bool wide_ok = blocks_ok && (m.hidden / tp_size) % 128 == 0;
  peers.two_shot = blocks_ok ? (wide_ok && m.hidden > 4096 ? 3 : 2) : 0;
  if (const char *e = getenv("DIS_TP_EXCHANGE")) {
    if (!strcmp(e, "oneshot")) peers.two_shot = 0;
    else if (!strcmp(e, "twoshot") && blocks_ok) peers.two_shot = 1;
    else if (!strcmp(e, "ll") && blocks_ok) peers.two_shot = 2;
    else if (!strcmp(e, "wide") && wide_ok) peers.two_shot = 3;
  }
  tp_custom = true;
  return DIS_OK;
}

Engine::~Engine() {
  if (cfg.device >= 0) cudaSetDevice(cfg.device);
  if (stream) cudaStreamSynchronize(stream);
  for (auto &g : graphs) cudaGraphExecDestroy(g.second);
  for (auto ev : prof_events) cudaEventDestroy(ev);
  if (comm && nccl) nccl->CommDestroy(comm);
  for (int r = 0; r < kTpMaxRanks; r++) if (peer_base[r]) cudaIpcCloseMemHandle(peer_base[r]);
  void *ptrs[] = {embed, wqkv, wo, wgu, wdown, lm_head, attn_norm, mlp_norm, final_norm, x, partial, logits, ar_buf,
                  attn_scratch, full_logits, gather_buf, cand_val, cand_val_all, cand_tok, cand_tok_all, cand_sum, cand_sum_all, xn, q, attn,
                  act, d_out_tokens, kv_pool, d_page_table, rope_cos, rope_sin, meta_dec.dev, meta_pre.dev, sym};
  for (void *p : ptrs) if (p) cudaFree(p);
  if (h_out_tokens) cudaFreeHost(h_out_tokens);
  if (h_page_table) cudaFreeHost(h_page_table);
  if (meta_dec.host) cudaFreeHost(meta_dec.host);
  if (meta_pre.host) cudaFreeHost(meta_pre.host);
  if (ev0) cudaEventDestroy(ev0);
  if (ev1) cudaEventDestroy(ev1);
  if (own_stream && stream) cudaStreamDestroy(stream);
}

// ------------------------------------------------------------------------------------ weights
bool Engine::tensor_slot(int kind, int layer, TensorSlot *o) {
  const size_t H = m.hidden, d = m.head_dim, l = (size_t)layer;
  const size_t qd = (size_t)m.n_heads * d, kd = (size_t)m.n_kv_heads * d;
  const bool per_layer = !(kind == DIS_T_EMBED || kind == DIS_T_FINAL_NORM || kind == DIS_T_LM_HEAD);
  if (per_layer && (layer < 0 || layer >= m.layers)) return false;
  TensorSlot s{};
  s.f32 = false;
  switch (kind) {
    case DIS_T_EMBED: s = {embed, (size_t)m.vocab, H, H, 0, 0, (size_t)m.vocab, H, false}; break;
    case DIS_T_LM_HEAD: s = {lm_head, (size_t)vocab_l, H, H, (size_t)vocab_off, 0, (size_t)m.vocab, H, false}; break;
    case DIS_T_FINAL_NORM: s = {final_norm, 1, H, H, 0, 0, 1, H, true}; break;
    case DIS_T_ATTN_NORM: s = {attn_norm + l * H, 1, H, H, 0, 0, 1, H, true}; break;
    case DIS_T_MLP_NORM: s = {mlp_norm + l * H, 1, H, H, 0, 0, 1, H, true}; break;
    case DIS_T_WQ: s = {wqkv + l * nqkv_l * H, (size_t)qd_l, H, H, (size_t)tp_rank * qd_l, 0, qd, H, false}; break;
    case DIS_T_WK: s = {wqkv + (l * nqkv_l + qd_l) * H, (size_t)kvd_l, H, H, (size_t)tp_rank * kvd_l, 0, kd, H, false}; break;
    case DIS_T_WV: s = {wqkv + (l * nqkv_l + qd_l + kvd_l) * H, (size_t)kvd_l, H, H, (size_t)tp_rank * kvd_l, 0, kd, H, false}; break;
    case DIS_T_WO: s = {wo + l * H * qd_l, H, (size_t)qd_l, (size_t)qd_l, 0, (size_t)tp_rank * qd_l, H, qd, false}; break;
    case DIS_T_WGATE: s = {wgu + l * 2 * inter_l * H, (size_t)inter_l, H, H, (size_t)tp_rank * inter_l, 0, (size_t)m.inter, H, false}; break;
    case DIS_T_WUP: s = {wgu + (l * 2 * inter_l + inter_l) * H, (size_t)inter_l, H, H, (size_t)tp_rank * inter_l, 0, (size_t)m.inter, H, false}; break;
    case DIS_T_WDOWN: s = {wdown + l * H * inter_l, H, (size_t)inter_l, (size_t)inter_l, 0, (size_t)tp_rank * inter_l, H, (size_t)m.inter, false}; break;
    default: return false;
  }
  *o = s;
  return true;
}

int Engine::load_tensor(int kind, int layer, const void *host, int dtype) {
  if (shut) return fail(DIS_ERR_SHUTDOWN, "engine is shut down");
  if (!host) return fail(DIS_ERR_INVALID_ARGUMENT, "load_tensor: null host pointer");
  if (dtype != DIS_F32 && dtype != DIS_BF16) return fail(DIS_ERR_INVALID_ARGUMENT, "load_tensor: dtype must be f32 or bf16");
  TensorSlot s;
  if (!tensor_slot(kind, layer, &s)) return fail(DIS_ERR_INVALID_ARGUMENT, "load_tensor: bad tensor kind or layer");
  ENG_CUDA(cudaSetDevice(cfg.device));
  const size_t n = s.rows * s.cols;
  if (s.f32) {
    std::vector<float> tmp(n);
    for (size_t r = 0; r < s.rows; r++)
      for (size_t c = 0; c < s.cols; c++) {
        const size_t src = (s.row0 + r) * s.full_cols + s.col0 + c;
        const float v = dtype == DIS_F32 ? static_cast<const float *>(host)[src]
                                         : bf16_bits_to_float(static_cast<const uint16_t *>(host)[src]);
        tmp[r * s.cols + c] = bf16_bits_to_float(float_to_bf16_bits(v));  // gains are bf16 values held in fp32
      }
    ENG_CUDA(cudaMemcpy(s.ptr, tmp.data(), n * 4, cudaMemcpyHostToDevice));
  } else {
    std::vector<uint16_t> tmp(n);
    for (size_t r = 0; r < s.rows; r++) {
      const size_t src0 = (s.row0 + r) * s.full_cols + s.col0;
      if (dtype == DIS_BF16) memcpy(&tmp[r * s.cols], static_cast<const uint16_t *>(host) + src0, s.cols * 2);
      else {
        const float *f = static_cast<const float *>(host) + src0;
        for (size_t c = 0; c < s.cols; c++) tmp[r * s.cols + c] = float_to_bf16_bits(f[c]);
      }
    }
    ENG_CUDA(cudaMemcpy(s.ptr, tmp.data(), n * 2, cudaMemcpyHostToDevice));
  }
  return DIS_OK;
}

// amplitude / offset convention of the synthetic model (independently restated in the oracle)
static void synth_params(const dis_model_config &m, int kind, float *amp, float *off) {
  const float a = 0.034641016f;  // 0.02 * sqrt(3): uniform(-a, a) has std 0.02
  *off = 0.f;
  switch (kind) {
    case DIS_T_ATTN_NORM: case DIS_T_MLP_NORM: case DIS_T_FINAL_NORM: *amp = 0.1f; *off = 1.0f; break;
    case DIS_T_WO: case DIS_T_WDOWN: *amp = a / sqrtf(2.0f * (float)m.layers); break;
    default: *amp = a; break;
  }
}

int Engine::load_synthetic(uint64_t seed) {
  if (shut) return fail(DIS_ERR_SHUTDOWN, "engine is shut down");
  ENG_CUDA(cudaSetDevice(cfg.device));
  uint16_t *tmp = nullptr;
  uint64_t dummy = 0;
  ENG_CUDA(dev_alloc(&tmp, (size_t)m.hidden, &dummy));
  const int globals[3] = {DIS_T_EMBED, DIS_T_FINAL_NORM, DIS_T_LM_HEAD};
  const int per_layer[9] = {DIS_T_ATTN_NORM, DIS_T_WQ, DIS_T_WK, DIS_T_WV, DIS_T_WO, DIS_T_MLP_NORM, DIS_T_WGATE, DIS_T_WUP, DIS_T_WDOWN};
  auto gen = [&](int kind, int layer) -> cudaError_t {
    TensorSlot s;
    tensor_slot(kind, layer, &s);
    float amp, off;
    synth_params(m, kind, &amp, &off);
    const bool pl = !(kind == DIS_T_EMBED || kind == DIS_T_FINAL_NORM || kind == DIS_T_LM_HEAD);
    const size_t n = s.rows * s.cols;
    const size_t idx0 = s.row0 * s.full_cols;
    if (s.f32) {
      RET_CUDA(launch_synth_bf16(tmp, n, seed, kind, pl ? layer : 0, amp, off, idx0, s.cols, s.full_cols, s.col0, stream));
      return launch_bf16_to_f32(tmp, static_cast<float *>(s.ptr), n, stream);
    }
    return launch_synth_bf16(static_cast<uint16_t *>(s.ptr), n, seed, kind, pl ? layer : 0, amp, off, idx0, s.cols,
                             s.full_cols, s.col0, stream);
  };
  for (int k : globals) ENG_CUDA(gen(k, 0));
  for (int l = 0; l < m.layers; l++)
    for (int k : per_layer) ENG_CUDA(gen(k, l));
  ENG_CUDA(cudaStreamSynchronize(stream));
  cudaFree(tmp);
  return DIS_OK;
}

int Engine::read_tensor(int kind, int layer, uint16_t *host, size_t *n) {
  TensorSlot s;
  if (!n || !tensor_slot(kind, layer, &s)) return fail(DIS_ERR_INVALID_ARGUMENT, "read_tensor: bad argument");
  const size_t need = s.rows * s.cols;
  if (!host || *n < need) { *n = need; return fail(DIS_ERR_INVALID_ARGUMENT, "read_tensor: buffer too small"); }
  *n = need;
  ENG_CUDA(cudaSetDevice(cfg.device));
  ENG_CUDA(cudaStreamSynchronize(stream));
  if (s.f32) {
    std::vector<float> tmp(need);
    ENG_CUDA(cudaMemcpy(tmp.data(), s.ptr, need * 4, cudaMemcpyDeviceToHost));
    for (size_t i = 0; i < need; i++) host[i] = float_to_bf16_bits(tmp[i]);
  } else {
    ENG_CUDA(cudaMemcpy(host, s.ptr, need * 2, cudaMemcpyDeviceToHost));
  }
  return DIS_OK;
}

int Engine::finalize() {
  if (shut) return fail(DIS_ERR_SHUTDOWN, "engine is shut down");
  loaded = true;
  return DIS_OK;
}

// ------------------------------------------------------------------------------------ sequences
bool Engine::ensure_pages(Sequence &s, int slot, int n_tokens) {
  const int need = (n_tokens + page_size - 1) / page_size;
  if (need > max_pages) return false;
  if (need - (int)s.pages.size() > pages_ix.available()) return false;  // nothing is evicted for a request that cannot be admitted
  while ((int)s.pages.size() < need) {
    int p;
    if (!pages_ix.alloc(&p)) return false;
    h_page_table[(size_t)slot * max_pages + s.pages.size()] = p;
    s.pages.push_back(p);
    pt_dirty = true;
  }
  return true;
}

void Engine::release_pages(Sequence &s) {
  pages_ix.release(s.pages);
  s.pages.clear();
}

int Engine::seq_add(const uint32_t *prompt, int n, const dis_sampling_params *p, int *slot_out) {
  if (shut) return fail(DIS_ERR_SHUTDOWN, "engine is shut down");
  if (!loaded) return fail(DIS_ERR_MODEL_NOT_LOADED, "model not loaded");
  if (!prompt || n <= 0 || !p || !slot_out) return fail(DIS_ERR_INVALID_ARGUMENT, "seq_add: empty prompt or null argument");
  if (n >= m.max_ctx) return fail(DIS_ERR_INVALID_ARGUMENT, "seq_add: prompt does not fit the context window");
  for (int i = 0; i < n; i++)
    if (prompt[i] >= (uint32_t)m.vocab) return fail(DIS_ERR_INVALID_ARGUMENT, "seq_add: token id out of vocabulary");
  if (p->temperature < 0.f || p->top_p < 0.f || p->top_p > 1.f || p->max_tokens < 1)
    return fail(DIS_ERR_INVALID_ARGUMENT, "seq_add: bad sampling parameters");
  int slot = -1;
  for (int i = 0; i < max_batch; i++) if (!seqs[(size_t)i].used) { slot = i; break; }
  if (slot < 0) return fail(DIS_ERR_OUT_OF_MEMORY, "seq_add: all sequence slots are in use");
  Sequence &s = seqs[(size_t)slot];
  s = Sequence();
  // prefix reuse (design.md:396 get_prefix, Property 9): the longest registered prefix, in whole pages, is mapped
  // into this sequence's page table; prefill starts behind it
  const int hit = pages_ix.match(prompt, n, &s.pages);
  for (int i = 0; i < hit; i++) h_page_table[(size_t)slot * max_pages + i] = s.pages[(size_t)i];
  if (hit) pt_dirty = true;
  s.cached = hit * page_size;
  s.prefill_pos = s.cached;
  if (!ensure_pages(s, slot, n + 1)) { release_pages(s); return fail(DIS_ERR_OUT_OF_MEMORY, "seq_add: out of KV pages"); }
  s.used = true;
  s.prompt.assign(prompt, prompt + n);
  s.max_tokens = p->max_tokens;
  s.temperature = p->temperature;
  s.top_p = p->top_p;
  s.seed = p->seed;
  size_t off = 0;
  for (int i = 0; i < p->n_stop; i++) {
    const int len = p->stop_lens ? p->stop_lens[i] : 0;
    if (len > 0 && p->stop_tokens) s.stop.emplace_back(p->stop_tokens + off, p->stop_tokens + off + len);
    off += (size_t)std::max(len, 0);
  }
  *slot_out = slot;
  return DIS_OK;
}

int Engine::seq_remove(int slot) {
  if (slot < 0 || slot >= max_batch || !seqs[(size_t)slot].used) return fail(DIS_ERR_INVALID_ARGUMENT, "seq_remove: no such sequence");
  Sequence &s = seqs[(size_t)slot];
  if (pages_ix.enabled() && s.prefilled) {
    // what this sequence generated becomes reusable too (a follow-up turn that quotes it): tokens whose K/V are
    // in the cache = prompt + all sampled tokens but the last
    std::vector<uint32_t> toks(s.prompt);
    toks.insert(toks.end(), s.gen.begin(), s.gen.end());
    if ((int)toks.size() > s.len) toks.resize((size_t)s.len);
    pages_ix.commit(toks.data(), (int)toks.size() / page_size, s.pages);
  }
  release_pages(s);
  s = Sequence();
  return DIS_OK;
}

void Engine::finish_check(Sequence &s, uint32_t tok) {
  s.generated++;
  tokens_generated++;
  s.last_token = tok;
  if (pages_ix.enabled()) s.gen.push_back(tok);
  s.tail.push_back(tok);
  size_t longest = 0;
  for (auto &st : s.stop) longest = std::max(longest, st.size());
  if (s.tail.size() > longest && longest > 0) s.tail.erase(s.tail.begin(), s.tail.end() - (long)longest);
  if (longest == 0) s.tail.clear();
  if (cfg.eos_token >= 0 && tok == (uint32_t)cfg.eos_token) { s.finish = DIS_FINISH_STOP; return; }
  for (auto &st : s.stop)
    if (st.size() <= s.tail.size() && std::equal(st.begin(), st.end(), s.tail.end() - (long)st.size())) { s.finish = DIS_FINISH_STOP_SEQUENCE; return; }
  if (s.generated >= s.max_tokens) { s.finish = DIS_FINISH_LENGTH; return; }
  if (s.len + 1 >= m.max_ctx) { s.finish = DIS_FINISH_LENGTH; return; }
}

// ------------------------------------------------------------------------------------ step plumbing
cudaError_t Engine::upload_step(int T, int R) {
  if (pt_dirty) {
    RET_CUDA(cudaMemcpyAsync(d_page_table, h_page_table, (size_t)max_batch * max_pages * sizeof(int), cudaMemcpyHostToDevice, stream));
    pt_dirty = false;
  }
  (void)T; (void)R;
  if (((++ar_step) & 0xFFFFFFu) == 0) ++ar_step;  // exchange epochs are (step << 8) + index: never 0, the value of an untouched flag
  cur_meta->h<uint32_t>(cur_meta->o_arstep)[0] = ar_step;
  ar_idx = 0;
  return cudaMemcpyAsync(cur_meta->dev, cur_meta->host, cur_meta->bytes, cudaMemcpyHostToDevice, stream);
}

int Engine::pick_attn_splits(int T, int max_ctx_in_batch) const {
  const int ctas = T * nkv_l;
  // two CTAs fit per SM; split the history only as far as everything still runs in ONE wave (measured on a TP=4
  // shard, 128 (token, kv head) pairs: 3 splits = 384 CTAs = two waves 2.08 ms per step, 2 splits 1.90 ms)
  int want = std::max(1, (2 * kNumSMs) / ctas);
  const int cap = std::max(1, max_ctx_in_batch / 128);
  want = std::min(std::min(want, cap), 32);
  if (max_attn_splits > 0) want = std::min(want, max_attn_splits);  // DIS_ATTN_MAX_SPLITS (A/B runs)
  while ((size_t)T * nh_l * want * (2 + m.head_dim) > attn_scratch_elems && want > 1) want--;
  return std::max(want, 1);
}

// Up to this many gate (= up) rows per rank the gate-up projection runs split-K over the fused [2I][K] weight with
// the SiLU-gate applied while the slices are summed; above it the single-pass SiLU-fused tile wins.  Measured per
// step: TP=2 shard (I = 7168, batch 64) 3.04 -> 2.6 ms; one GPU (I = 14336) batch 32 3.53 vs 3.53, batch 64 4.33 vs 4.27.
static int gu_split_max_i() { return 8192; }

static int tile_index(int T) { const int t = gemm_token_tile(T); return t == 32 ? 0 : (t == 64 ? 1 : (t == 128 ? 2 : 3)); }

// out = partial[S][T][N] of  X[T][K] * W[w_row0 : w_row0+N][K]^T
cudaError_t Engine::linear_partial(const TmaMap *xmaps, const uint16_t *xptr, const TmaMap &wmap, const uint16_t *wptr,
                                   int w_row0, int T, int N, int K, int *split_out) {
  const bool gemv = (cfg.flags & DIS_FLAG_FORCE_GEMV) || T <= gemv_max_t;
  if (gemv) {
    *split_out = 1;
    kernel_launches += (uint64_t)((T + kGemvMaxT - 1) / kGemvMaxT);
    return launch_gemv(xptr, wptr + (size_t)w_row0 * K, partial, T, N, K, stream);
  }
  int s = gemm_pick_split_k(T, N, K);
  if ((size_t)s * T * N > partial_elems) s = 1;
  *split_out = s;
  kernel_launches++;
  if (s == 1 && use_persist && gemm_use_persist(T, N, false)) return launch_gemm_persist_partial(wmap, xmaps[3], partial, T, N, K, w_row0, stream);
  return launch_gemm_partial(wmap, xmaps[tile_index(T)], partial, T, N, K, w_row0, s, stream);
}

// Tensor parallel: sum the split-K slices, all-reduce over ranks; afterwards the data is ar_buf[T][H].
cudaError_t Engine::allreduce_partials(int split, int T) {
  const long long n = (long long)T * m.hidden;
  prof_mark(P_ALLREDUCE);
  const float *src = ar_buf;
  if (split == 1) src = partial;  // prefill-sized GEMMs do not split K: NCCL reads the projection output where it lies
  else { RET_CUDA(launch_reduce_partials(partial, split, n, ar_buf, n, stream)); kernel_launches++; }
  kernel_launches++;
  const int r = nccl->AllReduce(src, ar_buf, (size_t)n, /*ncclFloat32*/ 7, /*ncclSum*/ 0, comm, stream);
  if (r != 0) { set_last_error(std::string("ncclAllReduce: ") + nccl->GetErrorString(r)); return cudaErrorUnknown; }
  return cudaSuccess;
}

// x (+)= the pending projection output (summed over split-K slices and, under tensor parallelism, over the
// ranks); xn = bf16(RMSNorm(x) * gain).  With row_map only `rows` gathered rows are normalised (logits rows).
cudaError_t Engine::consume_pending(const float *gain, const int *row_map, int rows, int T) {
  const int H = m.hidden;
  if (pend_needs_ar) {
    // fused exchange over peer memory: split-K reduce + all-reduce + residual + norm in one kernel
    const int idx = ar_idx++;
    RET_CUDA(launch_tp_add_rmsnorm(T <= ar_cap_rows ? peers : peers_big, x, pend_ptr, pend_count, pend_str, gain, xn,
                                   cur_meta->d<uint32_t>(cur_meta->o_arstep), (uint32_t)idx, idx & 1, T, H, m.rms_eps, stream));
    kernel_launches++;
    pend_needs_ar = false; pend_ptr = nullptr; pend_count = 0;
    if (!row_map) return cudaSuccess;
    kernel_launches++;  // x is complete: gather the rows that need logits
    return launch_add_rmsnorm(x, nullptr, nullptr, nullptr, 0, 0, gain, xn, row_map, rows, H, m.rms_eps, stream);
  }
  kernel_launches++;
  const cudaError_t r = launch_add_rmsnorm(x, nullptr, nullptr, pend_ptr, pend_count, pend_str, gain, xn, row_map, rows, H,
                                           m.rms_eps, stream);
  pend_ptr = nullptr; pend_count = 0;
  return r;
}

cudaError_t Engine::enqueue_layers(int T, int n_splits) {
  const int H = m.hidden, d = m.head_dim;
  const bool gemv = (cfg.flags & DIS_FLAG_FORCE_GEMV) || T <= gemv_max_t;
  const bool fused_ar = tp_size > 1 && tp_custom && (T <= ar_cap_rows || T <= ar_cap_big);  // exchange over peer memory, no NCCL kernel
  const bool pdl_ok = use_pdl && !profiling && (tp_size == 1 || fused_ar);
  // the PDL switch is thread-local state of the launch helper: whatever way this function is left, later
  // kernel-level calls on this host thread must not inherit it
  struct PdlOff { ~PdlOff() { pdl_enabled() = false; } } pdl_off;
  pend_ptr = nullptr; pend_count = 0; pend_str = 0; pend_needs_ar = false;
  auto skip = [&](int cls) { return n_qblocks == 0 && tp_size == 1 && (skip_mask >> cls) & 1u; };
  const uint32_t *tokens = cur_meta->d<uint32_t>(cur_meta->o_tokens);
  const int *pos = cur_meta->d<int>(cur_meta->o_pos), *ptrow = cur_meta->d<int>(cur_meta->o_ptrow), *ctx = cur_meta->d<int>(cur_meta->o_ctx);
  // a projection output becomes "pending": the next norm kernel adds it to the residual
  auto set_pending = [&](int split) -> cudaError_t {
    pend_ptr = partial; pend_count = split; pend_str = (long long)T * H;
    if (tp_size > 1) {
      if (fused_ar) pend_needs_ar = true;
      else { RET_CUDA(allreduce_partials(split, T)); pend_ptr = ar_buf; pend_count = 1; pend_str = 0; }
    }
    return cudaSuccess;
  };
  for (int l = 0; l < m.layers; l++) {
    prof_mark(P_NORM);
    if (l == 0) {
      pdl_enabled() = false;  // the step's first kernel follows a memcpy
      RET_CUDA(launch_add_rmsnorm(x, tokens, embed, nullptr, 0, 0, attn_norm, xn, nullptr, T, H, m.rms_eps, stream));
      kernel_launches++;
    } else {
      pdl_enabled() = pdl_ok;
      if (!skip(P_NORM)) RET_CUDA(consume_pending(attn_norm + (size_t)l * H, nullptr, T, T));
    }
    pdl_enabled() = pdl_ok;
    int s_qkv = 1;
    prof_mark(P_QKV);
    if (!skip(P_QKV)) RET_CUDA(linear_partial(xn_map, xn, wqkv_map, wqkv, l * nqkv_l, T, nqkv_l, H, &s_qkv));
    else s_qkv = gemm_pick_split_k(T, nqkv_l, H);
    uint16_t *pool = kv_pool + (size_t)l * kv_layer_stride;
    if (n_qblocks > 0) {
      prof_mark(P_ROPE);
      RET_CUDA(launch_rope_kv_write(partial, s_qkv, (long long)T * nqkv_l, q, pool, d_page_table, max_pages, ptrow, pos,
                                    rope_cos, rope_sin, T, nh_l, nkv_l, d, page_size, stream));
      prof_mark(P_ATTN);
      RET_CUDA((pf_tc ? launch_attn_prefill_tc : launch_attn_prefill)(kv_map, (long long)l * kv_rows_per_layer, q, d_page_table, max_pages,
                                   cur_meta->d<int>(cur_meta->o_qb_tok0), cur_meta->d<int>(cur_meta->o_qb_len),
                                   cur_meta->d<int>(cur_meta->o_qb_pos0), cur_meta->d<int>(cur_meta->o_qb_ptrow), attn,
                                   n_qblocks, nh_l, nkv_l, d, page_size, stream));
      kernel_launches += 2;
    } else {
      // decode: RoPE + KV write live inside the attention kernel (one kernel boundary less per layer)
      prof_mark(P_ATTN);
      if (!skip(P_ATTN)) RET_CUDA(launch_attn_decode_fused(kv_map, (long long)l * kv_rows_per_layer, pool, partial, s_qkv, (long long)T * nqkv_l,
                                        pos, rope_cos, rope_sin, d_page_table, max_pages, ptrow, ctx, attn, attn_scratch,
                                        n_splits, T, nh_l, nkv_l, d, page_size, stream));
      kernel_launches += n_splits > 1 ? 2 : 1;
    }
    int s_o = 1;
    prof_mark(P_OPROJ);
    if (!skip(P_OPROJ)) RET_CUDA(linear_partial(attn_map, attn, wo_map, wo, l * H, T, H, qd_l, &s_o));
    else s_o = gemm_pick_split_k(T, H, qd_l);
    RET_CUDA(set_pending(s_o));
    prof_mark(P_NORM);
    if (!skip(P_NORM)) RET_CUDA(consume_pending(mlp_norm + (size_t)l * H, nullptr, T, T));
    prof_mark(P_GATEUP);
    if (gemv) {
      RET_CUDA(launch_gemv_silu(xn, wgu + (size_t)l * 2 * inter_l * H, act, T, inter_l, H, stream));
      kernel_launches += (uint64_t)((T + kGemvMaxT - 1) / kGemvMaxT);
    } else if (use_persist && gemm_use_persist(T, inter_l, true)) {
      RET_CUDA(launch_gemm_persist_silu(wgu_map64, xn_map[3], act, T, inter_l, H, l * 2 * inter_l, stream));
      kernel_launches++;
    } else if (gemm_token_tile(T) >= 128) {
      // prefill is tensor-bound: full 128-row tiles (the balanced, shorter tiles would idle MMA rows)
      RET_CUDA(launch_gemm_silu(wgu_map128, xn_map[tile_index(T)], act, T, inter_l, H, l * 2 * inter_l, 128, stream));
      kernel_launches++;
    } else if (inter_l <= gu_split_max_i() && T <= 64 && (size_t)gemm_pick_split_k(T, 2 * inter_l, H) * T * 2 * inter_l <= partial_elems) {
      // small tensor-parallel shard: a single-pass SiLU-fused GEMM would leave each CTA a long, thin K loop;
      // split-K over the fused [2I][K] weight instead, SiLU-gate applied while the slices are summed
      const int sgu = gemm_pick_split_k(T, 2 * inter_l, H);
      RET_CUDA(launch_gemm_partial(wgu_map128, xn_map[tile_index(T)], partial, T, 2 * inter_l, H, l * 2 * inter_l, sgu, stream));
      RET_CUDA(launch_silu_mul_reduce(partial, sgu, (long long)T * 2 * inter_l, act, T, inter_l, stream));
      kernel_launches += 2;
    } else if (!skip(P_GATEUP)) {
      RET_CUDA(launch_gemm_silu(wgu_map, xn_map[tile_index(T)], act, T, inter_l, H, l * 2 * inter_l, gemm_pick_tile_rows(inter_l), stream));
      kernel_launches++;
    }
    int s_d = 1;
    prof_mark(P_DOWN);
    if (!skip(P_DOWN)) RET_CUDA(linear_partial(act_map, act, wdown_map, wdown, l * H, T, H, inter_l, &s_d));
    else s_d = gemm_pick_split_k(T, H, inter_l);
    RET_CUDA(set_pending(s_d));
  }
  pdl_enabled() = false;
  return cudaSuccess;
}

// final norm on the rows that need logits, lm_head, sampling
cudaError_t Engine::enqueue_head(int R, bool all_greedy) {
  const int H = m.hidden;
  prof_mark(P_HEAD_NORM);
  RET_CUDA(consume_pending(final_norm, rowmap_identity ? nullptr : cur_meta->d<int>(cur_meta->o_rowmap), R, last_T));
  const bool gemv = (cfg.flags & DIS_FLAG_FORCE_GEMV) || R <= gemv_max_t;
  prof_mark(P_LM_HEAD);
  if (gemv) { RET_CUDA(launch_gemv(xn, lm_head, logits, R, vocab_l, H, stream)); kernel_launches += (uint64_t)((R + kGemvMaxT - 1) / kGemvMaxT); }
  else { RET_CUDA(launch_gemm_partial(lm_map, xn_map[tile_index(R)], logits, R, vocab_l, H, 0, 1, stream)); kernel_launches++; }
  prof_mark(P_SAMPLE);
  const float *temp = cur_meta->d<float>(cur_meta->o_temp), *topp = cur_meta->d<float>(cur_meta->o_topp);
  const unsigned long long *seed = cur_meta->d<unsigned long long>(cur_meta->o_seed), *step = cur_meta->d<unsigned long long>(cur_meta->o_step);
  float *d_logprob = reinterpret_cast<float *>(d_out_tokens + max_batch);  // TokenEvent.logprob, models.rs:276
  if (tp_size == 1) {
    kernel_launches++;
    if (!all_greedy && sampler_buckets) {
      // temperature / nucleus rows: the 16-ary bucket search of k_tp_sample.cu as a single rank (8 passes over the row
      // instead of the 31 of sample_kernel's bisection; nothing is exchanged).  DIS_SAMPLER=bisect keeps sample_kernel.
      TpSampPeers solo{};
      solo.rank = 0; solo.tp = 1; solo.cap_rows = max_batch;
      return launch_tp_sample(solo, logits, vocab_l, R, vocab_l, temp, topp, seed, step, nullptr, 0, d_out_tokens, d_logprob, stream);
    }
    return launch_sample(logits, vocab_l, R, vocab_l, temp, topp, seed, step, d_out_tokens, 0, nullptr, nullptr, d_logprob, stream);
  }
  if (all_greedy && tp_custom) {
    // vocab-parallel argmax: local (value, id, sum-exp), exchanged and merged over peer memory in one kernel
    RET_CUDA(launch_sample(logits, vocab_l, R, vocab_l, nullptr, nullptr, nullptr, nullptr, cand_tok, vocab_off, cand_val, cand_sum, nullptr, stream));
    kernel_launches += 2;
    return launch_tp_argmax_merge(peers, cand_val, cand_tok, cand_sum, cur_meta->d<uint32_t>(cur_meta->o_arstep), R, d_out_tokens,
                                  d_logprob, stream);
  }
  if (all_greedy) {
    // vocab-parallel argmax: local (value, id, sum-exp) -> all-gather -> merge
    RET_CUDA(launch_sample(logits, vocab_l, R, vocab_l, nullptr, nullptr, nullptr, nullptr, cand_tok, vocab_off, cand_val, cand_sum, nullptr, stream));
    int r = nccl->AllGather(cand_val, cand_val_all, (size_t)R, /*ncclFloat32*/ 7, comm, stream);
    if (r == 0) r = nccl->AllGather(cand_tok, cand_tok_all, (size_t)R, /*ncclUint32*/ 3, comm, stream);
    if (r == 0) r = nccl->AllGather(cand_sum, cand_sum_all, (size_t)R, /*ncclFloat32*/ 7, comm, stream);
    if (r != 0) { set_last_error(std::string("ncclAllGather: ") + nccl->GetErrorString(r)); return cudaErrorUnknown; }
    kernel_launches += 5;
    return launch_argmax_merge(cand_val_all, cand_tok_all, cand_sum_all, tp_size, R, d_out_tokens, d_logprob, stream);
  }
  if (tp_custom && tp_sample_fused && R <= ar_cap_rows) {
    // temperature / nucleus rows (and the greedy rows beside them): the logits stay sharded, the ranks exchange a few
    // scalars per row over peer memory (k_tp_sample.cu) instead of all-gathering R * V fp32 logits
    kernel_launches++;
    return launch_tp_sample(samp_peers, logits, vocab_l, R, vocab_l, temp, topp, seed, step, cur_meta->d<uint32_t>(cur_meta->o_arstep), 0,
                            d_out_tokens, d_logprob, stream);
  }
  RET_CUDA(gather_full_logits(R));
  kernel_launches++;
  return launch_sample(full_logits, m.vocab, R, m.vocab, temp, topp, seed, step, d_out_tokens, 0, nullptr, nullptr, d_logprob, stream);
}

__global__ void interleave_logits_kernel(const float *__restrict__ gathered, float *__restrict__ full, int tp, int R,
                                         int vl) {
  // gathered [tp][R][vl] -> full [R][tp*vl]
  const size_t n = (size_t)tp * R * vl;
  for (size_t i = blockIdx.x * (size_t)blockDim.x + threadIdx.x; i < n; i += (size_t)gridDim.x * blockDim.x) {
    const size_t c = i % vl, r = (i / vl) % R, k = i / ((size_t)vl * R);
    full[r * (size_t)tp * vl + k * vl + c] = gathered[i];
  }
}

cudaError_t Engine::gather_full_logits(int R) {
  const int r = nccl->AllGather(logits, gather_buf, (size_t)R * vocab_l, /*ncclFloat32*/ 7, comm, stream);
  if (r != 0) { set_last_error(std::string("ncclAllGather: ") + nccl->GetErrorString(r)); return cudaErrorUnknown; }
  interleave_logits_kernel<<<kNumSMs * 4, 256, 0, stream>>>(gather_buf, full_logits, tp_size, R, vocab_l);
  kernel_launches += 2;
  return cudaGetLastError();
}

// ------------------------------------------------------------------------------------ prefill
int Engine::prefill(int n, const int *slots, uint32_t *out_tokens, int *out_finish) {
  if (shut) return fail(DIS_ERR_SHUTDOWN, "engine is shut down");
  if (!loaded) return fail(DIS_ERR_MODEL_NOT_LOADED, "model not loaded");
  if (n <= 0 || n > max_batch || !slots || !out_tokens || !out_finish) return fail(DIS_ERR_INVALID_ARGUMENT, "prefill: bad argument");
  const uint64_t stamp = ++call_stamp;  // a slot seen twice in one call carries this call's stamp already
  for (int i = 0; i < n; i++) {
    if (slots[i] < 0 || slots[i] >= max_batch || !seqs[(size_t)slots[i]].used || seqs[(size_t)slots[i]].prefilled)
      return fail(DIS_ERR_INVALID_ARGUMENT, "prefill: sequence missing or already prefilled");
    if (slot_stamp[(size_t)slots[i]] == stamp) return fail(DIS_ERR_INVALID_ARGUMENT, "prefill: duplicate slot");
    slot_stamp[(size_t)slots[i]] = stamp;
  }
  ENG_CUDA(cudaSetDevice(cfg.device));
  ENG_CUDA(cudaEventRecord(ev0, stream));
  cur_meta = &meta_pre;
  struct MetaReset { Engine *e; ~MetaReset() { e->cur_meta = &e->meta_dec; e->n_qblocks = 0; } } meta_reset{this};
  // flatten (sequence, position) pairs and cut them into chunks of at most cap_tokens
  int si = 0, sp = seqs[(size_t)slots[0]].prefill_pos;  // next sequence index / position inside it (behind what is cached already)
  bool all_greedy = true;
  for (int i = 0; i < n; i++) all_greedy = all_greedy && seqs[(size_t)slots[i]].temperature <= 0.f;
  while (si < n) {
    int T = 0, R = 0, QB = 0;
    std::vector<int> owners;
    while (si < n && T < cap_tokens) {
      Sequence &s = seqs[(size_t)slots[si]];
      const int total = (int)s.prompt.size();
      const int take = std::min(total - sp, cap_tokens - T);
      for (int k = 0; k < take; k++) {
        cur_meta->h<uint32_t>(cur_meta->o_tokens)[T + k] = s.prompt[(size_t)(sp + k)];
        cur_meta->h<int>(cur_meta->o_pos)[T + k] = sp + k;
        cur_meta->h<int>(cur_meta->o_ptrow)[T + k] = slots[si];
        cur_meta->h<int>(cur_meta->o_ctx)[T + k] = sp + k + 1;
      }
      // query blocks of up to 16 consecutive tokens for the prefill attention kernel
      for (int k = 0; k < take; k += pf_block) {
        cur_meta->h<int>(cur_meta->o_qb_tok0)[QB] = T + k;
        cur_meta->h<int>(cur_meta->o_qb_len)[QB] = std::min(pf_block, take - k);
        cur_meta->h<int>(cur_meta->o_qb_pos0)[QB] = sp + k;
        cur_meta->h<int>(cur_meta->o_qb_ptrow)[QB] = slots[si];
        QB++;
      }
      T += take;
      sp += take;
      if (sp == total) {
        cur_meta->h<int>(cur_meta->o_rowmap)[R] = T - 1;
        cur_meta->h<float>(cur_meta->o_temp)[R] = s.temperature;
        cur_meta->h<float>(cur_meta->o_topp)[R] = s.top_p;
        cur_meta->h<unsigned long long>(cur_meta->o_seed)[R] = s.seed;
        cur_meta->h<unsigned long long>(cur_meta->o_step)[R] = 0;
        owners.push_back(si);
        R++;
        si++;
        sp = si < n ? seqs[(size_t)slots[si]].prefill_pos : 0;
      }
    }
    n_qblocks = QB;
    last_T = T;
    rowmap_identity = false;
    ENG_CUDA(upload_step(T, R));
    ENG_CUDA(enqueue_layers(T, 1));
    if (R > 0) {
      ENG_CUDA(enqueue_head(R, all_greedy));
      ENG_CUDA(cudaMemcpyAsync(h_out_tokens, d_out_tokens, (size_t)(max_batch + R) * 4, cudaMemcpyDeviceToHost, stream));
    } else if (pend_needs_ar) {
      // A chunk that ends no prompt needs no logits, but its last exchange must still happen: the exchange buffers
      // alternate by the parity of the exchange index, which only lines up across forward passes if every pass
      // performs the same (even) number of exchanges, 2 per layer.
      ENG_CUDA(consume_pending(final_norm, nullptr, T, T));
    }
    ENG_CUDA(cudaStreamSynchronize(stream));  // the pinned metadata block is reused by the next chunk
    for (int r = 0; r < R; r++) {
      out_tokens[owners[(size_t)r]] = h_out_tokens[r];
      seqs[(size_t)slots[owners[(size_t)r]]].last_logprob = reinterpret_cast<const float *>(h_out_tokens + max_batch)[r];
    }
    last_rows = R;
  }
  ENG_CUDA(cudaEventRecord(ev1, stream));
  ENG_CUDA(cudaEventSynchronize(ev1));
  float ms = 0.f;
  ENG_CUDA(cudaEventElapsedTime(&ms, ev0, ev1));
  last_step_ms = ms;
  for (int i = 0; i < n; i++) {
    Sequence &s = seqs[(size_t)slots[i]];
    s.len = (int)s.prompt.size();
    s.prefill_pos = s.len;
    s.prefilled = true;
    pages_ix.commit(s.prompt.data(), s.len / page_size, s.pages);  // full prompt pages become reusable (no-op when off)
    finish_check(s, out_tokens[i]);
    out_finish[i] = s.finish;
  }
  return DIS_OK;
}

// ------------------------------------------------------------------------------------ mixed step
// How one forward pass is shared between running sequences and prompts that are still being read (chunked prefill,
// SURVEY.md §8 row f1): every decode row gets its one token; the remaining room -- at most `budget` prompt tokens and
// never more than `cap` tokens in the pass -- goes to the unfinished prompts IN LIST ORDER: the caller's list is the
// schedule (the batcher lists them by priority class, then shortest prompt first with aging, serving.cpp).  Pure function:
// exported as dis_serving_plan_step so that the rule is property-tested without a GPU.
void plan_mixed_step(int n, const int *remaining_prompt, int budget, int cap, int *take) {
  int decode_rows = 0;
  for (int i = 0; i < n; i++) decode_rows += remaining_prompt[i] <= 0 ? 1 : 0;
  long long room = std::max(0, std::min(budget, cap - decode_rows));
  for (int i = 0; i < n; i++) {
    if (remaining_prompt[i] <= 0) { take[i] = 1; continue; }
    const int t = (int)std::min<long long>(remaining_prompt[i], room);
    take[i] = t;
    room -= t;
  }
}

// One forward pass over decode rows and prompt chunks together.  A listed sequence that is already prefilled
// contributes its last sampled token (exactly what dis_engine_decode would do); one that is not contributes the next
// chunk of its prompt, and samples its first token in the pass that finishes the prompt.  out_produced[i] = 1 where
// out_tokens[i] holds a new token.  The pass runs the prefill kernels: a decode row is a one-token query block.
int Engine::step_mixed(int n, const int *slots, int max_prefill_tokens, uint32_t *out_tokens, int *out_finish, int *out_produced) {
  if (shut) return fail(DIS_ERR_SHUTDOWN, "engine is shut down");
  if (!loaded) return fail(DIS_ERR_MODEL_NOT_LOADED, "model not loaded");
  if (n <= 0 || n > max_batch || !slots || !out_tokens || !out_finish || !out_produced) return fail(DIS_ERR_INVALID_ARGUMENT, "step: bad argument");
  const uint64_t stamp = ++call_stamp;
  for (int i = 0; i < n; i++) {
    if (slots[i] < 0 || slots[i] >= max_batch || !seqs[(size_t)slots[i]].used) return fail(DIS_ERR_INVALID_ARGUMENT, "step: no such sequence");
    if (slot_stamp[(size_t)slots[i]] == stamp) return fail(DIS_ERR_INVALID_ARGUMENT, "step: duplicate slot");
    slot_stamp[(size_t)slots[i]] = stamp;
  }
  ENG_CUDA(cudaSetDevice(cfg.device));
  // rows that cannot run fail alone (Property 22, design.md:812-816)
  std::vector<char> live((size_t)n, 0);
  std::vector<int> idx;
  for (int i = 0; i < n; i++) {
    Sequence &s = seqs[(size_t)slots[i]];
    out_tokens[i] = 0; out_produced[i] = 0; out_finish[i] = s.finish;
    if (s.finish != DIS_FINISH_NONE) continue;
    if (s.prefilled && (s.len + 1 > m.max_ctx || !ensure_pages(s, slots[i], s.len + 1))) { s.finish = DIS_FINISH_FAILED; out_finish[i] = s.finish; continue; }
    live[(size_t)i] = 1;
    idx.push_back(i);
  }
  if (idx.empty()) { last_rows = 0; return DIS_OK; }
  std::vector<int> rem_live, take_live(idx.size());
  bool any_prompt = false;
  for (int i : idx) {
    const Sequence &s = seqs[(size_t)slots[i]];
    rem_live.push_back(s.prefilled ? 0 : (int)s.prompt.size() - s.prefill_pos);
    any_prompt = any_prompt || !s.prefilled;
  }
  if (!any_prompt) {
    // nothing but running sequences: this is a plain decode step -- the CUDA-graph path with the fused decode attention
    const int rc = decode(n, slots, out_tokens, out_finish);
    if (rc != DIS_OK) return rc;
    for (int i = 0; i < n; i++) out_produced[i] = live[(size_t)i] && out_finish[i] != DIS_FINISH_FAILED ? 1 : 0;
    return DIS_OK;
  }
  plan_mixed_step((int)idx.size(), rem_live.data(), max_prefill_tokens > 0 ? max_prefill_tokens : cap_tokens, cap_tokens, take_live.data());
  ENG_CUDA(cudaEventRecord(ev0, stream));
  cur_meta = &meta_pre;
  struct MetaReset { Engine *e; ~MetaReset() { e->cur_meta = &e->meta_dec; e->n_qblocks = 0; } } meta_reset{this};
  int T = 0, R = 0, QB = 0;
  bool all_greedy = true;
  for (size_t k = 0; k < idx.size(); k++) {
    const int i = idx[k];
    Sequence &s = seqs[(size_t)slots[i]];
    const int tk = take_live[k];
    if (tk <= 0) continue;  // no room for this prompt in this pass
    const bool dec = s.prefilled;
    const int p0 = dec ? s.len : s.prefill_pos;
    for (int j = 0; j < tk; j++) {
      cur_meta->h<uint32_t>(cur_meta->o_tokens)[T + j] = dec ? s.last_token : s.prompt[(size_t)(p0 + j)];
      cur_meta->h<int>(cur_meta->o_pos)[T + j] = p0 + j;
      cur_meta->h<int>(cur_meta->o_ptrow)[T + j] = slots[i];
      cur_meta->h<int>(cur_meta->o_ctx)[T + j] = p0 + j + 1;
    }
    for (int j = 0; j < tk; j += pf_block) {
      cur_meta->h<int>(cur_meta->o_qb_tok0)[QB] = T + j;
      cur_meta->h<int>(cur_meta->o_qb_len)[QB] = std::min(pf_block, tk - j);
      cur_meta->h<int>(cur_meta->o_qb_pos0)[QB] = p0 + j;
      cur_meta->h<int>(cur_meta->o_qb_ptrow)[QB] = slots[i];
      QB++;
    }
    T += tk;
    const bool ends = dec || p0 + tk == (int)s.prompt.size();
    if (ends) {
      cur_meta->h<int>(cur_meta->o_rowmap)[R] = T - 1;
      cur_meta->h<float>(cur_meta->o_temp)[R] = s.temperature;
      cur_meta->h<float>(cur_meta->o_topp)[R] = s.top_p;
      cur_meta->h<unsigned long long>(cur_meta->o_seed)[R] = s.seed;
      cur_meta->h<unsigned long long>(cur_meta->o_step)[R] = (unsigned long long)s.generated;
      all_greedy = all_greedy && s.temperature <= 0.f;
      R++;
    }
  }
  if (T == 0) { last_rows = 0; return DIS_OK; }
  n_qblocks = QB;
  last_T = T;
  rowmap_identity = false;
  ENG_CUDA(upload_step(T, R));
  ENG_CUDA(enqueue_layers(T, 1));
  if (R > 0) {
    ENG_CUDA(enqueue_head(R, all_greedy));
    ENG_CUDA(cudaMemcpyAsync(h_out_tokens, d_out_tokens, (size_t)(max_batch + R) * 4, cudaMemcpyDeviceToHost, stream));
  } else if (pend_needs_ar) {
    ENG_CUDA(consume_pending(final_norm, nullptr, T, T));  // keep the exchange count per pass even (see prefill)
  }
  ENG_CUDA(cudaEventRecord(ev1, stream));
  ENG_CUDA(cudaEventSynchronize(ev1));
  float ms = 0.f;
  ENG_CUDA(cudaEventElapsedTime(&ms, ev0, ev1));
  last_step_ms = ms;
  last_rows = R;
  // advance the sequences
  int r = 0;
  for (size_t k = 0; k < idx.size(); k++) {
    const int i = idx[k];
    Sequence &s = seqs[(size_t)slots[i]];
    const int tk = take_live[k];
    if (tk <= 0) continue;
    if (s.prefilled) s.len += 1;
    else {
      s.prefill_pos += tk;
      if (s.prefill_pos < (int)s.prompt.size()) continue;  // more chunks to come
      s.len = (int)s.prompt.size();
      s.prefilled = true;
      pages_ix.commit(s.prompt.data(), s.len / page_size, s.pages);
    }
    out_tokens[i] = h_out_tokens[r];
    s.last_logprob = reinterpret_cast<const float *>(h_out_tokens + max_batch)[r];
    finish_check(s, h_out_tokens[r]);
    out_finish[i] = s.finish;
    out_produced[i] = 1;
    r++;
  }
  return DIS_OK;
}

// ------------------------------------------------------------------------------------ decode
int Engine::decode(int n, const int *slots, uint32_t *out_tokens, int *out_finish) {
  if (shut) return fail(DIS_ERR_SHUTDOWN, "engine is shut down");
  if (!loaded) return fail(DIS_ERR_MODEL_NOT_LOADED, "model not loaded");
  if (n <= 0 || n > max_batch || !slots || !out_tokens || !out_finish) return fail(DIS_ERR_INVALID_ARGUMENT, "decode: bad argument");
  const uint64_t stamp = ++call_stamp;
  for (int i = 0; i < n; i++) {
    if (slots[i] < 0 || slots[i] >= max_batch) return fail(DIS_ERR_INVALID_ARGUMENT, "decode: slot out of range");
    const Sequence &s = seqs[(size_t)slots[i]];
    if (!s.used || !s.prefilled) return fail(DIS_ERR_INVALID_ARGUMENT, "decode: sequence missing or not prefilled");
    if (slot_stamp[(size_t)slots[i]] == stamp) return fail(DIS_ERR_INVALID_ARGUMENT, "decode: duplicate slot");
    slot_stamp[(size_t)slots[i]] = stamp;
  }
  ENG_CUDA(cudaSetDevice(cfg.device));
  // Sequences that cannot take another token (finished, out of context, out of pages) are left out
  // of the batch and reported individually: one bad request must not fail the batch
  // (design.md:812-816, Property 22).
  std::vector<int> live;
  int max_ctx_b = 1;
  bool all_greedy = true;
  for (int i = 0; i < n; i++) {
    Sequence &s = seqs[(size_t)slots[i]];
    out_tokens[i] = 0;
    if (s.finish != DIS_FINISH_NONE) { out_finish[i] = s.finish; continue; }
    if (s.len + 1 > m.max_ctx || !ensure_pages(s, slots[i], s.len + 1)) { s.finish = DIS_FINISH_FAILED; out_finish[i] = s.finish; continue; }
    const int r = (int)live.size();
    cur_meta->h<uint32_t>(cur_meta->o_tokens)[r] = s.last_token;
    cur_meta->h<int>(cur_meta->o_pos)[r] = s.len;
    cur_meta->h<int>(cur_meta->o_ptrow)[r] = slots[i];
    cur_meta->h<int>(cur_meta->o_ctx)[r] = s.len + 1;
    cur_meta->h<int>(cur_meta->o_rowmap)[r] = r;
    cur_meta->h<float>(cur_meta->o_temp)[r] = s.temperature;
    cur_meta->h<float>(cur_meta->o_topp)[r] = s.top_p;
    cur_meta->h<unsigned long long>(cur_meta->o_seed)[r] = s.seed;
    cur_meta->h<unsigned long long>(cur_meta->o_step)[r] = (unsigned long long)s.generated;
    all_greedy = all_greedy && s.temperature <= 0.f;
    max_ctx_b = std::max(max_ctx_b, s.len + 1);
    live.push_back(i);
  }
  const int T_live = (int)live.size();
  if (T_live == 0) { last_rows = 0; return DIS_OK; }
  // One CUDA graph exists per (rows, attention splits): a batch that drains one sequence at a time would capture a new
  // graph at every size.  Above 8 rows the step is padded to the next multiple of 8 with copies of the last live row
  // (same slot, token and position: the copies recompute and rewrite exactly the same K/V row and their tokens are
  // ignored), so at most max_batch / 8 + 8 sizes are ever captured.
  int T = T_live;
  if (T > 8 && !(cfg.flags & DIS_FLAG_NO_GRAPH) && !profiling) T = std::min((T + 7) / 8 * 8, max_batch);
  for (int r = T_live; r < T; r++) {
    const int q = T_live - 1;
    cur_meta->h<uint32_t>(cur_meta->o_tokens)[r] = cur_meta->h<uint32_t>(cur_meta->o_tokens)[q];
    cur_meta->h<int>(cur_meta->o_pos)[r] = cur_meta->h<int>(cur_meta->o_pos)[q];
    cur_meta->h<int>(cur_meta->o_ptrow)[r] = cur_meta->h<int>(cur_meta->o_ptrow)[q];
    cur_meta->h<int>(cur_meta->o_ctx)[r] = cur_meta->h<int>(cur_meta->o_ctx)[q];
    cur_meta->h<int>(cur_meta->o_rowmap)[r] = r;
    cur_meta->h<float>(cur_meta->o_temp)[r] = cur_meta->h<float>(cur_meta->o_temp)[q];
    cur_meta->h<float>(cur_meta->o_topp)[r] = cur_meta->h<float>(cur_meta->o_topp)[q];
    cur_meta->h<unsigned long long>(cur_meta->o_seed)[r] = cur_meta->h<unsigned long long>(cur_meta->o_seed)[q];
    cur_meta->h<unsigned long long>(cur_meta->o_step)[r] = cur_meta->h<unsigned long long>(cur_meta->o_step)[q];
  }
  const int n_splits = pick_attn_splits(T, max_ctx_b);
  last_T = T;
  rowmap_identity = true;
  ENG_CUDA(cudaEventRecord(ev0, stream));
  ENG_CUDA(upload_step(T, T));
  if ((cfg.flags & DIS_FLAG_NO_GRAPH) || profiling) {
    ENG_CUDA(enqueue_layers(T, n_splits));
    ENG_CUDA(enqueue_head(T, all_greedy));
    prof_mark(P_COUNT);
  } else {
    const uint64_t key = ((uint64_t)T << 9) | ((uint64_t)n_splits << 1) | (all_greedy ? 1u : 0u);
    auto it = graphs.find(key);
    if (it == graphs.end()) {
      const uint64_t before = kernel_launches;
      cudaGraph_t g = nullptr;
      ENG_CUDA(cudaStreamBeginCapture(stream, cudaStreamCaptureModeThreadLocal));
      cudaError_t e1 = enqueue_layers(T, n_splits);
      cudaError_t e2 = e1 == cudaSuccess ? enqueue_head(T, all_greedy) : e1;
      cudaError_t e3 = cudaStreamEndCapture(stream, &g);
      if (e2 != cudaSuccess || e3 != cudaSuccess) {
        if (g) cudaGraphDestroy(g);
        cudaGetLastError();
        healthy = false;
        return fail(DIS_ERR_INFERENCE_FAILED, std::string("decode graph capture failed: ") + (g_last_error.empty() ? cudaGetErrorString(e2 != cudaSuccess ? e2 : e3) : g_last_error));
      }
      cudaGraphExec_t ex = nullptr;
      ENG_CUDA(cudaGraphInstantiate(&ex, g, 0));
      cudaGraphDestroy(g);
      graph_launches[key] = kernel_launches - before;
      kernel_launches = before;
      it = graphs.emplace(key, ex).first;
    }
    ENG_CUDA(cudaGraphLaunch(it->second, stream));
    last_graph = it->second;
    kernel_launches += graph_launches[key];
  }
  ENG_CUDA(cudaMemcpyAsync(h_out_tokens, d_out_tokens, (size_t)(max_batch + T) * 4, cudaMemcpyDeviceToHost, stream));
  ENG_CUDA(cudaEventRecord(ev1, stream));
  ENG_CUDA(cudaEventSynchronize(ev1));
  float ms = 0.f;
  ENG_CUDA(cudaEventElapsedTime(&ms, ev0, ev1));
  last_step_ms = ms;
  last_rows = T_live;
  for (int r = 0; r < T_live; r++) {
    const int i = live[(size_t)r];
    Sequence &s = seqs[(size_t)slots[i]];
    s.len += 1;  // the consumed token is now in the cache
    out_tokens[i] = h_out_tokens[r];
    s.last_logprob = reinterpret_cast<const float *>(h_out_tokens + max_batch)[r];
    finish_check(s, h_out_tokens[r]);
    out_finish[i] = s.finish;
  }
  return DIS_OK;
}

int Engine::replay_last_step(int n, float *avg_ms) {
  if (!last_graph || n <= 0 || !avg_ms) return fail(DIS_ERR_INVALID_ARGUMENT, "replay_last_step: no captured decode step yet");
  ENG_CUDA(cudaSetDevice(cfg.device));
  ENG_CUDA(cudaEventRecord(ev0, stream));
  for (int i = 0; i < n; i++) ENG_CUDA(cudaGraphLaunch(last_graph, stream));
  ENG_CUDA(cudaEventRecord(ev1, stream));
  ENG_CUDA(cudaEventSynchronize(ev1));
  float ms = 0.f;
  ENG_CUDA(cudaEventElapsedTime(&ms, ev0, ev1));
  *avg_ms = ms / (float)n;
  return DIS_OK;
}

void Engine::prof_mark(int cls) {
  if (!profiling) return;
  const size_t i = prof_marks.size();
  if (i >= prof_events.size()) {
    cudaEvent_t ev;
    if (cudaEventCreate(&ev) != cudaSuccess) return;
    prof_events.push_back(ev);
  }
  cudaEventRecord(prof_events[i], stream);
  prof_marks.emplace_back(cls, (int)i);
}

// One eager decode step with a CUDA event before every kernel class; class_ms[P_COUNT] receives the
// summed device time per class (event to next event), class_launches how often each class ran.
int Engine::profile_decode(int n, const int *slots, uint32_t *out_tokens, int *out_finish, float *class_ms,
                           int *class_launches) {
  if (!class_ms || !class_launches) return fail(DIS_ERR_INVALID_ARGUMENT, "profile_decode: null output");
  profiling = true;
  prof_marks.clear();
  const int rc = decode(n, slots, out_tokens, out_finish);
  profiling = false;
  if (rc != DIS_OK) return rc;
  for (int c = 0; c < P_COUNT; c++) { class_ms[c] = 0.f; class_launches[c] = 0; }
  for (size_t i = 0; i + 1 < prof_marks.size(); i++) {
    float ms = 0.f;
    ENG_CUDA(cudaEventElapsedTime(&ms, prof_events[(size_t)prof_marks[i].second], prof_events[(size_t)prof_marks[i + 1].second]));
    class_ms[prof_marks[i].first] += ms;
    class_launches[prof_marks[i].first] += 1;
  }
  return DIS_OK;
}

int Engine::last_logits_host(float *host, int rows) {
  if (!host || rows <= 0 || rows > last_rows) return fail(DIS_ERR_INVALID_ARGUMENT, "last_logits: no such rows");
  ENG_CUDA(cudaSetDevice(cfg.device));
  if (tp_size == 1) {
    ENG_CUDA(cudaMemcpyAsync(host, logits, (size_t)rows * m.vocab * 4, cudaMemcpyDeviceToHost, stream));
  } else {
    ENG_CUDA(gather_full_logits(last_rows));
    ENG_CUDA(cudaMemcpyAsync(host, full_logits, (size_t)rows * m.vocab * 4, cudaMemcpyDeviceToHost, stream));
  }
  ENG_CUDA(cudaStreamSynchronize(stream));
  return DIS_OK;
}

int Engine::last_logprobs(int n, const int *slots, float *out) {
  if (n <= 0 || !slots || !out) return fail(DIS_ERR_INVALID_ARGUMENT, "last_logprobs: bad argument");
  for (int i = 0; i < n; i++) {
    if (slots[i] < 0 || slots[i] >= max_batch || !seqs[(size_t)slots[i]].used) return fail(DIS_ERR_INVALID_ARGUMENT, "last_logprobs: no such sequence");
    out[i] = seqs[(size_t)slots[i]].last_logprob;
  }
  return DIS_OK;
}

void prefix_fill_stats(const PrefixStats &s, dis_prefix_stats *o);  // prefix.cpp

int Engine::prefix_stats(dis_prefix_stats *o) {
  if (!o) return fail(DIS_ERR_INVALID_ARGUMENT, "prefix_stats: null argument");
  prefix_fill_stats(pages_ix.stats(), o);
  return DIS_OK;
}

int Engine::status(dis_worker_status *o) {
  if (!o) return fail(DIS_ERR_INVALID_ARGUMENT, "status: null output");
  memset(o, 0, sizeof(*o));
  int active = 0;
  for (auto &s : seqs) if (s.used) active++;
  o->active_sequences = active;
  o->is_healthy = healthy && !shut;
  o->memory_used = bytes_allocated;
  size_t free_b = 0, total_b = 0;
  if (cudaSetDevice(cfg.device) == cudaSuccess && cudaMemGetInfo(&free_b, &total_b) == cudaSuccess) o->memory_available = free_b;
  o->kv_pages_total = n_pages;
  o->kv_pages_free = pages_ix.available();  // free + registered pages that no live sequence maps
  o->tokens_generated = tokens_generated;
  o->last_step_ms = last_step_ms;
  o->kernel_launches = kernel_launches;
  return DIS_OK;
}

}  // namespace dis
