// worker.cpp -- InferenceWorker over the C ABI, plus dis_worker_infer (the same loop for C hosts).
#include "worker.hpp"

#include <algorithm>
#include <cstring>

namespace dis_worker {

void InferenceWorker::raise(int code, const char *what) const {
  const std::string msg = std::string(what) + ": " + (eng_ ? dis_engine_last_error(eng_) : dis_last_error());
  switch (code) {
    case DIS_ERR_MODEL_NOT_LOADED: throw WorkerError(WorkerError::ModelNotLoaded, msg);
    case DIS_ERR_SHUTDOWN: throw WorkerError(WorkerError::Shutdown, msg);
    case DIS_ERR_OUT_OF_MEMORY: throw WorkerError(WorkerError::OutOfMemory, msg);
    default: throw WorkerError(WorkerError::InferenceFailed, msg);
  }
}

InferenceWorker::~InferenceWorker() {
  if (eng_) dis_engine_destroy(eng_);
}

void InferenceWorker::initialize() {
  if (down_) throw WorkerError(WorkerError::Shutdown, "worker was shut down");
  if (eng_) return;
  dis_engine_config c;
  memset(&c, 0, sizeof(c));
  c.model = cfg_.model;
  if (cfg_.context_size) c.model.max_ctx = (int32_t)cfg_.context_size;
  c.device = cfg_.device;
  c.tp_rank = cfg_.tp_rank;
  c.tp_size = cfg_.tp_size;
  c.max_batch = (int32_t)cfg_.max_batch;
  c.kv_pages = cfg_.kv_pages;
  c.flags = cfg_.flags;
  c.eos_token = cfg_.eos_token;
  if (cfg_.nccl_id) memcpy(c.nccl_id, cfg_.nccl_id, 128);
  int rc = dis_engine_create(&c, &eng_);
  if (rc != DIS_OK) raise(rc, "initialize");
  rc = dis_engine_load_synthetic(eng_, cfg_.synthetic_seed);
  if (rc != DIS_OK) raise(rc, "initialize/load");
  rc = dis_engine_finalize(eng_);
  if (rc != DIS_OK) raise(rc, "initialize/finalize");
}

void InferenceWorker::shutdown() {
  down_ = true;
  if (eng_) { dis_engine_destroy(eng_); eng_ = nullptr; }
}

WorkerStatus InferenceWorker::status() const {
  WorkerStatus s;
  if (!eng_) return s;
  dis_worker_status r;
  if (dis_engine_status(eng_, &r) != DIS_OK) return s;
  s.active_batches = (size_t)r.active_sequences;
  s.memory_used = (size_t)r.memory_used;
  s.memory_available = (size_t)r.memory_available;
  s.is_healthy = r.is_healthy != 0;
  return s;
}

dis_model_config InferenceWorker::model_info() const {
  dis_model_config m;
  memset(&m, 0, sizeof(m));
  if (eng_) dis_engine_model_info(eng_, &m);
  else m = cfg_.model;
  return m;
}

BatchResult InferenceWorker::infer(const InferenceBatch &batch, const TokenSink &sink) const {
  if (down_) throw WorkerError(WorkerError::Shutdown, "worker was shut down");
  if (!eng_) throw WorkerError(WorkerError::ModelNotLoaded, "initialize() has not been called");
  const size_t n = batch.input_ids.size();
  BatchResult out;
  out.batch_id = batch.id;
  out.results.resize(n);
  if (n == 0) return out;
  // flatten into the padded [n][stride] layout of the C entry point
  size_t stride = 1;
  for (auto &r : batch.input_ids) stride = std::max(stride, r.size());
  std::vector<uint32_t> ids(n * stride, 0);
  std::vector<int32_t> lens(n), out_lens(n), out_fin(n);
  std::vector<dis_sampling_params> params(n);
  std::vector<std::vector<uint32_t>> stop_flat(n);
  std::vector<std::vector<int32_t>> stop_lens(n);
  size_t max_new = 1;
  for (size_t i = 0; i < n; i++) {
    const auto &row = batch.input_ids[i];
    size_t len = i < batch.requests.size() && batch.requests[i].original_length ? batch.requests[i].original_length : row.size();
    if (i < batch.attention_mask.size() && !batch.requests.empty() && batch.requests[i].original_length == 0) {
      len = 0;
      for (uint8_t mk : batch.attention_mask[i]) len += mk ? 1 : 0;
    }
    len = std::min(len, row.size());
    std::copy(row.begin(), row.begin() + (long)len, ids.begin() + (long)(i * stride));
    lens[i] = (int32_t)len;
    InferenceParams ip = i < batch.requests.size() ? batch.requests[i].params : InferenceParams();
    if (batch.max_new_tokens) ip.max_tokens = std::min(ip.max_tokens, batch.max_new_tokens);
    for (auto &s : ip.stop_sequences) { stop_flat[i].insert(stop_flat[i].end(), s.begin(), s.end()); stop_lens[i].push_back((int32_t)s.size()); }
    dis_sampling_params &p = params[i];
    p.max_tokens = (int32_t)std::max<size_t>(ip.max_tokens, 1);
    p.temperature = ip.temperature;
    p.top_p = ip.top_p;
    p.seed = ip.seed;
    p.stop_tokens = stop_flat[i].data();
    p.stop_lens = stop_lens[i].data();
    p.n_stop = (int32_t)stop_lens[i].size();
    max_new = std::max(max_new, (size_t)p.max_tokens);
    if (i < batch.requests.size()) out.results[i].request_id = batch.requests[i].id;
  }
  std::vector<uint32_t> toks(n * max_new, 0);
  struct Ctx { const TokenSink *sink; } ctx{&sink};
  auto tramp = [](void *user, int32_t idx, uint32_t tok, int32_t index, int32_t fin, float logprob) -> int {
    const TokenSink &s = *static_cast<Ctx *>(user)->sink;
    return s((size_t)idx, tok, (size_t)index, fin != DIS_FINISH_NONE, logprob) ? 0 : 1;
  };
  dis_batch_result br;
  const int rc = dis_worker_infer(eng_, (int32_t)n, ids.data(), (int32_t)stride, lens.data(), params.data(), toks.data(),
                                  (int32_t)max_new, out_lens.data(), out_fin.data(), &br, sink ? +tramp : nullptr, &ctx);
  if (rc != DIS_OK) raise(rc, "infer");
  for (size_t i = 0; i < n; i++) {
    RequestResult &r = out.results[i];
    r.tokens.assign(toks.begin() + (long)(i * max_new), toks.begin() + (long)(i * max_new + (size_t)out_lens[i]));
    r.prompt_tokens = (size_t)lens[i];
    r.completion_tokens = r.tokens.size();
    r.failed = out_fin[i] == DIS_FINISH_FAILED;
    r.finish_reason = out_fin[i] == DIS_FINISH_STOP ? FinishReason::Stop
                      : out_fin[i] == DIS_FINISH_STOP_SEQUENCE ? FinishReason::StopSequence : FinishReason::Length;
  }
  out.inference_time = std::chrono::duration<double, std::milli>(br.inference_ms);
  out.time_to_first_token = std::chrono::duration<double, std::milli>(br.ttft_ms);
  out.tokens_generated = (size_t)br.tokens_generated;
  return out;
}

}  // namespace dis_worker

// ---- the same generation loop for C hosts ---------------------------------------------------------
extern "C" int dis_worker_infer(dis_engine *e, int32_t n, const uint32_t *input_ids, int32_t stride, const int32_t *lens,
                                const dis_sampling_params *params, uint32_t *out_tokens, int32_t max_new_cap,
                                int32_t *out_lens, int32_t *out_finish, dis_batch_result *result,
                                dis_token_callback on_token, void *user) {
  if (!e || n <= 0 || !input_ids || !lens || !params || !out_tokens || !out_lens || !out_finish || max_new_cap <= 0)
    return DIS_ERR_INVALID_ARGUMENT;
  const auto t0 = std::chrono::steady_clock::now();
  std::vector<int32_t> slot((size_t)n, -1);
  uint64_t generated = 0;
  double ttft = 0.0;
  int rc = DIS_OK;
  // Admit: a request that cannot be admitted fails alone (Property 22, design.md:812-816).
  std::vector<int32_t> live_idx, live_slot;
  for (int32_t i = 0; i < n; i++) {
    out_lens[i] = 0;
    out_finish[i] = DIS_FINISH_NONE;
    dis_sampling_params p = params[i];
    p.max_tokens = std::min(p.max_tokens, max_new_cap);
    const int r = dis_engine_seq_add(e, input_ids + (size_t)i * stride, lens[i], &p, &slot[(size_t)i]);
    if (r == DIS_ERR_MODEL_NOT_LOADED || r == DIS_ERR_SHUTDOWN) { rc = r; break; }
    if (r != DIS_OK) { out_finish[i] = DIS_FINISH_FAILED; slot[(size_t)i] = -1; continue; }
    live_idx.push_back(i);
    live_slot.push_back(slot[(size_t)i]);
  }
  auto cleanup = [&]() { for (int32_t s : slot) if (s >= 0) dis_engine_seq_remove(e, s); };
  if (rc != DIS_OK) { cleanup(); return rc; }
  std::vector<uint32_t> toks(live_idx.size());
  std::vector<int32_t> fin(live_idx.size());
  std::vector<float> lps(live_idx.size());
  auto deliver = [&]() {
    if (on_token && !live_slot.empty()) dis_engine_last_logprobs(e, (int32_t)live_slot.size(), live_slot.data(), lps.data());
    std::vector<int32_t> ni, ns;
    for (size_t k = 0; k < live_idx.size(); k++) {
      const int32_t i = live_idx[k];
      bool keep = true;
      if (fin[k] == DIS_FINISH_FAILED) { out_finish[i] = DIS_FINISH_FAILED; keep = false; }  // produced no token
      else {
        out_tokens[(size_t)i * max_new_cap + out_lens[i]] = toks[k];
        out_lens[i]++;
        generated++;
        out_finish[i] = fin[k];
        if (on_token && on_token(user, i, toks[k], out_lens[i] - 1, fin[k], lps[k]) != 0 && fin[k] == DIS_FINISH_NONE) {
          out_finish[i] = DIS_FINISH_STOP;  // dropped by the host
          keep = false;
        }
        if (fin[k] != DIS_FINISH_NONE) keep = false;
      }
      if (keep) { ni.push_back(i); ns.push_back(live_slot[k]); }
    }
    live_idx.swap(ni);
    live_slot.swap(ns);
  };
  if (!live_idx.empty()) {
    rc = dis_engine_prefill(e, (int32_t)live_slot.size(), live_slot.data(), toks.data(), fin.data());
    if (rc != DIS_OK) { cleanup(); return rc; }
    ttft = std::chrono::duration<double, std::milli>(std::chrono::steady_clock::now() - t0).count();
    deliver();
  }
  while (!live_idx.empty()) {
    rc = dis_engine_decode(e, (int32_t)live_slot.size(), live_slot.data(), toks.data(), fin.data());
    if (rc != DIS_OK) { cleanup(); return rc; }
    deliver();
  }
  cleanup();
  if (result) {
    result->inference_ms = std::chrono::duration<double, std::milli>(std::chrono::steady_clock::now() - t0).count();
    result->ttft_ms = ttft;
    result->tokens_generated = generated;
  }
  return DIS_OK;
}
