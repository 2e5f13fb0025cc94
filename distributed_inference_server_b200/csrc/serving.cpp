// serving.cpp -- the callers either side of the hot path, host side, above the C ABI:
//
//   RequestQueue       three FIFO lanes drained High -> Normal -> Low, admission refused while the queue is over
//                      capacity or in back-pressure, hysteresis between a high and a low watermark, expiry of
//                      requests that waited too long.  Behaviour restated from the reference's
//                      PriorityQueueManager (crates/core/src/queue.rs:103-126 enqueue, :130-158 dequeue_batch,
//                      :198-226 remove_expired, :235-249 watermarks; defaults :24-33) -- a new implementation,
//                      with an explicit clock so that it can be tested without sleeping.
//   ContinuousBatcher  the scheduler the reference only has as a window batcher in its spec (design.md:227-267,
//                      requirements.md:45-48): every step admits queued requests into free engine slots (KV pages
//                      permitting) and runs ONE forward pass in which every running sequence decodes a token and
//                      the admitted prompts advance by a bounded chunk (dis_engine_step; shortest prompt first inside a
//                      priority class, with aging), retires finished
//                      sequences and reports one event per token (TokenEvent, models.rs:268-288).
#include <algorithm>
#include <chrono>
#include <cstring>
#include <deque>
#include <mutex>
#include <string>
#include <unordered_map>
#include <vector>

#include "../../include/dis_engine.h"

namespace {

// RequestID = Uuid (crates/core/src/types.rs:7): 16 opaque bytes, compared bytewise
struct ReqId {
  uint8_t b[16];
  bool operator==(const ReqId &o) const { return memcmp(b, o.b, 16) == 0; }
};
static ReqId make_id(const uint8_t *p) { ReqId r; memcpy(r.b, p, 16); return r; }

struct Queued {
  ReqId id;
  int priority;
  double enqueued_at;
  std::vector<uint32_t> prompt;
  dis_sampling_params params;
  std::vector<uint32_t> stop_tokens;
  std::vector<int32_t> stop_lens;
};

class RequestQueue {
 public:
  explicit RequestQueue(const dis_queue_config &c) : cfg_(c) {}

  // queue.rs:103-126: refused while back-pressure is on or the queue is at capacity
  bool enqueue(Queued &&r) {
    if (backpressure_ || total() >= (size_t)cfg_.max_queue_size) return false;
    lane(r.priority).push_back(std::move(r));
    update();
    return true;
  }
  // queue.rs:130-158: High first, then Normal, then Low, FIFO inside a lane
  bool pop(Queued *out) {
    for (int p = DIS_PRIORITY_HIGH; p >= DIS_PRIORITY_LOW; p--) {
      auto &q = lanes_[p];
      if (!q.empty()) { *out = std::move(q.front()); q.pop_front(); update(); return true; }
    }
    return false;
  }
  // a request that could not be admitted yet (no KV pages) goes back to the head of its lane
  void push_front(Queued &&r) { lane(r.priority).push_front(std::move(r)); update(); }
  // queue.rs:198-226: everything that waited longer than the timeout leaves the queue
  void remove_expired(double now, std::vector<Queued> *expired) {
    for (auto &q : lanes_) {
      for (size_t i = 0; i < q.size();) {
        if (now - q[i].enqueued_at > cfg_.request_timeout_s) { expired->push_back(std::move(q[i])); q.erase(q.begin() + (long)i); }
        else i++;
      }
    }
    update();
  }
  // a queued request whose client went away (requirements.md:85)
  bool remove(const ReqId &id, Queued *out) {
    for (auto &q : lanes_)
      for (size_t i = 0; i < q.size(); i++)
        if (q[i].id == id) { *out = std::move(q[i]); q.erase(q.begin() + (long)i); update(); return true; }
    return false;
  }
  size_t total() const { return lanes_[0].size() + lanes_[1].size() + lanes_[2].size(); }
  size_t depth(int p) const { return lanes_[p].size(); }
  bool accepting() const { return !backpressure_; }

 private:
  std::deque<Queued> &lane(int p) { return lanes_[p < 0 ? 0 : (p > 2 ? 2 : p)]; }
  // queue.rs:235-249: on above the high watermark, off only once below the low watermark
  void update() {
    const size_t t = total();
    if (backpressure_) { if (t < (size_t)cfg_.low_watermark) backpressure_ = false; }
    else if (t > (size_t)cfg_.high_watermark) backpressure_ = true;
  }
  dis_queue_config cfg_;
  std::deque<Queued> lanes_[3];
  bool backpressure_ = false;
};

double steady_now() {
  return std::chrono::duration<double>(std::chrono::steady_clock::now().time_since_epoch()).count();
}

struct Active {
  ReqId id;
  int32_t slot;
  int32_t produced;
  int32_t prompt_tokens;
  double enqueued_at;
  int32_t priority;
};

// Which admitted prompt is read first inside a pass.  The admission QUEUE keeps the reference's order (priority lane,
// then arrival: crates/core/src/queue.rs:130-158); this is the order among requests that already hold a slot, which the
// reference does not define.  Shortest prompt first inside a priority class minimises the mean time to first token (a
// 128-token prompt does not wait behind an 8192-token one that arrived a moment earlier), and `aging` tokens of
// precedence per second of waiting bound how long a long prompt can be overtaken.  Ties keep arrival order.
void order_prompts(int n, const int32_t *priority, const int32_t *prompt_tokens, const double *waited_s, double aging,
                   int32_t *order) {
  std::vector<int32_t> idx((size_t)std::max(n, 0));
  for (int i = 0; i < n; i++) idx[(size_t)i] = i;
  std::stable_sort(idx.begin(), idx.end(), [&](int32_t a, int32_t b) {
    if (priority[a] != priority[b]) return priority[a] > priority[b];  // High = 2 first
    const double ka = (double)prompt_tokens[a] - aging * waited_s[a], kb = (double)prompt_tokens[b] - aging * waited_s[b];
    return ka < kb;
  });
  for (int i = 0; i < n; i++) order[i] = idx[(size_t)i];
}

}  // namespace

struct dis_serving {
  dis_engine *eng = nullptr;
  dis_queue_config cfg{};
  RequestQueue queue;
  std::vector<Active> active;
  std::deque<dis_serving_event> pending;
  std::mutex mu;
  int32_t max_active = 0;
  int32_t prefill_budget = 4096;  // prompt tokens read per scheduler step (beside one token for every running sequence)
  int32_t prefill_order = 1;      // 0: admission order, 1: shortest prompt first inside a priority class, with aging
  double aging_tokens_per_s = 4096.0;
  std::string last_error;
  explicit dis_serving(const dis_queue_config &c) : cfg(c), queue(c) {}
};

static void emit(dis_serving *s, const ReqId &id, int kind, uint32_t token, int32_t index, int32_t finish, int32_t prompt_tokens,
                 double waited_ms, float logprob = 0.f) {
  dis_serving_event e;
  memset(&e, 0, sizeof(e));
  memcpy(e.request_id, id.b, 16);
  e.kind = kind; e.token = token; e.index = index; e.finish_reason = finish;
  e.prompt_tokens = prompt_tokens; e.waited_ms = waited_ms; e.logprob = logprob;
  s->pending.push_back(e);
}

extern "C" {

int dis_serving_create(dis_engine *engine, const dis_queue_config *cfg, int32_t max_active, dis_serving **out) {
  if (!out) return DIS_ERR_INVALID_ARGUMENT;
  dis_queue_config c;
  if (cfg) c = *cfg;
  else { c.high_watermark = 1000; c.low_watermark = 500; c.max_queue_size = 2000; c.request_timeout_s = 30.0; }  // queue.rs:24-33
  if (c.low_watermark > c.high_watermark || c.max_queue_size <= 0) return DIS_ERR_INVALID_ARGUMENT;
  dis_serving *s = new dis_serving(c);
  s->eng = engine;  // NULL: queue-only (admission control can be exercised without a GPU)
  s->max_active = max_active > 0 ? max_active : 32;
  *out = s;
  return DIS_OK;
}

int dis_serving_destroy(dis_serving *s) {
  if (!s) return DIS_ERR_INVALID_ARGUMENT;
  if (s->eng) for (auto &a : s->active) dis_engine_seq_remove(s->eng, a.slot);
  delete s;
  return DIS_OK;
}

const char *dis_serving_last_error(dis_serving *s) { return s ? s->last_error.c_str() : ""; }

int dis_serving_submit_at(dis_serving *s, const uint8_t request_id[16], const uint32_t *prompt, int32_t n_prompt,
                          const dis_sampling_params *params, int32_t priority, double now_s) {
  if (!s || !request_id || !prompt || n_prompt <= 0 || !params) return DIS_ERR_INVALID_ARGUMENT;
  std::lock_guard<std::mutex> lk(s->mu);
  Queued q;
  q.id = make_id(request_id); q.priority = priority; q.enqueued_at = now_s;
  q.prompt.assign(prompt, prompt + n_prompt);
  q.params = *params;
  size_t off = 0;
  for (int i = 0; i < params->n_stop; i++) {
    const int len = params->stop_lens ? params->stop_lens[i] : 0;
    for (int k = 0; k < len; k++) q.stop_tokens.push_back(params->stop_tokens[off + k]);
    q.stop_lens.push_back(len);
    off += (size_t)std::max(len, 0);
  }
  if (!s->queue.enqueue(std::move(q))) { s->last_error = "queue full"; return DIS_ERR_QUEUE_FULL; }
  return DIS_OK;
}

int dis_serving_submit(dis_serving *s, const uint8_t request_id[16], const uint32_t *prompt, int32_t n_prompt,
                       const dis_sampling_params *params, int32_t priority) {
  return dis_serving_submit_at(s, request_id, prompt, n_prompt, params, priority, steady_now());
}

int dis_serving_set_prefill_budget(dis_serving *s, int32_t tokens) {
  if (!s || tokens < 1) return DIS_ERR_INVALID_ARGUMENT;
  std::lock_guard<std::mutex> lk(s->mu);
  s->prefill_budget = tokens;
  return DIS_OK;
}

int dis_serving_set_prefill_order(dis_serving *s, int32_t mode, double aging_tokens_per_s) {
  if (!s || (mode != 0 && mode != 1) || !(aging_tokens_per_s >= 0.0)) return DIS_ERR_INVALID_ARGUMENT;
  std::lock_guard<std::mutex> lk(s->mu);
  s->prefill_order = mode;
  s->aging_tokens_per_s = aging_tokens_per_s;
  return DIS_OK;
}

int dis_serving_order_prompts(int32_t n, const int32_t *priority, const int32_t *prompt_tokens, const double *waited_s,
                              double aging_tokens_per_s, int32_t *order) {
  if (n < 0 || (n > 0 && (!priority || !prompt_tokens || !waited_s || !order)) || !(aging_tokens_per_s >= 0.0)) return DIS_ERR_INVALID_ARGUMENT;
  order_prompts(n, priority, prompt_tokens, waited_s, aging_tokens_per_s, order);
  return DIS_OK;
}

int dis_serving_depth(dis_serving *s, dis_queue_depth *out) {
  if (!s || !out) return DIS_ERR_INVALID_ARGUMENT;
  std::lock_guard<std::mutex> lk(s->mu);
  out->low = (int32_t)s->queue.depth(DIS_PRIORITY_LOW);
  out->normal = (int32_t)s->queue.depth(DIS_PRIORITY_NORMAL);
  out->high = (int32_t)s->queue.depth(DIS_PRIORITY_HIGH);
  out->total = (int32_t)s->queue.total();
  out->accepting = s->queue.accepting() ? 1 : 0;
  out->active = (int32_t)s->active.size();
  return DIS_OK;
}

// Queue-only helper (tests of the admission order without an engine): pops up to `max` requests in
// scheduling order and returns their ids.
int dis_serving_debug_pop(dis_serving *s, int32_t max, uint8_t *ids /* [max][16] */, int32_t *n) {
  if (!s || !ids || !n) return DIS_ERR_INVALID_ARGUMENT;
  std::lock_guard<std::mutex> lk(s->mu);
  *n = 0;
  Queued q;
  while (*n < max && s->queue.pop(&q)) { memcpy(ids + (size_t)*n * 16, q.id.b, 16); (*n)++; }
  return DIS_OK;
}

// The client of a request went away (requirements.md:85): a queued request leaves the queue, a running one gives
// its slot and KV pages back at once.  Either way one DONE event with FinishReason::Stop closes the stream.
int dis_serving_cancel(dis_serving *s, const uint8_t request_id[16]) {
  if (!s || !request_id) return DIS_ERR_INVALID_ARGUMENT;
  std::lock_guard<std::mutex> lk(s->mu);
  const ReqId id = make_id(request_id);
  Queued q;
  if (s->queue.remove(id, &q)) {
    emit(s, id, DIS_EVENT_DONE, 0, 0, DIS_FINISH_STOP, (int32_t)q.prompt.size(), 0.0);
    return DIS_OK;
  }
  for (size_t i = 0; i < s->active.size(); i++)
    if (s->active[i].id == id) {
      const Active a = s->active[i];
      if (s->eng) dis_engine_seq_remove(s->eng, a.slot);
      s->active.erase(s->active.begin() + (long)i);
      emit(s, id, DIS_EVENT_DONE, 0, a.produced, DIS_FINISH_STOP, a.prompt_tokens, 0.0);
      return DIS_OK;
    }
  s->last_error = "cancel: no such request";
  return DIS_ERR_INVALID_ARGUMENT;
}

static int drain(dis_serving *s, dis_serving_event *events, int32_t cap, int32_t *n_events) {
  *n_events = 0;
  while (*n_events < cap && !s->pending.empty()) { events[(*n_events)++] = s->pending.front(); s->pending.pop_front(); }
  return DIS_OK;
}

static int step_impl(dis_serving *s, double now_s, bool clock_injected, dis_serving_event *events, int32_t cap, int32_t *n_events) {
  if (!s || !events || !n_events || cap <= 0) return DIS_ERR_INVALID_ARGUMENT;
  std::lock_guard<std::mutex> lk(s->mu);
  if (!s->pending.empty()) return drain(s, events, cap, n_events);  // deliver what is owed first
  // 1. requests that waited too long (ApiError::Timeout -> HTTP 408, crates/core/src/error.rs:46)
  std::vector<Queued> expired;
  s->queue.remove_expired(now_s, &expired);
  for (auto &q : expired) emit(s, q.id, DIS_EVENT_EXPIRED, 0, 0, DIS_FINISH_FAILED, (int32_t)q.prompt.size(), (now_s - q.enqueued_at) * 1e3);
  if (!s->eng) return drain(s, events, cap, n_events);

  // 2. admission: free slots, High -> Normal -> Low; a request is admitted when the engine can reserve its KV pages
  Queued q;
  while ((int32_t)s->active.size() < s->max_active && s->queue.pop(&q)) {
    dis_sampling_params p = q.params;
    p.stop_tokens = q.stop_tokens.empty() ? nullptr : q.stop_tokens.data();
    p.stop_lens = q.stop_lens.empty() ? nullptr : q.stop_lens.data();
    p.n_stop = (int32_t)q.stop_lens.size();
    int32_t slot = -1;
    const int rc = dis_engine_seq_add(s->eng, q.prompt.data(), (int32_t)q.prompt.size(), &p, &slot);
    if (rc == DIS_ERR_OUT_OF_MEMORY && !s->active.empty()) {
      s->queue.push_front(std::move(q));  // wait for running sequences to free KV pages / slots
      break;
    }
    if (rc == DIS_ERR_MODEL_NOT_LOADED || rc == DIS_ERR_SHUTDOWN) { s->queue.push_front(std::move(q)); s->last_error = dis_engine_last_error(s->eng); return rc; }
    if (rc != DIS_OK) {  // a bad request fails alone (Property 22, design.md:812-816)
      emit(s, q.id, DIS_EVENT_DONE, 0, 0, DIS_FINISH_FAILED, (int32_t)q.prompt.size(), (now_s - q.enqueued_at) * 1e3);
      continue;
    }
    s->active.push_back(Active{q.id, slot, 0, (int32_t)q.prompt.size(), q.enqueued_at, q.priority});
  }
  if (s->active.empty()) return drain(s, events, cap, n_events);
  // 3. ONE forward pass for everybody: each running sequence decodes a token and, in the same pass, the prompts still
  // being read advance by up to prefill_budget tokens in admission order (chunked prefill).  Running sequences are
  // never stalled behind a long prompt, and a prompt is never starved by the running ones.
  if (s->prefill_order == 1) {
    // running sequences first (a decode row each, order immaterial), then the prompts still being read in the order
    // order_prompts gives them; dis_engine_step hands out the pass's prompt budget in list order
    std::vector<Active> running, reading;
    for (auto &a : s->active) (a.produced > 0 ? running : reading).push_back(a);
    if (reading.size() > 1) {
      std::vector<int32_t> pr, pt, ord(reading.size());
      std::vector<double> wt;
      for (auto &a : reading) { pr.push_back(a.priority); pt.push_back(a.prompt_tokens); wt.push_back(std::max(0.0, now_s - a.enqueued_at)); }
      order_prompts((int)reading.size(), pr.data(), pt.data(), wt.data(), s->aging_tokens_per_s, ord.data());
      for (int32_t i : ord) running.push_back(reading[(size_t)i]);
    } else {
      for (auto &a : reading) running.push_back(a);
    }
    s->active.swap(running);
  }
  std::vector<int32_t> slots, fin(s->active.size()), produced(s->active.size());
  std::vector<uint32_t> toks(s->active.size());
  for (auto &a : s->active) slots.push_back(a.slot);
  const int rc = dis_engine_step(s->eng, (int32_t)slots.size(), slots.data(), s->prefill_budget, toks.data(), fin.data(), produced.data());
  if (rc != DIS_OK) {
    // the pass failed as a whole: sequences that have not produced anything yet get their DONE/FAILED event, so no
    // request id vanishes; the others stay (the caller sees the error and decides)
    s->last_error = dis_engine_last_error(s->eng);
    std::vector<Active> keep;
    for (auto &a : s->active) {
      if (a.produced > 0) { keep.push_back(a); continue; }
      dis_engine_seq_remove(s->eng, a.slot);
      emit(s, a.id, DIS_EVENT_DONE, 0, 0, DIS_FINISH_FAILED, a.prompt_tokens, (now_s - a.enqueued_at) * 1e3);
    }
    s->active.swap(keep);
    return rc;
  }
  std::vector<Active> keep;
  for (size_t i = 0; i < s->active.size(); i++) {
    Active a = s->active[i];
    if (fin[i] == DIS_FINISH_FAILED && !produced[i]) {  // produced no token in this pass
      emit(s, a.id, DIS_EVENT_DONE, 0, a.produced, DIS_FINISH_FAILED, a.prompt_tokens, 0.0);
      dis_engine_seq_remove(s->eng, a.slot);
      continue;
    }
    if (!produced[i]) { keep.push_back(a); continue; }  // its prompt is still being read
    float lp = 0.f;
    dis_engine_last_logprobs(s->eng, 1, &a.slot, &lp);
    // under the injected clock of the *_at variants the step's own duration is not visible: TTFT = time queued
    const double t_first = clock_injected ? now_s : steady_now();
    emit(s, a.id, DIS_EVENT_TOKEN, toks[i], a.produced, fin[i], a.prompt_tokens, a.produced == 0 ? (t_first - a.enqueued_at) * 1e3 : 0.0, lp);
    a.produced++;
    if (fin[i] != DIS_FINISH_NONE) {
      emit(s, a.id, DIS_EVENT_DONE, 0, a.produced, fin[i], a.prompt_tokens, 0.0);
      dis_engine_seq_remove(s->eng, a.slot);
    } else keep.push_back(a);
  }
  s->active.swap(keep);
  return drain(s, events, cap, n_events);
}

int dis_serving_step_at(dis_serving *s, double now_s, dis_serving_event *events, int32_t cap, int32_t *n_events) {
  return step_impl(s, now_s, true, events, cap, n_events);
}
int dis_serving_step(dis_serving *s, dis_serving_event *events, int32_t cap, int32_t *n_events) {
  return step_impl(s, steady_now(), false, events, cap, n_events);
}

}  // extern "C"
