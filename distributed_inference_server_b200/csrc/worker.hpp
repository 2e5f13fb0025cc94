// worker.hpp -- C++ mirror of the reference's spec'd worker interface, written above the C ABI only
// (it includes nothing but include/dis_engine.h), the way a Rust `impl InferenceWorker` in
// crates/inference would be.  Names and argument meaning follow
// .kiro/specs/distributed-inference-server/design.md:310-361:
//   WorkerConfig (316-322)  InferenceWorker{initialize, infer, shutdown, status, model_info} (334-342)
//   InferenceBatch (241-248)  InferenceParams (666-671)  BatchResult / RequestResult (345-360)
// and crates/core: FinishReason (models.rs:26-32), WorkerError (error.rs:115-128),
// RequestID = Uuid (types.rs:7).
#pragma once
#include <array>
#include <chrono>
#include <cstdint>
#include <functional>
#include <stdexcept>
#include <string>
#include <vector>

#include "../../include/dis_engine.h"

namespace dis_worker {

enum class FinishReason { Stop, Length, StopSequence };  // models.rs:26-32

// error.rs:115-128
struct WorkerError : std::runtime_error {
  enum Kind { ModelNotLoaded, InferenceFailed, Shutdown, OutOfMemory } kind;
  WorkerError(Kind k, const std::string &m) : std::runtime_error(m), kind(k) {}
};

using RequestId = std::array<uint8_t, 16>;  // uuid::Uuid

struct InferenceParams {  // design.md:666-671
  size_t max_tokens = 256;   // models.rs:294 default
  float temperature = 1.0f;  // models.rs:298
  float top_p = 1.0f;        // models.rs:302
  std::vector<std::vector<uint32_t>> stop_sequences;
  uint64_t seed = 0;  // not in the reference: makes sampling reproducible
};

struct BatchedRequest {  // design.md:258-266 (channels replaced by the per-token callback of infer)
  RequestId id{};
  size_t original_length = 0;
  size_t padded_length = 0;
  InferenceParams params;
};

struct InferenceBatch {  // design.md:241-248
  RequestId id{};
  std::vector<BatchedRequest> requests;
  std::vector<std::vector<uint32_t>> input_ids;      // right-padded to a common length
  std::vector<std::vector<uint8_t>> attention_mask;  // 1 = real token
  size_t max_new_tokens = 0;
};

struct RequestResult {  // design.md:353-360 (`text` needs a tokenizer the reference does not have)
  RequestId request_id{};
  std::vector<uint32_t> tokens;
  FinishReason finish_reason = FinishReason::Length;
  bool failed = false;  // Property 22: a failed request does not fail its batch
  size_t prompt_tokens = 0;
  size_t completion_tokens = 0;
};

struct BatchResult {  // design.md:345-350
  RequestId batch_id{};
  std::vector<RequestResult> results;
  std::chrono::duration<double, std::milli> inference_time{0};
  std::chrono::duration<double, std::milli> time_to_first_token{0};
  size_t tokens_generated = 0;
};

struct WorkerStatus {  // design.md:290-297
  size_t active_batches = 0;
  size_t memory_used = 0;
  size_t memory_available = 0;
  bool is_healthy = false;
};

struct WorkerConfig {  // design.md:316-322 + the fields a GPU engine needs
  dis_model_config model{};
  size_t context_size = 0;  // 0 = model.max_ctx
  int device = 0;
  int tp_rank = 0, tp_size = 1;
  size_t max_batch = 32;  // requirements.md:46
  int64_t kv_pages = 0;
  uint32_t flags = 0;
  int32_t eos_token = -1;
  uint64_t synthetic_seed = 0;  // random-init weights; real weights come through load_tensor()
  const uint8_t *nccl_id = nullptr;
};

// Called once per generated token (TokenEvent::Token{token, index, logprob}, models.rs:272-277).  Return false
// to drop the request (client disconnected, requirements.md:85).
using TokenSink = std::function<bool(size_t request_index, uint32_t token, size_t index, bool finished, float logprob)>;

class InferenceWorker {
 public:
  explicit InferenceWorker(const WorkerConfig &cfg) : cfg_(cfg) {}
  ~InferenceWorker();
  InferenceWorker(const InferenceWorker &) = delete;
  InferenceWorker &operator=(const InferenceWorker &) = delete;

  void initialize();                                         // design.md:337
  BatchResult infer(const InferenceBatch &batch, const TokenSink &sink = nullptr) const;  // design.md:338
  void shutdown();                                           // design.md:339
  WorkerStatus status() const;                               // design.md:340
  dis_model_config model_info() const;                       // design.md:341
  dis_engine *raw() const { return eng_; }                   // for load_tensor before initialize() finalises

 private:
  [[noreturn]] void raise(int code, const char *what) const;
  WorkerConfig cfg_;
  dis_engine *eng_ = nullptr;
  bool down_ = false;
};

}  // namespace dis_worker
