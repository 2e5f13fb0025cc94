//! inference-sys -- hand-written bindings for include/dis_engine.h (bindgen would produce the same).
//! UNVERIFIED: written without a Rust toolchain; see ../../README.md.
#![allow(non_camel_case_types)]
use std::os::raw::{c_char, c_int, c_void};

pub const DIS_OK: c_int = 0;
pub const DIS_FLAG_PREFIX_CACHE: u32 = 8;

#[repr(C)]
#[derive(Clone, Copy, Default)]
pub struct dis_model_config {
    pub vocab: i32, pub hidden: i32, pub layers: i32, pub n_heads: i32, pub n_kv_heads: i32,
    pub head_dim: i32, pub inter: i32, pub max_ctx: i32, pub rope_theta: f32, pub rms_eps: f32,
}

#[repr(C)]
pub struct dis_engine_config {
    pub model: dis_model_config, pub device: i32, pub tp_rank: i32, pub tp_size: i32, pub max_batch: i32,
    pub max_step_tokens: i32, pub page_size: i32, pub kv_pages: i64, pub kv_mem_fraction: f32, pub flags: u32,
    pub eos_token: i32, pub stream: *mut c_void, pub nccl_id: [u8; 128],
}

#[repr(C)]
pub struct dis_sampling_params {
    pub max_tokens: i32, pub temperature: f32, pub top_p: f32, pub seed: u64,
    pub stop_tokens: *const u32, pub stop_lens: *const i32, pub n_stop: i32,
}

#[repr(C)]
#[derive(Default)]
pub struct dis_worker_status {
    pub active_sequences: i32, pub is_healthy: i32, pub memory_used: u64, pub memory_available: u64,
    pub kv_pages_total: i64, pub kv_pages_free: i64, pub tokens_generated: u64, pub last_step_ms: f64,
    pub kernel_launches: u64,
}

#[repr(C)]
#[derive(Default)]
pub struct dis_prefix_stats {
    pub cached_pages: i64, pub evictable_pages: i64, pub free_pages: i64, pub lookups: u64, pub hits: u64,
    pub lookup_tokens: u64, pub hit_tokens: u64, pub evictions: u64,
}

pub enum dis_engine {}
pub enum dis_serving {}

/// Admission queue of the continuous batcher: the watermarks and timeout of `PriorityQueueManager`
/// (crates/core/src/queue.rs:103-158).
#[repr(C)]
#[derive(Clone, Copy)]
pub struct dis_queue_config {
    pub high_watermark: i32, pub low_watermark: i32, pub max_queue_size: i32, pub request_timeout_s: f64,
}

#[repr(C)]
#[derive(Clone, Copy, Default)]
pub struct dis_queue_depth { pub high: i32, pub normal: i32, pub low: i32, pub total: i32, pub accepting: i32, pub active: i32 }

pub const DIS_EVENT_TOKEN: i32 = 0;
pub const DIS_EVENT_DONE: i32 = 1;
pub const DIS_EVENT_EXPIRED: i32 = 2;

/// One streamed item: `TokenEvent::Token { index, logprob }` / `TokenEvent::Done` (crates/core/src/models.rs:268-288)
/// for the request whose `Uuid` bytes (`RequestID`, crates/core/src/types.rs:7) are in `request_id`.
#[repr(C)]
#[derive(Clone, Copy)]
pub struct dis_serving_event {
    pub request_id: [u8; 16], pub kind: i32, pub token: u32, pub index: i32, pub finish_reason: i32,
    pub prompt_tokens: i32, pub waited_ms: f64, pub logprob: f32,
}

extern "C" {
    pub fn dis_abi_version() -> c_int;
    pub fn dis_last_error() -> *const c_char;
    pub fn dis_engine_create(cfg: *const dis_engine_config, out: *mut *mut dis_engine) -> c_int;
    pub fn dis_engine_destroy(e: *mut dis_engine) -> c_int;
    pub fn dis_engine_last_error(e: *mut dis_engine) -> *const c_char;
    pub fn dis_engine_load_tensor(e: *mut dis_engine, kind: c_int, layer: c_int, host: *const c_void, dtype: c_int) -> c_int;
    pub fn dis_engine_load_synthetic(e: *mut dis_engine, seed: u64) -> c_int;
    pub fn dis_engine_finalize(e: *mut dis_engine) -> c_int;
    pub fn dis_engine_seq_add(e: *mut dis_engine, prompt: *const u32, n: i32, p: *const dis_sampling_params, slot: *mut i32) -> c_int;
    pub fn dis_engine_seq_remove(e: *mut dis_engine, slot: i32) -> c_int;
    pub fn dis_engine_prefill(e: *mut dis_engine, n: i32, slots: *const i32, out_tokens: *mut u32, out_finish: *mut i32) -> c_int;
    pub fn dis_engine_decode(e: *mut dis_engine, n: i32, slots: *const i32, out_tokens: *mut u32, out_finish: *mut i32) -> c_int;
    /// One forward pass for running sequences (a decode row each) and unfinished prompts (a bounded chunk, in list order).
    pub fn dis_engine_step(e: *mut dis_engine, n: i32, slots: *const i32, max_prefill_tokens: i32, out_tokens: *mut u32,
                           out_finish: *mut i32, out_produced: *mut i32) -> c_int;
    /// TokenEvent::Token.logprob (crates/core/src/models.rs:272-277) of each listed sequence's newest token.
    pub fn dis_engine_last_logprobs(e: *mut dis_engine, n: i32, slots: *const i32, out_logprobs: *mut f32) -> c_int;
    pub fn dis_engine_status(e: *mut dis_engine, out: *mut dis_worker_status) -> c_int;
    pub fn dis_engine_prefix_stats(e: *mut dis_engine, out: *mut dis_prefix_stats) -> c_int;
    pub fn dis_engine_model_info(e: *mut dis_engine, out: *mut dis_model_config) -> c_int;
    pub fn dis_tp_unique_id(out: *mut u8) -> c_int;
    pub fn dis_engine_tp_export(e: *mut dis_engine, out: *mut u8) -> c_int;
    pub fn dis_engine_tp_import(e: *mut dis_engine, handles: *const u8) -> c_int;

    // ---- continuous batching above the engine (optional: a host that keeps its own PriorityQueueManager drives
    // dis_engine_step itself; one that does not gets admission by priority, chunked prefill and streaming from these)
    pub fn dis_serving_create(engine: *mut dis_engine, cfg: *const dis_queue_config, max_active: i32, out: *mut *mut dis_serving) -> c_int;
    pub fn dis_serving_destroy(s: *mut dis_serving) -> c_int;
    pub fn dis_serving_last_error(s: *mut dis_serving) -> *const c_char;
    /// request_id: `Uuid::as_bytes()`; priority: DIS_PRIORITY_LOW = 0, NORMAL = 1, HIGH = 2 (the lanes of crates/core/src/queue.rs).
    pub fn dis_serving_submit(s: *mut dis_serving, request_id: *const u8, prompt: *const u32, n_prompt: i32,
                              params: *const dis_sampling_params, priority: i32) -> c_int;
    /// Client went away (requirements.md:85): the request leaves the queue or frees its slot and KV pages.
    pub fn dis_serving_cancel(s: *mut dis_serving, request_id: *const u8) -> c_int;
    /// One scheduler step = one forward pass; fills up to `cap` events.
    pub fn dis_serving_step(s: *mut dis_serving, events: *mut dis_serving_event, cap: i32, n_events: *mut i32) -> c_int;
    pub fn dis_serving_depth(s: *mut dis_serving, out: *mut dis_queue_depth) -> c_int;
    pub fn dis_serving_set_prefill_budget(s: *mut dis_serving, tokens: i32) -> c_int;
    /// mode 0: prompts advance in admission order; 1 (default): shortest prompt first inside a priority class, a waiting
    /// request gaining `aging_tokens_per_s` tokens of precedence per second.
    pub fn dis_serving_set_prefill_order(s: *mut dis_serving, mode: i32, aging_tokens_per_s: f64) -> c_int;
}
