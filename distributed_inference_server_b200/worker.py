"""Host-side mirror of the reference's worker interface, over the C ABI only.

Same names, argument meaning and error behaviour as the spec'd Rust items
(.kiro/specs/distributed-inference-server/design.md): WorkerConfig (316-322), InferenceWorker
{initialize, infer, shutdown, status, model_info} (334-342), InferenceBatch (241-248), InferenceParams
(666-671), BatchResult / RequestResult (345-360); FinishReason from crates/core/src/models.rs:26-32,
WorkerError from crates/core/src/error.rs:115-128.  The reference's own worker file is empty
(crates/inference/src/worker.rs:1).
"""
from __future__ import annotations

import ctypes as C
import enum
import uuid
from dataclasses import dataclass, field
from typing import Callable, List, Optional, Sequence

import numpy as np

from . import binding as B


@dataclass(frozen=True)
class ModelShape:
    vocab: int
    hidden: int
    layers: int
    n_heads: int
    n_kv_heads: int
    head_dim: int
    inter: int
    max_ctx: int = 8192
    rope_theta: float = 500000.0
    rms_eps: float = 1e-5

    def c(self) -> B.ModelConfig:
        return B.ModelConfig(self.vocab, self.hidden, self.layers, self.n_heads, self.n_kv_heads, self.head_dim,
                             self.inter, self.max_ctx, self.rope_theta, self.rms_eps)


TINYLLAMA = ModelShape(32000, 2048, 22, 32, 4, 64, 5632, 2048, 10000.0, 1e-5)
LLAMA3_8B = ModelShape(128256, 4096, 32, 32, 8, 128, 14336, 8192, 500000.0, 1e-5)
LLAMA3_70B = ModelShape(128256, 8192, 80, 64, 8, 128, 28672, 8192, 500000.0, 1e-5)


class FinishReason(enum.Enum):  # models.rs:26-32 (serde snake_case)
    Stop = "stop"
    Length = "length"
    StopSequence = "stop_sequence"


class WorkerError(Exception):  # error.rs:115-128
    ModelNotLoaded = "ModelNotLoaded"
    InferenceFailed = "InferenceFailed"
    Shutdown = "Shutdown"
    OutOfMemory = "OutOfMemory"

    def __init__(self, kind: str, message: str = ""):
        self.kind = kind
        text = {"ModelNotLoaded": "Model not loaded", "InferenceFailed": f"Inference failed: {message}",
                "Shutdown": "Worker shutdown", "OutOfMemory": "Out of memory"}[kind]  # error.rs Display strings
        super().__init__(text if kind == "InferenceFailed" or not message else f"{text} ({message})")


_KIND = {B.ERR_MODEL_NOT_LOADED: WorkerError.ModelNotLoaded, B.ERR_INFERENCE_FAILED: WorkerError.InferenceFailed,
         B.ERR_SHUTDOWN: WorkerError.Shutdown, B.ERR_OUT_OF_MEMORY: WorkerError.OutOfMemory,
         B.ERR_INVALID_ARGUMENT: WorkerError.InferenceFailed}


@dataclass
class InferenceParams:  # design.md:666-671; defaults models.rs:294-304
    max_tokens: int = 256
    temperature: float = 1.0
    top_p: float = 1.0
    stop_sequences: List[List[int]] = field(default_factory=list)
    seed: int = 0  # not in the reference; makes sampling reproducible


@dataclass
class BatchedRequest:  # design.md:258-266
    id: uuid.UUID = field(default_factory=uuid.uuid4)
    original_length: int = 0
    padded_length: int = 0
    params: InferenceParams = field(default_factory=InferenceParams)


@dataclass
class InferenceBatch:  # design.md:241-248
    input_ids: List[List[int]]
    attention_mask: Optional[List[List[int]]] = None
    requests: Optional[List[BatchedRequest]] = None
    max_new_tokens: int = 0
    id: uuid.UUID = field(default_factory=uuid.uuid4)


@dataclass
class RequestResult:  # design.md:353-360
    request_id: uuid.UUID
    tokens: List[int]
    finish_reason: FinishReason
    prompt_tokens: int
    completion_tokens: int
    failed: bool = False
    text: str = ""  # needs a tokenizer; the reference has none (validator.rs:60-65 counts bytes/4)


@dataclass
class BatchResult:  # design.md:345-350
    batch_id: uuid.UUID
    results: List[RequestResult]
    inference_time: float  # seconds
    tokens_generated: int
    time_to_first_token: float = 0.0


@dataclass
class WorkerStatus:  # design.md:290-297
    active_batches: int
    memory_used: int
    memory_available: int
    is_healthy: bool
    kv_pages_total: int = 0
    kv_pages_free: int = 0
    tokens_generated: int = 0
    last_step_ms: float = 0.0
    kernel_launches: int = 0


@dataclass
class WorkerConfig:  # design.md:316-322 + GPU-engine fields
    model: ModelShape
    model_path: Optional[str] = None  # resolved by the host into load_tensor calls; None = synthetic weights
    context_size: int = 0
    device: int = 0
    tp_rank: int = 0
    tp_size: int = 1
    max_batch: int = 32
    max_step_tokens: int = 0
    page_size: int = 0
    kv_pages: int = 0
    kv_mem_fraction: float = 0.0
    flags: int = 0
    eos_token: int = -1
    synthetic_seed: int = 0
    stream: int = 0
    nccl_id: Optional[bytes] = None


class Engine:
    """Thin RAII wrapper over the dis_engine_* step-wise entry points."""

    def __init__(self, cfg: WorkerConfig):
        self.lib = B.load()
        c = B.EngineConfig()
        c.model = cfg.model.c()
        if cfg.context_size:
            c.model.max_ctx = cfg.context_size
        c.device, c.tp_rank, c.tp_size = cfg.device, cfg.tp_rank, cfg.tp_size
        c.max_batch, c.max_step_tokens, c.page_size = cfg.max_batch, cfg.max_step_tokens, cfg.page_size
        c.kv_pages, c.kv_mem_fraction, c.flags, c.eos_token = cfg.kv_pages, cfg.kv_mem_fraction, cfg.flags, cfg.eos_token
        c.stream = cfg.stream or None
        if cfg.nccl_id:
            C.memmove(c.nccl_id, cfg.nccl_id, 128)
        self.h = C.c_void_p()
        rc = self.lib.dis_engine_create(C.byref(c), C.byref(self.h))
        if rc != B.OK:
            self.h = None
            raise WorkerError(_KIND[rc], self.lib.dis_last_error().decode())
        self.vocab = cfg.model.vocab

    def _check(self, rc: int):
        if rc != B.OK:
            raise WorkerError(_KIND[rc], self.lib.dis_engine_last_error(self.h).decode())

    def close(self):
        if getattr(self, "h", None):
            self.lib.dis_engine_destroy(self.h)
            self.h = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    def load_synthetic(self, seed: int):
        self._check(self.lib.dis_engine_load_synthetic(self.h, seed))

    def load_tensor(self, kind: int, layer: int, arr: np.ndarray):
        arr = np.ascontiguousarray(arr)
        if arr.dtype == np.float32:
            dt = B.F32
        elif arr.dtype == np.uint16:
            dt = B.BF16
        else:
            raise TypeError("load_tensor: float32 values or uint16 bf16 bit patterns")
        self._check(self.lib.dis_engine_load_tensor(self.h, kind, layer, arr.ctypes.data, dt))

    def read_tensor(self, kind: int, layer: int) -> np.ndarray:
        n = C.c_size_t(0)
        self.lib.dis_engine_read_tensor(self.h, kind, layer, None, C.byref(n))
        out = np.empty(n.value, np.uint16)
        self._check(self.lib.dis_engine_read_tensor(self.h, kind, layer, out.ctypes.data, C.byref(n)))
        return out

    def tp_export(self) -> bytes:
        """64-byte CUDA IPC handle of this rank's symmetric exchange buffer."""
        buf = (C.c_uint8 * 64)()
        self._check(self.lib.dis_engine_tp_export(self.h, buf))
        return bytes(buf)

    def tp_import(self, handles):
        """handles: the tp_size exported handles in rank order; switches the all-reduce to peer memory."""
        blob = b"".join(handles)
        arr = (C.c_uint8 * len(blob)).from_buffer_copy(blob)
        self._check(self.lib.dis_engine_tp_import(self.h, arr))

    def finalize(self):
        self._check(self.lib.dis_engine_finalize(self.h))

    def seq_add(self, prompt: Sequence[int], params: InferenceParams) -> int:
        p = np.ascontiguousarray(prompt, dtype=np.uint32)
        sp, keep = _sampling(params)
        slot = C.c_int32(-1)
        self._check(self.lib.dis_engine_seq_add(self.h, p.ctypes.data, len(p), C.byref(sp), C.byref(slot)))
        return slot.value

    def seq_remove(self, slot: int):
        self._check(self.lib.dis_engine_seq_remove(self.h, slot))

    def _step(self, fn, slots):
        s = np.ascontiguousarray(slots, dtype=np.int32)
        toks = np.zeros(len(s), np.uint32)
        fin = np.zeros(len(s), np.int32)
        self._check(fn(self.h, len(s), s.ctypes.data, toks.ctypes.data, fin.ctypes.data))
        return toks, fin

    def prefill(self, slots):
        return self._step(self.lib.dis_engine_prefill, slots)

    def decode(self, slots):
        return self._step(self.lib.dis_engine_decode, slots)

    def step(self, slots, max_prefill_tokens: int = 0):
        """One mixed pass (dis_engine_step): (tokens, finish, produced) per listed sequence."""
        s = np.ascontiguousarray(slots, dtype=np.int32)
        toks = np.zeros(len(s), np.uint32)
        fin = np.zeros(len(s), np.int32)
        prod = np.zeros(len(s), np.int32)
        self._check(self.lib.dis_engine_step(self.h, len(s), s.ctypes.data, int(max_prefill_tokens), toks.ctypes.data,
                                             fin.ctypes.data, prod.ctypes.data))
        return toks, fin, prod

    def profile_decode(self, slots):
        """One eager decode step timed per kernel class: (tokens, finish, {class: (ms, launches)})."""
        s = np.ascontiguousarray(slots, dtype=np.int32)
        toks = np.zeros(len(s), np.uint32)
        fin = np.zeros(len(s), np.int32)
        ms = np.zeros(len(B.PROF_CLASSES), np.float32)
        cnt = np.zeros(len(B.PROF_CLASSES), np.int32)
        self._check(self.lib.dis_engine_profile_decode(self.h, len(s), s.ctypes.data, toks.ctypes.data, fin.ctypes.data,
                                                       ms.ctypes.data, cnt.ctypes.data))
        return toks, fin, {k: (float(ms[i]), int(cnt[i])) for i, k in enumerate(B.PROF_CLASSES)}

    def replay_last_step(self, n: int) -> float:
        """Average device milliseconds of the last decode step's CUDA graph replayed n times back to back."""
        ms = C.c_float()
        self._check(self.lib.dis_engine_replay_last_step(self.h, int(n), C.byref(ms)))
        return float(ms.value)

    def last_logprobs(self, slots) -> np.ndarray:
        """TokenEvent.logprob of the token each listed sequence got from its latest prefill/decode call."""
        s = np.ascontiguousarray(slots, dtype=np.int32)
        out = np.zeros(len(s), np.float32)
        self._check(self.lib.dis_engine_last_logprobs(self.h, len(s), s.ctypes.data, out.ctypes.data))
        return out

    def last_logits(self, n_rows: int) -> np.ndarray:
        out = np.empty((n_rows, self.vocab), np.float32)
        self._check(self.lib.dis_engine_last_logits(self.h, out.ctypes.data, n_rows))
        return out

    def status(self) -> WorkerStatus:
        s = B.WorkerStatus()
        self._check(self.lib.dis_engine_status(self.h, C.byref(s)))
        return WorkerStatus(s.active_sequences, s.memory_used, s.memory_available, bool(s.is_healthy), s.kv_pages_total,
                            s.kv_pages_free, s.tokens_generated, s.last_step_ms, s.kernel_launches)


    def prefix_stats(self) -> dict:
        """CacheStats of the prefix cache (design.md:404-411): pages registered / evictable / free, requests and
        tokens looked up and reused, evictions."""
        s = B.PrefixStats()
        self._check(self.lib.dis_engine_prefix_stats(self.h, C.byref(s)))
        return {n: int(getattr(s, n)) for n, _ in B.PrefixStats._fields_}


def _sampling(p: InferenceParams):
    sp = B.SamplingParams()
    sp.max_tokens, sp.temperature, sp.top_p, sp.seed = int(p.max_tokens), float(p.temperature), float(p.top_p), int(p.seed)
    flat = np.array([t for s in p.stop_sequences for t in s], dtype=np.uint32)
    lens = np.array([len(s) for s in p.stop_sequences], dtype=np.int32)
    sp.n_stop = len(lens)
    sp.stop_tokens = flat.ctypes.data_as(C.POINTER(C.c_uint32)) if len(flat) else None
    sp.stop_lens = lens.ctypes.data_as(C.POINTER(C.c_int32)) if len(lens) else None
    return sp, (flat, lens)


class InferenceWorker:
    """design.md:334-342.  `infer` blocks; an async host wraps it in spawn_blocking."""

    def __init__(self, config: WorkerConfig):
        self.config = config
        self.engine: Optional[Engine] = None
        self._shutdown = False

    def initialize(self, tensors=None) -> None:
        """Create the engine and load weights: `tensors` yields (kind, layer, ndarray) or, when None,
        random-init weights are generated on the device from config.synthetic_seed."""
        if self._shutdown:
            raise WorkerError(WorkerError.Shutdown)
        if self.engine is not None:
            return
        eng = Engine(self.config)
        if tensors is None:
            eng.load_synthetic(self.config.synthetic_seed)
        else:
            for kind, layer, arr in tensors:
                eng.load_tensor(kind, layer, arr)
        eng.finalize()
        self.engine = eng

    def shutdown(self) -> None:
        self._shutdown = True
        if self.engine is not None:
            self.engine.close()
            self.engine = None

    def status(self) -> WorkerStatus:
        if self.engine is None:
            return WorkerStatus(0, 0, 0, False)
        return self.engine.status()

    def model_info(self) -> ModelShape:
        return self.config.model

    def infer(self, batch: InferenceBatch,
              on_token: Optional[Callable[..., bool]] = None) -> BatchResult:
        """Run a padded batch to completion.  N requests in -> N results out (design.md:806-810); a
        request that cannot be served is marked failed without failing the batch (design.md:812-816).
        on_token(request_index, token, index, finished[, logprob]) is TokenEvent::Token (models.rs:272-277); a
        callback that takes five arguments also receives the token's log-probability."""
        if self._shutdown:
            raise WorkerError(WorkerError.Shutdown)
        if self.engine is None:
            raise WorkerError(WorkerError.ModelNotLoaded)
        n = len(batch.input_ids)
        reqs = batch.requests or [BatchedRequest() for _ in range(n)]
        if n == 0:
            return BatchResult(batch.id, [], 0.0, 0)
        lens = np.zeros(n, np.int32)
        for i, row in enumerate(batch.input_ids):
            if reqs[i].original_length:
                lens[i] = reqs[i].original_length
            elif batch.attention_mask is not None:
                lens[i] = int(np.sum(np.asarray(batch.attention_mask[i]) != 0))
            else:
                lens[i] = len(row)
        stride = max(1, max(len(r) for r in batch.input_ids))
        ids = np.zeros((n, stride), np.uint32)
        for i, row in enumerate(batch.input_ids):
            ids[i, :len(row)] = row
        params = (B.SamplingParams * n)()
        keep = []
        max_new = 1
        for i in range(n):
            p = reqs[i].params
            if batch.max_new_tokens:
                p = InferenceParams(min(p.max_tokens, batch.max_new_tokens), p.temperature, p.top_p, p.stop_sequences, p.seed)
            sp, k = _sampling(p)
            params[i] = sp
            keep.append(k)
            max_new = max(max_new, int(p.max_tokens))
        toks = np.zeros((n, max_new), np.uint32)
        out_lens = np.zeros(n, np.int32)
        fin = np.zeros(n, np.int32)
        res = B.BatchResult()
        cb = None
        if on_token is not None:
            import inspect
            try:
                want_lp = len(inspect.signature(on_token).parameters) >= 5
            except (TypeError, ValueError):
                want_lp = False

            def _cb(_user, idx, tok, index, f, lp):
                args = (idx, tok, index, f != B.FINISH_NONE) + ((lp,) if want_lp else ())
                return 0 if on_token(*args) else 1
            cb = B.TOKEN_CALLBACK(_cb)
        eng = self.engine
        rc = eng.lib.dis_worker_infer(eng.h, n, ids.ctypes.data, stride, lens.ctypes.data, params, toks.ctypes.data,
                                      max_new, out_lens.ctypes.data, fin.ctypes.data, C.byref(res),
                                      C.cast(cb, C.c_void_p) if cb else None, None)
        eng._check(rc)
        results = []
        fr = {B.FINISH_STOP: FinishReason.Stop, B.FINISH_STOP_SEQUENCE: FinishReason.StopSequence}
        for i in range(n):
            t = toks[i, :out_lens[i]].tolist()
            results.append(RequestResult(reqs[i].id, t, fr.get(int(fin[i]), FinishReason.Length), int(lens[i]), len(t),
                                         failed=int(fin[i]) == B.FINISH_FAILED))
        return BatchResult(batch.id, results, res.inference_ms / 1e3, int(res.tokens_generated), res.ttft_ms / 1e3)


class QueueFull(Exception):
    """QueueError::Full (crates/core/src/error.rs:79-89): the admission queue refuses the request (HTTP 503)."""


class ContinuousBatcher:
    """Admission queue + continuous batching over one engine (dis_serving_*).  `engine=None` gives the queue alone.

    Queue semantics follow crates/core/src/queue.rs (High -> Normal -> Low FIFO lanes, watermark back-pressure,
    expiry); events follow TokenEvent (crates/core/src/models.rs:268-288)."""

    def __init__(self, engine: Optional[Engine], max_active: int = 32, high_watermark: int = 1000, low_watermark: int = 500,
                 max_queue_size: int = 2000, request_timeout_s: float = 30.0):
        self.lib = B.load()
        self.engine = engine
        cfg = B.QueueConfig(high_watermark, low_watermark, max_queue_size, request_timeout_s)
        self.h = C.c_void_p()
        rc = self.lib.dis_serving_create(engine.h if engine is not None else None, C.byref(cfg), max_active, C.byref(self.h))
        if rc != B.OK:
            raise ValueError("dis_serving_create: bad queue configuration")
        self._events = (B.ServingEvent * 1024)()
        self._ids = {}  # 16-byte wire id -> the object the caller submitted (int, uuid.UUID or bytes)

    def _wire_id(self, request_id) -> bytes:
        """RequestID is a Uuid (crates/core/src/types.rs:7): 16 bytes on the wire.  Ints are accepted for tests."""
        if isinstance(request_id, uuid.UUID):
            b = request_id.bytes
        elif isinstance(request_id, (bytes, bytearray)):
            b = bytes(request_id)
            if len(b) != 16:
                raise ValueError("request id must be 16 bytes")
        elif isinstance(request_id, str):
            import hashlib
            b = hashlib.blake2b(request_id.encode(), digest_size=16).digest()
        else:
            b = int(request_id).to_bytes(16, "little")
        self._ids.setdefault(b, request_id)
        return b

    def close(self):
        if getattr(self, "h", None):
            self.lib.dis_serving_destroy(self.h)
            self.h = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    def submit(self, request_id, prompt: Sequence[int], params: InferenceParams, priority: int = B.PRIORITY_NORMAL,
               now: Optional[float] = None) -> None:
        p = np.ascontiguousarray(prompt, dtype=np.uint32)
        sp, keep = _sampling(params)
        rid = self._wire_id(request_id)
        if now is None:
            rc = self.lib.dis_serving_submit(self.h, rid, p.ctypes.data, len(p), C.byref(sp), priority)
        else:
            rc = self.lib.dis_serving_submit_at(self.h, rid, p.ctypes.data, len(p), C.byref(sp), priority, now)
        if rc == B.ERR_QUEUE_FULL:
            raise QueueFull()
        if rc != B.OK:
            raise WorkerError(_KIND.get(rc, WorkerError.InferenceFailed), "submit")

    def set_prefill_budget(self, tokens: int) -> None:
        if self.lib.dis_serving_set_prefill_budget(self.h, int(tokens)) != B.OK:
            raise ValueError("prefill budget must be >= 1")

    def set_prefill_order(self, shortest_first: bool = True, aging_tokens_per_s: float = 4096.0) -> None:
        """Order in which admitted prompts are read inside a pass (dis_serving_set_prefill_order)."""
        if self.lib.dis_serving_set_prefill_order(self.h, 1 if shortest_first else 0, float(aging_tokens_per_s)) != B.OK:
            raise ValueError("aging must be >= 0")

    def cancel(self, request_id) -> bool:
        """Client disconnect (requirements.md:85); False if the request is neither queued nor running."""
        return self.lib.dis_serving_cancel(self.h, self._wire_id(request_id)) == B.OK

    def step(self, now: Optional[float] = None):
        """One scheduler iteration; returns the list of events it produced."""
        n = C.c_int32(0)
        if now is None:
            rc = self.lib.dis_serving_step(self.h, self._events, len(self._events), C.byref(n))
        else:
            rc = self.lib.dis_serving_step_at(self.h, now, self._events, len(self._events), C.byref(n))
        if rc != B.OK:
            raise WorkerError(_KIND.get(rc, WorkerError.InferenceFailed), self.lib.dis_serving_last_error(self.h).decode())
        self.last_logprobs = [e.logprob for e in self._events[:n.value]]  # parallel to the returned list
        return [(self._ids.get(bytes(e.request_id), bytes(e.request_id)), e.kind, e.token, e.index, e.finish_reason,
                 e.prompt_tokens, e.waited_ms) for e in self._events[:n.value]]

    def depth(self) -> B.QueueDepth:
        d = B.QueueDepth()
        self.lib.dis_serving_depth(self.h, C.byref(d))
        return d

    def debug_pop(self, max_count: int):
        ids = (C.c_uint8 * (16 * max_count))()
        n = C.c_int32(0)
        self.lib.dis_serving_debug_pop(self.h, max_count, ids, C.byref(n))
        raw = bytes(ids)
        return [self._ids.get(raw[16 * i:16 * i + 16], raw[16 * i:16 * i + 16]) for i in range(n.value)]

    def run_until_idle(self, max_steps: int = 1_000_000):
        """Drive the scheduler until queue and engine are empty; returns {request_id: (tokens, finish_reason)}."""
        out = {}
        for _ in range(max_steps):
            evs = self.step()
            for rid, kind, tok, idx, fin, ptoks, waited in evs:
                rec = out.setdefault(rid, {"tokens": [], "finish": None, "ttft_ms": None, "prompt_tokens": ptoks})
                if kind == B.EVENT_TOKEN:
                    rec["tokens"].append(tok)
                    if idx == 0:
                        rec["ttft_ms"] = waited
                else:
                    rec["finish"] = fin if kind == B.EVENT_DONE else "expired"
            d = self.depth()
            if d.total == 0 and d.active == 0 and not evs:
                break
        return out
