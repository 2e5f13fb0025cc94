"""ctypes binding of include/dis_engine.h.  No compute happens in Python and there is no fallback:
if libdis_engine.so is missing the import fails, and every compute entry point fails without a
B200-class CUDA device."""
from __future__ import annotations

import ctypes as C
import os

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(_HERE, "libdis_engine.so")

OK, ERR_MODEL_NOT_LOADED, ERR_INFERENCE_FAILED, ERR_SHUTDOWN, ERR_OUT_OF_MEMORY, ERR_INVALID_ARGUMENT, ERR_QUEUE_FULL = range(7)
PRIORITY_LOW, PRIORITY_NORMAL, PRIORITY_HIGH = 0, 1, 2
EVENT_TOKEN, EVENT_DONE, EVENT_EXPIRED = 0, 1, 2
FINISH_NONE, FINISH_STOP, FINISH_LENGTH, FINISH_STOP_SEQUENCE, FINISH_FAILED = range(5)
(T_EMBED, T_ATTN_NORM, T_WQ, T_WK, T_WV, T_WO, T_MLP_NORM, T_WGATE, T_WUP, T_WDOWN, T_FINAL_NORM,
 T_LM_HEAD) = range(12)
F32, BF16 = 0, 1
FLAG_NO_GRAPH, FLAG_FORCE_GEMV, FLAG_NO_PDL, FLAG_PREFIX_CACHE = 1, 2, 4, 8
PROF_CLASSES = ("norm", "qkv", "rope", "attn", "oproj", "gateup", "down", "allreduce", "head_norm", "lm_head",
                "sample")


class ModelConfig(C.Structure):
    _fields_ = [(n, C.c_int32) for n in
                ("vocab", "hidden", "layers", "n_heads", "n_kv_heads", "head_dim", "inter", "max_ctx")] + \
               [("rope_theta", C.c_float), ("rms_eps", C.c_float)]


class EngineConfig(C.Structure):
    _fields_ = [("model", ModelConfig), ("device", C.c_int32), ("tp_rank", C.c_int32), ("tp_size", C.c_int32),
                ("max_batch", C.c_int32), ("max_step_tokens", C.c_int32), ("page_size", C.c_int32),
                ("kv_pages", C.c_int64), ("kv_mem_fraction", C.c_float), ("flags", C.c_uint32),
                ("eos_token", C.c_int32), ("stream", C.c_void_p), ("nccl_id", C.c_uint8 * 128)]


class SamplingParams(C.Structure):
    _fields_ = [("max_tokens", C.c_int32), ("temperature", C.c_float), ("top_p", C.c_float), ("seed", C.c_uint64),
                ("stop_tokens", C.POINTER(C.c_uint32)), ("stop_lens", C.POINTER(C.c_int32)), ("n_stop", C.c_int32)]


class WorkerStatus(C.Structure):
    _fields_ = [("active_sequences", C.c_int32), ("is_healthy", C.c_int32), ("memory_used", C.c_uint64),
                ("memory_available", C.c_uint64), ("kv_pages_total", C.c_int64), ("kv_pages_free", C.c_int64),
                ("tokens_generated", C.c_uint64), ("last_step_ms", C.c_double), ("kernel_launches", C.c_uint64)]


class BatchResult(C.Structure):
    _fields_ = [("inference_ms", C.c_double), ("ttft_ms", C.c_double), ("tokens_generated", C.c_uint64)]


class QueueConfig(C.Structure):
    _fields_ = [("high_watermark", C.c_int32), ("low_watermark", C.c_int32), ("max_queue_size", C.c_int32),
                ("request_timeout_s", C.c_double)]


class QueueDepth(C.Structure):
    _fields_ = [(n, C.c_int32) for n in ("high", "normal", "low", "total", "accepting", "active")]


class PrefixStats(C.Structure):
    _fields_ = [(n, C.c_int64) for n in ("cached_pages", "evictable_pages", "free_pages")] + \
               [(n, C.c_uint64) for n in ("lookups", "hits", "lookup_tokens", "hit_tokens", "evictions")]


class ServingEvent(C.Structure):
    _fields_ = [("request_id", C.c_uint8 * 16), ("kind", C.c_int32), ("token", C.c_uint32), ("index", C.c_int32),
                ("finish_reason", C.c_int32), ("prompt_tokens", C.c_int32), ("waited_ms", C.c_double),
                ("logprob", C.c_float)]


TOKEN_CALLBACK = C.CFUNCTYPE(C.c_int, C.c_void_p, C.c_int32, C.c_uint32, C.c_int32, C.c_int32, C.c_float)

u16p, u32p, i32p, f32p, u64p, vp = (C.POINTER(C.c_uint16), C.POINTER(C.c_uint32), C.POINTER(C.c_int32),
                                    C.POINTER(C.c_float), C.POINTER(C.c_uint64), C.c_void_p)

# name -> (restype, argtypes): every symbol include/dis_engine.h declares
PROTOTYPES = {
    "dis_abi_version": (C.c_int, []),
    "dis_build_id": (C.c_char_p, []),
    "dis_engine_create": (C.c_int, [C.POINTER(EngineConfig), C.POINTER(vp)]),
    "dis_engine_destroy": (C.c_int, [vp]),
    "dis_last_error": (C.c_char_p, []),
    "dis_engine_last_error": (C.c_char_p, [vp]),
    "dis_engine_load_tensor": (C.c_int, [vp, C.c_int, C.c_int, vp, C.c_int]),
    "dis_engine_load_synthetic": (C.c_int, [vp, C.c_uint64]),
    "dis_engine_read_tensor": (C.c_int, [vp, C.c_int, C.c_int, vp, C.POINTER(C.c_size_t)]),
    "dis_engine_finalize": (C.c_int, [vp]),
    "dis_engine_seq_add": (C.c_int, [vp, vp, C.c_int32, C.POINTER(SamplingParams), i32p]),
    "dis_engine_seq_remove": (C.c_int, [vp, C.c_int32]),
    "dis_engine_prefill": (C.c_int, [vp, C.c_int32, vp, vp, vp]),
    "dis_engine_decode": (C.c_int, [vp, C.c_int32, vp, vp, vp]),
    "dis_engine_step": (C.c_int, [vp, C.c_int32, vp, C.c_int32, vp, vp, vp]),
    "dis_engine_last_logits": (C.c_int, [vp, vp, C.c_int32]),
    "dis_engine_last_logprobs": (C.c_int, [vp, C.c_int32, vp, vp]),
    "dis_engine_profile_decode": (C.c_int, [vp, C.c_int32, vp, vp, vp, vp, vp]),
    "dis_engine_replay_last_step": (C.c_int, [vp, C.c_int32, f32p]),
    "dis_engine_status": (C.c_int, [vp, C.POINTER(WorkerStatus)]),
    "dis_engine_model_info": (C.c_int, [vp, C.POINTER(ModelConfig)]),
    "dis_tp_unique_id": (C.c_int, [vp]),
    "dis_engine_tp_export": (C.c_int, [vp, vp]),
    "dis_engine_tp_import": (C.c_int, [vp, vp]),
    "dis_worker_infer": (C.c_int, [vp, C.c_int32, vp, C.c_int32, vp, C.POINTER(SamplingParams), vp, C.c_int32, vp, vp,
                                   C.POINTER(BatchResult), vp, vp]),
    "dis_serving_create": (C.c_int, [vp, C.POINTER(QueueConfig), C.c_int32, C.POINTER(vp)]),
    "dis_serving_destroy": (C.c_int, [vp]),
    "dis_serving_last_error": (C.c_char_p, [vp]),
    "dis_serving_submit": (C.c_int, [vp, vp, vp, C.c_int32, C.POINTER(SamplingParams), C.c_int32]),
    "dis_serving_submit_at": (C.c_int, [vp, vp, vp, C.c_int32, C.POINTER(SamplingParams), C.c_int32, C.c_double]),
    "dis_serving_cancel": (C.c_int, [vp, vp]),
    "dis_serving_step": (C.c_int, [vp, C.POINTER(ServingEvent), C.c_int32, i32p]),
    "dis_serving_step_at": (C.c_int, [vp, C.c_double, C.POINTER(ServingEvent), C.c_int32, i32p]),
    "dis_serving_depth": (C.c_int, [vp, C.POINTER(QueueDepth)]),
    "dis_serving_set_prefill_budget": (C.c_int, [vp, C.c_int32]),
    "dis_serving_set_prefill_order": (C.c_int, [vp, C.c_int32, C.c_double]),
    "dis_serving_order_prompts": (C.c_int, [C.c_int32, vp, vp, vp, C.c_double, vp]),
    "dis_serving_plan_step": (C.c_int, [C.c_int32, vp, C.c_int32, C.c_int32, vp]),
    "dis_serving_debug_pop": (C.c_int, [vp, C.c_int32, vp, i32p]),
    "dis_engine_prefix_stats": (C.c_int, [vp, C.POINTER(PrefixStats)]),
    "dis_prefix_index_create": (C.c_int, [C.c_int32, C.c_int64, C.POINTER(vp)]),
    "dis_prefix_index_destroy": (C.c_int, [vp]),
    "dis_prefix_index_admit": (C.c_int, [vp, vp, C.c_int32, vp, C.c_int32, i32p, i32p]),
    "dis_prefix_index_commit": (C.c_int, [vp, vp, C.c_int32, vp, C.c_int32]),
    "dis_prefix_index_release": (C.c_int, [vp, vp, C.c_int32]),
    "dis_prefix_index_stats": (C.c_int, [vp, C.POINTER(PrefixStats)]),
    "dis_k_add_rmsnorm": (C.c_int, [vp, vp, vp, vp, C.c_int32, C.c_int64, vp, vp, vp, C.c_int32, C.c_int32,
                                    C.c_float, vp]),
    "dis_k_gemm": (C.c_int, [vp, vp, vp, C.c_int32, C.c_int32, C.c_int32, C.c_int32, vp]),
    "dis_k_gemm_pick_split": (C.c_int, [C.c_int32, C.c_int32, C.c_int32]),
    "dis_k_gemm_silu": (C.c_int, [vp, vp, vp, C.c_int32, C.c_int32, C.c_int32, vp]),
    "dis_k_gemv": (C.c_int, [vp, vp, vp, C.c_int32, C.c_int32, C.c_int32, vp]),
    "dis_k_gemv_silu": (C.c_int, [vp, vp, vp, C.c_int32, C.c_int32, C.c_int32, vp]),
    "dis_k_rope_kv_write": (C.c_int, [vp, C.c_int32, C.c_int64, vp, vp, vp, C.c_int32, vp, vp, vp, vp, C.c_int32,
                                      C.c_int32, C.c_int32, C.c_int32, C.c_int32, vp]),
    "dis_k_attn_scratch_elems": (C.c_size_t, [C.c_int32, C.c_int32, C.c_int32, C.c_int32]),
    "dis_k_attn_paged": (C.c_int, [vp, vp, C.c_int64, vp, C.c_int32, vp, vp, vp, vp, C.c_int32, C.c_int32, C.c_int32,
                                   C.c_int32, C.c_int32, C.c_int32, vp]),
    "dis_k_attn_paged_static": (C.c_int, [vp, vp, C.c_int64, vp, C.c_int32, vp, vp, vp, C.c_int32, C.c_int32, C.c_int32, C.c_int32,
                                          C.c_int32, C.c_int32, vp]),
    "dis_k_attn_prefill": (C.c_int, [vp, vp, C.c_int64, vp, C.c_int32, vp, vp, vp, vp, vp, C.c_int32, C.c_int32, C.c_int32,
                                     C.c_int32, C.c_int32, vp]),
    "dis_k_attn_prefill_block_tokens": (C.c_int, [C.c_int32, C.c_int32, C.c_int32]),
    "dis_k_sample_vocab_parallel_area_bytes": (C.c_size_t, [C.c_int32, C.c_int32]),
    "dis_k_sample_vocab_parallel": (C.c_int, [vp, C.c_int32, C.c_int32, C.c_int32, vp, vp, vp, vp, C.c_uint32, vp, C.c_int32, vp, vp, vp]),
    "dis_k_sample": (C.c_int, [vp, C.c_int64, C.c_int32, C.c_int32, vp, vp, vp, vp, vp, vp, vp]),
    "dis_k_synth_bf16": (C.c_int, [vp, C.c_size_t, C.c_uint64, C.c_int32, C.c_int32, C.c_float, C.c_float, vp]),
}

_lib = None


def load() -> C.CDLL:
    """Load the C-ABI library.  Raises if it has not been built: there is no Python/CPU fallback."""
    global _lib
    if _lib is None:
        if not os.path.exists(LIB_PATH):
            raise ImportError(f"{LIB_PATH} is missing: run `python -m distributed_inference_server_b200.build` "
                              "(or __graft_entry__.build()); this package has no CPU fallback")
        lib = C.CDLL(LIB_PATH)
        for name, (res, args) in PROTOTYPES.items():
            fn = getattr(lib, name)
            fn.restype = res
            fn.argtypes = args
        _lib = lib
    return _lib
