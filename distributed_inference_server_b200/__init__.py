"""B200-native decoder forward path for its-me-ojas/distributed-inference-server.

csrc/ holds the hand-written sm_100a kernels, the engine and the C ABI (include/dis_engine.h);
binding.py is the ctypes binding; worker.py mirrors the reference's spec'd InferenceWorker interface
(.kiro/specs/distributed-inference-server/design.md:310-361) on top of that ABI.
"""
from . import binding  # noqa: F401
from .worker import (BatchResult, BatchedRequest, FinishReason, InferenceBatch, InferenceParams,  # noqa: F401
                     InferenceWorker, RequestResult, WorkerConfig, WorkerError, WorkerStatus, Engine, ContinuousBatcher, QueueFull,
                     LLAMA3_8B, LLAMA3_70B, TINYLLAMA, ModelShape)
