"""Helpers for the -m gpu parity tests: torch is used for device memory only."""
import ctypes as C

import numpy as np
import torch

from distributed_inference_server_b200 import binding as B
from oracle import oracle as O

DEV = "cuda:0"
_KEEP = []  # device tensors created by the helpers stay alive until the test ends (see conftest.py)


def _keep(t):
    _KEEP.append(t)
    return t


def lib():
    return B.load()


def check(rc):
    assert rc == B.OK, lib().dis_last_error().decode()


def dev_bf16(a: np.ndarray) -> torch.Tensor:
    """fp32 numpy (bf16-representable values) -> device bf16 tensor."""
    return _keep(torch.from_numpy(np.ascontiguousarray(a, np.float32)).to(DEV).to(torch.bfloat16).contiguous())


def dev_f32(a: np.ndarray) -> torch.Tensor:
    return _keep(torch.from_numpy(np.ascontiguousarray(a, np.float32)).to(DEV).contiguous())


def dev_i32(a) -> torch.Tensor:
    return _keep(torch.from_numpy(np.ascontiguousarray(a, np.int32)).to(DEV).contiguous())


def host(t: torch.Tensor) -> np.ndarray:
    return t.detach().float().cpu().numpy()


def ptr(t):
    return C.c_void_p(t.data_ptr()) if t is not None else None


def sync():
    torch.cuda.synchronize()


def rand_bf16(rng, shape, scale=1.0):
    return O.bf16_round((rng.standard_normal(shape) * scale).astype(np.float32))


def shape_from_npz(z) -> O.ModelShape:
    return O.ModelShape(*[int(v) for v in z["shape"]], float(z["rope_theta"]), float(z["rms_eps"]))


def to_engine_shape(s: O.ModelShape):
    from distributed_inference_server_b200 import ModelShape
    return ModelShape(s.vocab, s.hidden, s.layers, s.n_heads, s.n_kv_heads, s.head_dim, s.inter, s.max_ctx,
                      s.rope_theta, s.rms_eps)


def oracle_tensors(m: "O.Oracle"):
    """(kind, layer, ndarray) triples in the engine's tensor numbering (same enum order)."""
    for kind, layer, arr in m.tensors():
        yield kind, layer, arr
