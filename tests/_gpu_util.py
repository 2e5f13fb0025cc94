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


# ---- guard bands (the stand-in for compute-sanitizer's memcheck, which is closed on this GPU pool): every device buffer
# the kernel tests hand to the C ABI sits between two 4 KiB bands of 0xA5 bytes; conftest.py checks the bands after each
# test, so a kernel that writes before or behind a buffer it was given (an output, a scratch area, an input) fails the test.
GUARD = 4096
_GUARDED = []  # (raw uint8 tensor, payload bytes, label)


def _guarded(shape, dtype, label=""):
    shape = (int(shape),) if np.isscalar(shape) else tuple(int(v) for v in shape)
    n = int(np.prod(shape, dtype=np.int64)) * torch.empty((), dtype=dtype).element_size()
    pad = (-n) % 256
    raw = torch.full((GUARD + n + pad + GUARD,), 0xA5, dtype=torch.uint8, device=DEV)
    _GUARDED.append((raw, n, label or f"{tuple(shape)} {dtype}"))
    return _keep(raw[GUARD:GUARD + n].view(dtype).view(shape))


def empty(shape, dtype=torch.float32, device=None):
    """torch.empty between guard bands (the payload starts as the 0xA5 pattern: finite in every dtype used here)."""
    return _guarded(shape, dtype)


def zeros(shape, dtype=torch.float32, device=None):
    t = _guarded(shape, dtype)
    t.zero_()
    return t


def _guarded_copy(src: torch.Tensor) -> torch.Tensor:
    t = _guarded(src.shape, src.dtype)
    t.copy_(src)
    return t


def check_guards():
    """Called by conftest.py after every test; returns the labels of the buffers whose bands were written to."""
    bad = []
    if _GUARDED:
        torch.cuda.synchronize()
    for raw, n, label in _GUARDED:
        lo, hi = raw[:GUARD], raw[GUARD + n:]
        if not (bool((lo == 0xA5).all()) and bool((hi == 0xA5).all())):
            below = int((lo != 0xA5).sum()); above = int((hi != 0xA5).sum())
            bad.append(f"{label}: {below} byte(s) written below, {above} behind")
    _GUARDED.clear()
    return bad


def lib():
    return B.load()


def check(rc):
    assert rc == B.OK, lib().dis_last_error().decode()


def dev_bf16(a: np.ndarray) -> torch.Tensor:
    """fp32 numpy (bf16-representable values) -> device bf16 tensor."""
    return _guarded_copy(torch.from_numpy(np.ascontiguousarray(a, np.float32)).to(DEV).to(torch.bfloat16))


def dev_f32(a: np.ndarray) -> torch.Tensor:
    return _guarded_copy(torch.from_numpy(np.ascontiguousarray(a, np.float32)).to(DEV))


def dev_i32(a) -> torch.Tensor:
    return _guarded_copy(torch.from_numpy(np.ascontiguousarray(a, np.int32)).to(DEV))


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
