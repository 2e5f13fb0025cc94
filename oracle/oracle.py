"""ctypes wrapper around oracle/llama_oracle.c.  TEST INFRASTRUCTURE ONLY.

Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference legs may
import this module; the product package never does.  Parity is unpinned by the reference (it
has no forward pass: crates/inference/src/worker.rs:1); the restatement is pinned against
transformers' LlamaForCausalLM through tests/golden/ (see oracle/make_golden.py).
"""
from __future__ import annotations

import ctypes as C
import os
import subprocess
from dataclasses import dataclass

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
_SRC = os.path.join(_HERE, "llama_oracle.c")
_LIB = os.path.join(_HERE, "_build", "liboracle.so")

EMBED, ATTN_NORM, WQ, WK, WV, WO, MLP_NORM, WGATE, WUP, WDOWN, FINAL_NORM, LM_HEAD = range(12)
PER_LAYER = (ATTN_NORM, WQ, WK, WV, WO, MLP_NORM, WGATE, WUP, WDOWN)
GLOBAL = (EMBED, FINAL_NORM, LM_HEAD)


class _Cfg(C.Structure):
    _fields_ = [(n, C.c_int32) for n in
                ("vocab", "hidden", "layers", "n_heads", "n_kv_heads", "head_dim", "inter", "max_ctx")] + \
               [("rope_theta", C.c_float), ("rms_eps", C.c_float)]


@dataclass(frozen=True)
class ModelShape:
    vocab: int
    hidden: int
    layers: int
    n_heads: int
    n_kv_heads: int
    head_dim: int
    inter: int
    max_ctx: int = 2048
    rope_theta: float = 500000.0
    rms_eps: float = 1e-5

    def c(self) -> _Cfg:
        return _Cfg(self.vocab, self.hidden, self.layers, self.n_heads, self.n_kv_heads, self.head_dim,
                    self.inter, self.max_ctx, self.rope_theta, self.rms_eps)


TINYLLAMA = ModelShape(32000, 2048, 22, 32, 4, 64, 5632, 2048, 10000.0, 1e-5)
LLAMA3_8B = ModelShape(128256, 4096, 32, 32, 8, 128, 14336, 8192, 500000.0, 1e-5)
LLAMA3_70B = ModelShape(128256, 8192, 80, 64, 8, 128, 28672, 8192, 500000.0, 1e-5)


def build(force: bool = False) -> str:
    """Compile the C oracle (gcc, OpenMP).  Portable flags: the .so is built where nvcc is and
    travels to the GPU box."""
    if not force and os.path.exists(_LIB) and os.path.getmtime(_LIB) >= os.path.getmtime(_SRC):
        return _LIB
    os.makedirs(os.path.dirname(_LIB), exist_ok=True)
    cmd = ["gcc", "-O3", "-fopenmp", "-mavx2", "-mfma", "-fPIC", "-shared", "-fvisibility=hidden",
           "-o", _LIB, _SRC, "-lm"]
    subprocess.run(cmd, check=True)
    return _LIB


_lib = None


def lib() -> C.CDLL:
    global _lib
    if _lib is None:
        _lib = C.CDLL(build())
        L = _lib
        fp = C.POINTER(C.c_float)
        ip = C.POINTER(C.c_int)
        L.orc_create.restype = C.c_void_p
        L.orc_create.argtypes = [C.POINTER(_Cfg), C.c_int]
        L.orc_destroy.argtypes = [C.c_void_p]
        L.orc_tensor.restype = fp
        L.orc_tensor.argtypes = [C.c_void_p, C.c_int, C.c_int]
        L.orc_tensor_elems.restype = C.c_size_t
        L.orc_tensor_elems.argtypes = [C.POINTER(_Cfg), C.c_int]
        L.orc_fill_synthetic.argtypes = [C.c_void_p, C.c_uint64, C.c_int]
        L.orc_synth_tensor.argtypes = [fp, C.c_size_t, C.c_uint64, C.c_int, C.c_int, C.c_float, C.c_float, C.c_int]
        L.orc_synth_params.argtypes = [C.POINTER(_Cfg), C.c_int, fp, fp]
        L.orc_uniform.restype = C.c_float
        L.orc_uniform.argtypes = [C.c_uint64, C.c_uint64, C.c_uint64]
        L.orc_seq_new.restype = C.c_int
        L.orc_seq_new.argtypes = [C.c_void_p]
        L.orc_seq_free.argtypes = [C.c_void_p, C.c_int]
        L.orc_seq_fill_synthetic.argtypes = [C.c_void_p, C.c_int, C.c_int, C.c_uint64]
        L.orc_forward.restype = C.c_int
        L.orc_forward.argtypes = [C.c_void_p, C.c_int, ip, ip, ip, ip, fp, fp]
        L.orc_generate.restype = C.c_int
        L.orc_generate.argtypes = [C.c_void_p, C.c_int, ip, ip, C.c_int, C.c_int, C.c_float, C.c_float,
                                   C.POINTER(C.c_uint64), C.c_int, ip, ip, fp, fp, C.c_int]
        L.orc_rmsnorm.argtypes = [fp, fp, fp, C.c_int, C.c_int, C.c_float, C.c_int]
        L.orc_linear.argtypes = [fp, fp, fp, C.c_int, C.c_int, C.c_int]
        L.orc_rope_table.argtypes = [fp, fp, C.c_int, C.c_int, C.c_float]
        L.orc_rope.argtypes = [fp, ip, C.c_int, C.c_int, C.c_int, fp, fp]
        L.orc_attention_one.argtypes = [fp, fp, fp, fp, C.c_int, C.c_int, C.c_int, C.c_int, C.c_size_t]
        L.orc_silu_mul.argtypes = [fp, fp, fp, C.c_size_t, C.c_int]
        L.orc_argmax.restype = C.c_int
        L.orc_argmax.argtypes = [fp, C.c_int, fp]
        L.orc_sample.restype = C.c_int
        L.orc_sample.argtypes = [fp, C.c_int, C.c_float, C.c_float, C.c_uint64, C.c_uint64, fp, fp]
        L.orc_bf16_round_inplace.argtypes = [fp, C.c_size_t]
        L.orc_num_threads.restype = C.c_int
        L.orc_set_num_threads.argtypes = [C.c_int]
    return _lib


def _fp(a: np.ndarray):
    assert a.dtype == np.float32 and a.flags["C_CONTIGUOUS"]
    return a.ctypes.data_as(C.POINTER(C.c_float))


def _ip(a: np.ndarray):
    assert a.dtype == np.int32 and a.flags["C_CONTIGUOUS"]
    return a.ctypes.data_as(C.POINTER(C.c_int))


def bf16_round(a: np.ndarray) -> np.ndarray:
    out = np.ascontiguousarray(a, dtype=np.float32).copy()
    lib().orc_bf16_round_inplace(_fp(out.reshape(-1)), out.size)
    return out


def to_bf16_bits(a: np.ndarray) -> np.ndarray:
    """fp32 values that are already bf16-representable -> uint16 bit patterns."""
    return (np.ascontiguousarray(a, dtype=np.float32).view(np.uint32) >> 16).astype(np.uint16)


def from_bf16_bits(a: np.ndarray) -> np.ndarray:
    return (a.astype(np.uint32) << 16).view(np.float32)


def tensor_shape(s: ModelShape, kind: int):
    H, qd, kd = s.hidden, s.n_heads * s.head_dim, s.n_kv_heads * s.head_dim
    return {EMBED: (s.vocab, H), LM_HEAD: (s.vocab, H), ATTN_NORM: (H,), MLP_NORM: (H,), FINAL_NORM: (H,),
            WQ: (qd, H), WK: (kd, H), WV: (kd, H), WO: (H, qd), WGATE: (s.inter, H), WUP: (s.inter, H),
            WDOWN: (H, s.inter)}[kind]


def synth_tensor(s: ModelShape, seed: int, kind: int, layer: int, round_bf16: bool = True) -> np.ndarray:
    amp, off = C.c_float(), C.c_float()
    cfg = s.c()
    lib().orc_synth_params(C.byref(cfg), kind, C.byref(amp), C.byref(off))
    shp = tensor_shape(s, kind)
    out = np.empty(int(np.prod(shp)), dtype=np.float32)
    lib().orc_synth_tensor(_fp(out), out.size, seed, kind, layer if kind in PER_LAYER else 0, amp.value,
                           off.value, int(round_bf16))
    return out.reshape(shp)


class Oracle:
    """One model instance with its own KV histories."""

    def __init__(self, shape: ModelShape, act_round: bool = False):
        self.shape = shape
        self._cfg = shape.c()
        self._m = lib().orc_create(C.byref(self._cfg), int(act_round))
        self.act_round = act_round

    def close(self):
        if self._m:
            lib().orc_destroy(self._m)
            self._m = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    def tensor(self, kind: int, layer: int = 0) -> np.ndarray:
        """Writable view of the model's storage for one tensor."""
        shp = tensor_shape(self.shape, kind)
        p = lib().orc_tensor(self._m, kind, layer)
        return np.ctypeslib.as_array(p, shape=(int(np.prod(shp)),)).reshape(shp)

    def fill_synthetic(self, seed: int, round_bf16: bool = True):
        lib().orc_fill_synthetic(self._m, seed, int(round_bf16))

    def tensors(self):
        for k in GLOBAL:
            yield k, 0, self.tensor(k, 0)
        for l in range(self.shape.layers):
            for k in PER_LAYER:
                yield k, l, self.tensor(k, l)

    def seq_new(self) -> int:
        return lib().orc_seq_new(self._m)

    def seq_free(self, sid: int):
        lib().orc_seq_free(self._m, sid)

    def seq_fill_synthetic(self, sid: int, length: int, seed: int):
        lib().orc_seq_fill_synthetic(self._m, sid, length, seed)

    def forward(self, tokens, seq_ids, positions, want_logits=None, want_hidden=False):
        tokens = np.ascontiguousarray(tokens, dtype=np.int32)
        seq_ids = np.ascontiguousarray(seq_ids, dtype=np.int32)
        positions = np.ascontiguousarray(positions, dtype=np.int32)
        T = tokens.size
        if want_logits is None:
            want_logits = np.ones(T, dtype=bool)
        want_logits = np.asarray(want_logits, dtype=bool)
        rows = np.full(T, -1, dtype=np.int32)
        rows[want_logits] = np.arange(int(want_logits.sum()), dtype=np.int32)
        logits = np.zeros((int(want_logits.sum()), self.shape.vocab), dtype=np.float32)
        hidden = np.zeros((T, self.shape.hidden), dtype=np.float32) if want_hidden else None
        rc = lib().orc_forward(self._m, T, _ip(tokens), _ip(seq_ids), _ip(positions), _ip(rows),
                               _fp(logits) if logits.size else None, _fp(hidden) if want_hidden else None)
        if rc != 0:
            raise ValueError("orc_forward: bad argument")
        return (logits, hidden) if want_hidden else logits

    def generate(self, prompts, max_new, temperature=0.0, top_p=1.0, seeds=None, eos_id=-1, n_logit_steps=0):
        """prompts: list of int lists.  Returns dict(tokens=list of lists, margins, logits)."""
        B = len(prompts)
        lens = np.array([len(p) for p in prompts], dtype=np.int32)
        stride = int(lens.max())
        buf = np.zeros((B, stride), dtype=np.int32)
        for b, p in enumerate(prompts):
            buf[b, :len(p)] = p
        out = np.zeros((B, max_new), dtype=np.int32)
        out_lens = np.zeros(B, dtype=np.int32)
        margins = np.zeros((B, max_new), dtype=np.float32)
        logits = np.zeros((B, n_logit_steps, self.shape.vocab), dtype=np.float32) if n_logit_steps else None
        sd = (C.c_uint64 * B)(*([0] * B if seeds is None else [int(x) for x in seeds]))
        rc = lib().orc_generate(self._m, B, _ip(buf), _ip(lens), stride, max_new, temperature, top_p, sd, eos_id,
                                _ip(out), _ip(out_lens), _fp(margins), _fp(logits) if n_logit_steps else None,
                                n_logit_steps)
        if rc != 0:
            raise ValueError("orc_generate: bad argument")
        return {"tokens": [out[b, :out_lens[b]].tolist() for b in range(B)], "margins": margins,
                "logits": logits, "lens": out_lens}


# ---- stand-alone primitives (kernel-level parity checks) -------------------------------------

def rmsnorm(x, g, eps, round_out=False):
    x = np.ascontiguousarray(x, np.float32); g = np.ascontiguousarray(g, np.float32)
    y = np.empty_like(x)
    lib().orc_rmsnorm(_fp(x), _fp(g), _fp(y), x.shape[0], x.shape[1], eps, int(round_out))
    return y


def linear(x, w):
    x = np.ascontiguousarray(x, np.float32); w = np.ascontiguousarray(w, np.float32)
    y = np.empty((x.shape[0], w.shape[0]), np.float32)
    lib().orc_linear(_fp(x), _fp(w), _fp(y), x.shape[0], w.shape[0], x.shape[1])
    return y


def rope_table(max_ctx, d, theta):
    cs = np.empty((max_ctx, d // 2), np.float32); sn = np.empty_like(cs)
    lib().orc_rope_table(_fp(cs), _fp(sn), max_ctx, d, theta)
    return cs, sn


def rope(x, positions, n_heads, d, cs, sn):
    y = np.ascontiguousarray(x, np.float32).copy()
    pos = np.ascontiguousarray(positions, np.int32)
    lib().orc_rope(_fp(y), _ip(pos), y.shape[0], n_heads, d, _fp(cs), _fp(sn))
    return y


def attention_one(q, K, V, n_heads, n_kv, d):
    """q [n_heads*d]; K,V [ctx, n_kv*d] -> [n_heads*d]."""
    q = np.ascontiguousarray(q, np.float32); K = np.ascontiguousarray(K, np.float32)
    V = np.ascontiguousarray(V, np.float32)
    out = np.empty(n_heads * d, np.float32)
    lib().orc_attention_one(_fp(q), _fp(K), _fp(V), _fp(out), K.shape[0], n_heads, n_kv, d, n_kv * d)
    return out


def silu_mul(g, u, round_out=False):
    g = np.ascontiguousarray(g, np.float32); u = np.ascontiguousarray(u, np.float32)
    y = np.empty_like(g)
    lib().orc_silu_mul(_fp(g.reshape(-1)), _fp(u.reshape(-1)), _fp(y.reshape(-1)), g.size, int(round_out))
    return y


def argmax(logits):
    logits = np.ascontiguousarray(logits, np.float32)
    m = C.c_float()
    i = lib().orc_argmax(_fp(logits), logits.size, C.byref(m))
    return i, m.value


def sample(logits, temperature, top_p, seed, step):
    logits = np.ascontiguousarray(logits, np.float32)
    tau, mass = C.c_float(), C.c_float()
    tok = lib().orc_sample(_fp(logits), logits.size, temperature, top_p, seed, step, C.byref(tau), C.byref(mass))
    return tok, tau.value, mass.value


def uniform(seed, key, idx):
    return lib().orc_uniform(seed, key, idx)


def num_threads():
    return lib().orc_num_threads()


def set_num_threads(n: int):
    lib().orc_set_num_threads(int(n))
