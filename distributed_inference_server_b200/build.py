"""Build libdis_engine.so in-tree with nvcc for sm_100a (no JIT cache: the .so travels with the repo)."""
from __future__ import annotations

import hashlib
import os
import subprocess
import sys
from concurrent.futures import ThreadPoolExecutor

HERE = os.path.dirname(os.path.abspath(__file__))
CSRC = os.path.join(HERE, "csrc")
OBJ = os.path.join(CSRC, "_build")
LIB = os.path.join(HERE, "libdis_engine.so")
SOURCES = ["k_elementwise.cu", "k_gemv.cu", "k_attn_paged.cu", "k_attn_prefill_tc.cu", "k_gemm_tcgen05.cu", "k_gemm_persist.cu", "k_tp.cu", "k_tp_sample.cu", "engine.cu", "api.cpp",
           "tp.cpp", "worker.cpp", "serving.cpp", "prefix.cpp"]
NVCC = os.environ.get("NVCC", "/usr/local/cuda/bin/nvcc")
FLAGS = ["-gencode", "arch=compute_100a,code=sm_100a", "-lineinfo", "-O3", "-std=c++17", "-Xcompiler", "-fPIC",
         "-Xcompiler", "-fvisibility=hidden", "-Xcompiler", "-Wall", "-Xcompiler", "-Wno-unused-function",
         "--expt-relaxed-constexpr"]


def source_id() -> str:
    """sha256 over every source file, header and compiler flag that goes into libdis_engine.so.  The library carries it
    (dis_build_id()): a test compares the two, so a stale binary cannot pass for the current sources."""
    h = hashlib.sha256()
    h.update(" ".join(FLAGS).encode())
    files = [os.path.join(CSRC, f) for f in sorted(os.listdir(CSRC)) if f.endswith((".cu", ".cpp", ".h", ".hpp", ".cuh"))]
    inc = os.path.join(os.path.dirname(HERE), "include")
    files += [os.path.join(inc, f) for f in sorted(os.listdir(inc)) if f.endswith(".h")]
    for f in files:
        h.update(os.path.basename(f).encode())
        h.update(open(f, "rb").read())
    return h.hexdigest()[:32]


def _deps_mtime() -> float:
    t = 0.0
    for root in (CSRC, os.path.join(os.path.dirname(HERE), "include")):
        for f in os.listdir(root):
            if f.endswith((".h", ".hpp", ".cuh")):
                t = max(t, os.path.getmtime(os.path.join(root, f)))
    return t


def _compile(src: str, verbose: bool) -> str:
    obj = os.path.join(OBJ, src + ".o")
    path = os.path.join(CSRC, src)
    if os.path.exists(obj) and os.path.getmtime(obj) >= max(os.path.getmtime(path), _deps_mtime()):
        return obj
    cmd = [NVCC, *FLAGS, "-x", "cu", "-c", path, "-o", obj]
    if verbose:
        cmd.insert(1, "-Xptxas=-v")
    r = subprocess.run(cmd, capture_output=True, text=True)
    if r.returncode != 0:
        raise RuntimeError(f"nvcc failed for {src}:\n{r.stdout}\n{r.stderr}")
    if verbose or r.stderr.strip():
        sys.stderr.write(r.stderr)
    return obj


def build(force: bool = False, verbose: bool = False) -> str:
    os.makedirs(OBJ, exist_ok=True)
    if force:
        for f in os.listdir(OBJ):
            os.remove(os.path.join(OBJ, f))
    with ThreadPoolExecutor(max_workers=8) as ex:
        objs = list(ex.map(lambda s: _compile(s, verbose), SOURCES))
    # the build id is its own tiny object, regenerated whenever the sources' hash changes
    sid = source_id()
    id_src, id_obj = os.path.join(OBJ, "build_id.cpp"), os.path.join(OBJ, "build_id.cpp.o")
    text = f'extern "C" __attribute__((visibility("default"))) const char *dis_build_id(void) {{ return "{sid}"; }}\n'
    if not os.path.exists(id_src) or open(id_src).read() != text or not os.path.exists(id_obj):
        open(id_src, "w").write(text)
        r = subprocess.run(["g++", "-O2", "-fPIC", "-c", id_src, "-o", id_obj], capture_output=True, text=True)
        if r.returncode != 0:
            raise RuntimeError(f"build id failed:\n{r.stderr}")
    objs.append(id_obj)
    if force or not os.path.exists(LIB) or any(os.path.getmtime(o) > os.path.getmtime(LIB) for o in objs):
        cmd = [NVCC, "-shared", "-o", LIB, *objs, "-ldl", "-gencode", "arch=compute_100a,code=sm_100a"]
        r = subprocess.run(cmd, capture_output=True, text=True)
        if r.returncode != 0:
            raise RuntimeError(f"link failed:\n{r.stdout}\n{r.stderr}")
    return LIB


if __name__ == "__main__":
    print(build(force="--force" in sys.argv, verbose="-v" in sys.argv))
