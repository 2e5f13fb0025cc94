"""CPU-side checks of the boundary: the library loads, exports every symbol include/dis_engine.h
declares, the ctypes structs match the header's layout, and compute calls fail loudly without a GPU
(there is no CPU fallback)."""
import ctypes as C
import os
import re
import subprocess

import pytest

from distributed_inference_server_b200 import binding as B
from distributed_inference_server_b200 import Engine, ModelShape, WorkerConfig, WorkerError

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
HEADER = os.path.join(ROOT, "include", "dis_engine.h")


def declared_symbols():
    src = open(HEADER).read()
    return sorted(set(re.findall(r"DIS_API\s+[\w\s\*]+?\b(dis_\w+)\s*\(", src)))


def test_library_exports_every_declared_symbol():
    lib = B.load()
    names = declared_symbols()
    assert len(names) >= 25
    for n in names:
        assert hasattr(lib, n), n
        assert n in B.PROTOTYPES, f"binding.py lacks a prototype for {n}"
    assert lib.dis_abi_version() == 2


def test_struct_layout_matches_header(tmp_path):
    probe = tmp_path / "probe.c"
    probe.write_text('#include <stdio.h>\n#include "dis_engine.h"\nint main(){printf("%zu %zu %zu %zu %zu %zu %zu %zu %zu\\n",'
                     'sizeof(dis_model_config),sizeof(dis_engine_config),sizeof(dis_sampling_params),'
                     'sizeof(dis_worker_status),sizeof(dis_batch_result),offsetof(dis_engine_config,nccl_id),'
                     'sizeof(dis_serving_event),offsetof(dis_serving_event,kind),offsetof(dis_serving_event,logprob));return 0;}\n')
    exe = tmp_path / "probe"
    subprocess.run(["gcc", "-I", os.path.join(ROOT, "include"), str(probe), "-o", str(exe)], check=True)
    sizes = [int(x) for x in subprocess.run([str(exe)], capture_output=True, text=True, check=True).stdout.split()]
    assert sizes == [C.sizeof(B.ModelConfig), C.sizeof(B.EngineConfig), C.sizeof(B.SamplingParams),
                     C.sizeof(B.WorkerStatus), C.sizeof(B.BatchResult), B.EngineConfig.nccl_id.offset,
                     C.sizeof(B.ServingEvent), B.ServingEvent.kind.offset, B.ServingEvent.logprob.offset]
    assert B.ServingEvent.kind.offset == 16           # RequestID = Uuid, 16 bytes (crates/core/src/types.rs:7)


def test_header_is_plain_c():
    r = subprocess.run(["gcc", "-std=c99", "-pedantic", "-Wall", "-Werror", "-fsyntax-only", "-x", "c", HEADER],
                       capture_output=True, text=True)
    assert r.returncode == 0, r.stderr


def test_no_cpu_fallback():
    import torch
    if torch.cuda.is_available():
        pytest.skip("GPU present")
    with pytest.raises(WorkerError) as ei:
        Engine(WorkerConfig(ModelShape(1024, 256, 2, 4, 1, 128, 512, 128)))
    assert "no CUDA device" in str(ei.value)
    lib = B.load()
    assert lib.dis_k_gemv(None, None, None, 1, 8, 8, None) == B.ERR_INFERENCE_FAILED
    assert b"no CPU fallback" in lib.dis_last_error()


def test_product_never_imports_the_oracle():
    pkg = os.path.join(ROOT, "distributed_inference_server_b200")
    for dirpath, _, files in os.walk(pkg):
        for f in files:
            if f.endswith((".py", ".cu", ".cpp", ".h", ".hpp", ".cuh")):
                text = open(os.path.join(dirpath, f), errors="ignore").read()
                assert "liboracle" not in text and "from oracle" not in text and "import oracle" not in text, f


def test_build_recipes_list_the_same_sources():
    """The nvcc source list exists three times: build.py (what is built and tested here), the Rust build.rs a maintainer
    of the reference would add, and its copy in INTEGRATION.md.  They must not drift apart."""
    from distributed_inference_server_b200 import build as b
    want = list(b.SOURCES)
    for path in ("integration/rust/inference-sys/build.rs", "INTEGRATION.md"):
        text = open(os.path.join(ROOT, path)).read()
        block = text[text.index("let files = ["):]
        block = block[:block.index("];")]
        assert re.findall(r'"([\w.]+\.(?:cu|cpp))"', block) == want, path


def test_library_was_built_from_the_current_sources():
    """Binary provenance: libdis_engine.so is git-ignored yet travels to the GPU box, so it must say what it was built
    from -- dis_build_id() is the hash of every source, header and compiler flag, and has to match the tree."""
    from distributed_inference_server_b200 import build as b
    assert B.load().dis_build_id().decode() == b.source_id()


def test_rust_sys_shim_mirrors_the_header_field_for_field():
    """integration/rust cannot be compiled here (no cargo), so the one thing that can drift silently -- a #[repr(C)]
    struct whose fields no longer match the header's, or an extern fn the header does not declare -- is checked as text:
    every struct of the shim has the header's field names in the header's order, and every `pub fn dis_*` exists in
    the header with the same number of parameters."""
    import re
    hdr = open(os.path.join(ROOT, "include", "dis_engine.h")).read()
    hdr = re.sub(r"/\*.*?\*/", " ", hdr, flags=re.S)
    rs = open(os.path.join(ROOT, "integration", "rust", "inference-sys", "src", "lib.rs")).read()
    rs = re.sub(r"//[^\n]*", " ", rs)
    h_structs = {}
    for body, name in re.findall(r"typedef struct \w*\s*\{(.*?)\}\s*(\w+)\s*;", hdr, flags=re.S):
        fields = []
        for decl in body.split(";"):
            decl = decl.strip()
            if not decl:
                continue
            first, *rest = decl.split(",")
            names = [re.sub(r"\[.*?\]", "", first).split()[-1].lstrip("*")] + [re.sub(r"\[.*?\]", "", r).strip().lstrip("*") for r in rest]
            fields += names
        h_structs[name] = fields
    aliases = {"dis_worker_status": "dis_engine_status"}
    n = 0
    for name, body in re.findall(r"pub struct (\w+)\s*\{(.*?)\}", rs, flags=re.S):
        fields = re.findall(r"pub (\w+)\s*:", body)
        want = h_structs.get(name) or h_structs.get(aliases.get(name, ""))
        assert want is not None, f"{name}: no such struct in the header"
        assert fields == want, (name, fields, want)
        n += 1
    assert n >= 7
    h_fns = {m[0]: m[1] for m in re.findall(r"DIS_API [^;(]*?(dis_\w+)\s*\(([^;]*?)\)\s*;", hdr, flags=re.S)}
    def count(params):
        params = params.strip()
        return 0 if params in ("", "void") else len(params.split(","))
    for fn, params in re.findall(r"pub fn (dis_\w+)\s*\((.*?)\)\s*(?:->|;)", rs, flags=re.S):
        assert fn in h_fns, f"{fn} is not declared in include/dis_engine.h"
        assert count(params) == count(h_fns[fn]), (fn, params, h_fns[fn])
