#!/usr/bin/env python
"""bench.py -- decode throughput of the B200-native decoder engine on BASELINE.json's metric.

    python bench.py --gpus N --steps K --warmup W            (N>1: launched under torch.distributed.run)
    python bench.py --impl reference --gpus N --steps K --warmup W

A "step" is one decode step of the whole batch (one new token per sequence).  Workloads:
  N = 1 : BASELINE configs[1] -- Llama-3-8B bf16, batch 32, 512-token prompts, one B200
  N > 1 : BASELINE configs[2] -- Llama-3-8B bf16, tensor parallel TP = N over NVLink, batch 64
Weights are random-init (counter-hash, seed 0) and prompts uniform random tokens: there is no network
for checkpoints or datasets.  `value` is timed with CUDA events on the engine's stream around the K
steps; `e2e` is the same loop on the host clock through the C-ABI decode call, whose host buffers
(slot list in, token ids out) cross PCIe inside the timed region every step.

Other lines (diagnostics and the remaining BASELINE configs, never the default):
  --batch 1 / --model 70b            configs[1] batch 1, configs[3] Llama-3-70B TP = 8 batch 32
  --workload serving                 configs[4]: 256 requests, prompts 128-8192, through dis_serving_*
  --workload prefix                  TTFT with prefix KV reuse
  --shard-of N                       one rank's shard of a TP = N run on one GPU, no exchange
  --tp-exchange / --nccl-allreduce   A/B of the exchange kernels (k_tp.cu) against each other and NCCL
Extra keys: roofline (dominant kernel, from an eager step timed per kernel class), step_roofline (whole step),
graph_only_ms_last_step (the last step's CUDA graph replayed back to back), ttft_* (prefill of the whole batch and of
one request, measured AFTER the decode run), kernel_class_ms_per_step, clocks; N > 1 adds parity (the TP engines'
tokens and logits against a TP=1 engine on rank 0, checked before anything is timed) and tp1_same_batch /
scaling_vs_tp1_same_batch (the same batch on one GPU: the like-for-like denominator of the scaling curve).
Order of the run: [TP parity] -> one prefill of the batch -> `--settle-steps` untimed decode steps (power state) ->
W warm-up steps -> K timed steps -> graph replay, per-class profile, TTFT runs -> CPU legs.

The reference has no implementation of this path (crates/inference/src/worker.rs:1), so the
`--impl reference` arm and `cpu_baseline` time the CPU oracle port (oracle/llama_oracle.c) on the
host cores: the same workload at full depth (every reported step is run) when the host has the 32 GB + KV the fp32
model needs, else 2 of 32 layers scaled (the line then says so).  `cpu_config0` is BASELINE configs[0] (TinyLlama
shape, full depth) on the same oracle.  All CPU legs run on rank 0 outside the process group.
"""
from __future__ import annotations

import argparse
import json
import os
import statistics
import subprocess
import sys
import threading
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

import numpy as np  # noqa: E402


def peaks():
    p = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(p):
        d = json.load(open(p))
        return float(d["hbm_gbs"]), "measured (MEASURED_PEAKS.json hbm_gbs)"
    return 6650.0, "fallback (B200_PROFILING.md 6.65 TB/s)"


def tensor_peak():
    """Dense bf16 TFLOP/s a kernel inside a long step can sustain (the prefill is hundreds of milliseconds of GEMMs)."""
    p = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(p):
        d = json.load(open(p))
        if "bf16_tflops_sustained" in d:
            return float(d["bf16_tflops_sustained"]), "measured (MEASURED_PEAKS.json bf16_tflops_sustained)"
    return 1400.0, "fallback (B200_PROFILING.md: about 1.4 PFLOP/s sustained cuBLAS bf16)"


# --------------------------------------------------------------------------------------- byte model
def model_bytes(shape, tp):
    """Algorithmic bytes per decode step per GPU: every streamed weight once + KV."""
    H, d, L = shape.hidden, shape.head_dim, shape.layers
    qd, kd = shape.n_heads * d, shape.n_kv_heads * d
    per_layer = H * qd + 2 * H * kd + qd * H + 3 * H * shape.inter + 2 * H
    p_stream = L * per_layer + shape.vocab * H + H
    kv_per_token = L * 2 * kd * 2
    return 2 * p_stream / tp, kv_per_token / tp


# --------------------------------------------------------------------------------------- clocks
class ClockSampler:
    """Polls NVML (SM clock, power, clock-event reasons) every 20 ms on a thread; only samples taken
    between mark_start() and mark_end() -- the timed region -- are summarised."""
    REASONS = {0x8: "hw_slowdown", 0x40: "hw_thermal_slowdown", 0x20: "sw_thermal_slowdown", 0x4: "sw_power_cap"}

    def __init__(self, index):
        self.index, self.rows, self.stop_flag, self.t0, self.t1, self.thread, self.max = index, [], False, None, None, None, None

    def start(self):
        try:
            import pynvml
            pynvml.nvmlInit()
            h = pynvml.nvmlDeviceGetHandleByIndex(self.index)
            self.max = pynvml.nvmlDeviceGetMaxClockInfo(h, pynvml.NVML_CLOCK_SM)
        except Exception:
            return

        def loop():
            while not self.stop_flag:
                try:
                    sm = pynvml.nvmlDeviceGetClockInfo(h, pynvml.NVML_CLOCK_SM)
                    rs = pynvml.nvmlDeviceGetCurrentClocksEventReasons(h)
                    pw = pynvml.nvmlDeviceGetPowerUsage(h) / 1000.0
                    self.rows.append((time.perf_counter(), sm, rs, pw))
                except Exception:
                    pass
                time.sleep(0.02)
        self.thread = threading.Thread(target=loop, daemon=True)
        self.thread.start()

    def mark_start(self):
        self.t0 = time.perf_counter()

    def mark_end(self):
        self.t1 = time.perf_counter()

    def stop(self):
        self.stop_flag = True
        if self.thread:
            self.thread.join(timeout=1.0)
        inside = [r for r in self.rows if self.t0 is not None and self.t0 <= r[0] <= (self.t1 or 1e30)]
        use = inside or self.rows[-3:]
        reasons = set()
        for r in use:
            for bit, name in self.REASONS.items():
                if r[2] & bit:
                    reasons.add(name)
        return {"sm_mhz": statistics.median([r[1] for r in use]) if use else None, "sm_max_mhz": self.max,
                "reasons": sorted(reasons), "power_w_max": max([r[3] for r in use]) if use else None,
                "samples": len(inside)}


# --------------------------------------------------------------------------------------- CPU arm
def host_cores():
    try:
        return len(os.sched_getaffinity(0))
    except Exception:
        return os.cpu_count() or 1


def mem_available_gb():
    try:
        for line in open("/proc/meminfo"):
            if line.startswith("MemAvailable:"):
                return int(line.split()[1]) / 1e6
    except Exception:
        pass
    return 0.0


class CpuDecoder:
    """The fp32 CPU oracle (oracle/llama_oracle.c) set up for decode steps of the GPU arm's workload: Llama-3-8B
    width, batch `batch`, synthetic `ctx`-token KV histories.  FULL depth (32 layers, 32 GB of fp32 weights) when
    the host has the memory for it -- every reported step is then really run, nothing is extrapolated; otherwise
    2 of the 32 layers and the figure is scaled, which `sample` then says in words."""

    def __init__(self, batch, ctx, n_steps, full=None):
        from oracle import oracle as O
        self.O = O
        O.set_num_threads(host_cores())   # torch.distributed.run exports OMP_NUM_THREADS=1: never inherit that
        s8 = O.LLAMA3_8B
        kvd = s8.n_kv_heads * s8.head_dim
        need_gb = 4 * (s8.layers * (2 * s8.hidden * s8.hidden + 2 * s8.hidden * kvd + 3 * s8.hidden * s8.inter)
                       + 2 * s8.vocab * s8.hidden) / 1e9 + batch * (ctx + n_steps + 8) * kvd * 2 * 4 * s8.layers / 1e9
        self.full = (mem_available_gb() > need_gb + 8.0) if full is None else full
        self.layers = s8.layers if self.full else 2
        self.batch, self.ctx = batch, ctx
        shape = O.ModelShape(s8.vocab, s8.hidden, self.layers, s8.n_heads, s8.n_kv_heads, s8.head_dim,
                             s8.inter, ctx + n_steps + 8, s8.rope_theta, s8.rms_eps)
        t0 = time.perf_counter()
        self.m = O.Oracle(shape, act_round=False)
        self.m.fill_synthetic(0, True)
        self.sids = [self.m.seq_new() for _ in range(batch)]
        for s in self.sids:
            self.m.seq_fill_synthetic(s, ctx, 7)
        self.setup_s = time.perf_counter() - t0
        self.rng = np.random.default_rng(1)
        self.it = 0
        self.cores = O.num_threads()
        self.head_s = None

    def step(self):
        """One decode step of the whole batch; returns seconds (scaled to 32 layers when running a layer sample)."""
        toks = self.rng.integers(0, self.m.shape.vocab, self.batch)
        pos = np.full(self.batch, self.ctx + self.it, np.int32)
        self.it += 1
        if self.full:
            t0 = time.perf_counter()
            self.m.forward(toks, self.sids, pos, want_logits=np.ones(self.batch, bool))
            return time.perf_counter() - t0
        t0 = time.perf_counter()
        self.m.forward(toks, self.sids, pos, want_logits=np.zeros(self.batch, bool))   # layers only
        t1 = time.perf_counter()
        if self.head_s is None:  # layers + head at the same positions (K/V rows are simply rewritten)
            self.m.forward(toks, self.sids, pos, want_logits=np.ones(self.batch, bool))
            self.head_s = max(time.perf_counter() - t1 - (t1 - t0), 0.0)
        return (t1 - t0) * (32 / self.layers) + self.head_s

    def describe(self, n_timed):
        depth = ("all 32 layers + lm_head, every step really run" if self.full else
                 f"{self.layers} of 32 layers timed and SCALED to 32 (+ lm_head timed once): host memory too small for the 32 GB fp32 model")
        return (f"oracle/llama_oracle.c (fp32, OpenMP, {self.cores} threads): {n_timed} decode step(s) of batch {self.batch} at ctx "
                f"{self.ctx}, Llama-3-8B shape, {depth}; the reference itself has no forward pass (crates/inference/src/worker.rs:1)")

    def close(self):
        self.m.close()


def cpu_baseline_leg(batch, ctx):
    """cpu_baseline of the GPU arm's line: a bounded sample (about 20-30 s) of the same workload on the host cores."""
    dec = CpuDecoder(batch, ctx, n_steps=4)
    dec.step()                                   # warm-up
    ts = [dec.step() for _ in range(2)]
    step_s = statistics.median(ts)
    out = {"value": batch / step_s, "unit": "tokens/s", "cores": dec.cores, "kind": "port", "sample": dec.describe(len(ts)),
           "ms_per_step": step_s * 1e3, "extrapolated": not dec.full}
    dec.close()
    return out


def cpu_config0_leg(new_tokens=32):
    """BASELINE configs[0]: TinyLlama-1.1B shape, random-init fp32, one request, 128-token prompt, greedy, on the CPU
    oracle at FULL depth (22 layers).  Bounded: `new_tokens` decode steps instead of 128 (steady-state decode rate)."""
    from oracle import oracle as O
    O.set_num_threads(host_cores())
    shape = O.ModelShape(*[getattr(O.TINYLLAMA, f) for f in ("vocab", "hidden", "layers", "n_heads", "n_kv_heads", "head_dim", "inter")],
                         256, O.TINYLLAMA.rope_theta, O.TINYLLAMA.rms_eps)
    m = O.Oracle(shape, act_round=False)
    m.fill_synthetic(0, True)
    prompt = np.random.default_rng(1).integers(0, shape.vocab, 128)
    sid = m.seq_new()
    t0 = time.perf_counter()
    lg = m.forward(prompt, [sid] * 128, np.arange(128), want_logits=np.arange(128) == 127)
    t1 = time.perf_counter()
    tok = int(np.argmax(lg[0]))
    for i in range(new_tokens):
        lg = m.forward([tok], [sid], [128 + i])
        tok = int(np.argmax(lg[0]))
    t2 = time.perf_counter()
    m.close()
    return {"workload": "BASELINE configs[0]: TinyLlama-1.1B-shape Llama decoder (22 layers), random-init fp32, one request, greedy, "
                        f"128-token prompt -> {new_tokens} new tokens (bounded from 128), CPU oracle port, full depth, nothing extrapolated",
            "prefill_s": t1 - t0, "decode_tokens_per_s": new_tokens / (t2 - t1), "cores": O.num_threads(), "kind": "port",
            "parity": "the GPU engine is checked against this oracle at this shape by tests/test_baseline_configs_gpu.py"}


def run_reference(args):
    """`--impl reference`: the reference's CPU implementation of the path on the host cores.  The reference has none
    (crates/inference/src/worker.rs:1 is a comment), so this is the CPU oracle port, on the GPU arm's config.  Every one
    of the W + K steps is run; a step is one decode step of the whole batch."""
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return 0
    batch = args.batch or (32 if args.gpus == 1 else 64)
    t0 = time.perf_counter()
    dec = CpuDecoder(batch, args.prompt_len, n_steps=args.steps + args.warmup)
    for _ in range(args.warmup):
        dec.step()
    ts = [dec.step() for _ in range(args.steps)]
    step_s = sum(ts) / max(len(ts), 1)
    base = {"value": batch / step_s, "unit": "tokens/s", "cores": dec.cores, "kind": "port", "sample": dec.describe(len(ts)),
            "extrapolated": not dec.full, "setup_s": dec.setup_s}
    cfg = workload_config(args.gpus, batch, args.prompt_len)
    cfg["cpu_arm"] = f"run on {dec.cores} host threads; the tensor-parallel exchange does not exist on the CPU"
    if not dec.full:
        cfg["workload"] += " [CPU arm: 2 of 32 layers timed per step and scaled to 32; host memory too small for full depth]"
    line = {"impl": "reference", "metric": "decode tokens/s Llama-3-8B", "value": base["value"], "unit": "tokens/s",
            "n_gpus": args.gpus, "steps": len(ts), "warmup": args.warmup, "ms_per_step": step_s * 1e3,
            "higher_is_better": True, "scaling": "strong" if args.gpus > 1 else "weak", "vs_baseline": None, "dtype": "f32",
            "data": "synthetic", "config": cfg, "cpu_baseline": base,
            "e2e": {"value": base["value"], "unit": "tokens/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
            "gpu_launches": 0, "wall_s": time.perf_counter() - t0}
    dec.close()
    print(json.dumps(line))
    return 0


def workload_config(n_gpus, batch, prompt_len, model="8b"):
    if model == "70b":
        w = f"BASELINE configs[3]: Llama-3-70B bf16, tensor parallel TP={n_gpus}, batch {batch} decode, {prompt_len}-token prompts, paged KV"
        return {"workload": w, "batch": batch, "prompt_len": prompt_len, "parallelism": f"tp{n_gpus}",
                "tp_exchange": "fused peer-memory all-reduce (k_tp.cu)", "l2": "working set far larger than L2; no explicit flush"}
    if n_gpus == 1:
        w = f"BASELINE configs[1]: Llama-3-8B bf16, 1 B200, batch {batch} decode, {prompt_len}-token prompts"
    else:
        w = f"BASELINE configs[2]: Llama-3-8B bf16, tensor parallel TP={n_gpus} over NVLink, batch {batch} decode, {prompt_len}-token prompts"
    return {"workload": w, "batch": batch, "prompt_len": prompt_len, "parallelism": f"tp{n_gpus}",
            "tp_exchange": None if n_gpus == 1 else "fused peer-memory all-reduce (k_tp.cu); NCCL with --nccl-allreduce",
            "l2": "per-step working set (15 GB of weights + KV) is far larger than the 126 MB L2; no explicit flush"}


def roofline_traffic(kernel_key):
    """dram__bytes_read.sum + dram__bytes_write.sum per launch of the dominant kernel, from the committed summary of the
    ncu --set full capture (profiles/roofline_traffic.json, written by scripts/summarize_profile.py) -- a profiler is
    never run inside the bench.  None when there is no capture for this kernel/shape."""
    p = os.path.join(ROOT, "profiles", "roofline_traffic.json")
    try:
        d = json.load(open(p)).get(kernel_key)
        if d:
            return int(d["dram_bytes_read"]) + int(d["dram_bytes_write"]), d.get("source")
    except Exception:
        pass
    return None, None


# --------------------------------------------------------------------------------------- GPU arm
def tp_parity_check(eng, make_engine, prompts, P, rank, world, dist, torch, gloo):
    """Before anything is timed: the tensor-parallel engines' first three greedy tokens of the whole batch (prefill + two
    decode steps) and the logits of the first decode step, against a TP=1 engine with the same weights and prompts built
    on rank 0 (torch.distributed is the plumbing).  Logits are compared for the sequences whose history is still the same
    at that step.  Returns (parity dict on rank 0 / None, the TP=1 engine + its slots on rank 0 for the same-batch timing)."""
    def three_steps(e):
        slots = [e.seq_add(p, P) for p in prompts]
        toks = [e.prefill(slots)[0].astype(np.int64)]
        toks.append(e.decode(slots)[0].astype(np.int64))
        logits = e.last_logits(len(slots))                   # first decode step; gathered over ranks (every rank takes part)
        toks.append(e.decode(slots)[0].astype(np.int64))
        return slots, np.stack(toks), logits

    slots, got, logits = three_steps(eng)
    for s in slots:
        eng.seq_remove(s)
    mine = torch.tensor(got, device="cuda")
    every = [torch.zeros_like(mine) for _ in range(world)]
    dist.all_gather(every, mine)
    ranks_agree = all(bool(torch.equal(every[0], t)) for t in every)
    lg_sum = torch.tensor([float(np.abs(logits.astype(np.float64)).sum())], device="cuda", dtype=torch.float64)
    lg_all = [torch.zeros_like(lg_sum) for _ in range(world)]
    dist.all_gather(lg_all, lg_sum)
    ranks_agree = ranks_agree and all(float(t) == float(lg_all[0]) for t in lg_all)
    parity, one, one_slots = None, None, None
    if rank == 0:
        def errors(a, b, rows):
            if not rows.any():
                return None, None
            d = (a[rows] - b[rows]).astype(np.float64)
            return float(np.linalg.norm(d) / np.linalg.norm(b[rows].astype(np.float64))), float(np.abs(d).max() / np.abs(b[rows]).max())

        one = make_engine(1, 0)
        one_slots, ref, ref_logits = three_steps(one)
        # the noise floor of bf16 on THIS model, measured rather than assumed: the same TP=1 engine with the prompts read in
        # 512-token chunks instead of 8192 (128-token split-K GEMM tiles instead of the persistent 256-token ones, so other
        # fp32 summation orders and bf16 roundings in the KV cache).  A random-init 32-layer decoder does not damp rounding noise: two equally valid bf16 evaluations
        # differ by a few percent of the logit scale, and the tensor-parallel engines are held to twice that floor.
        other = make_engine(1, 0, chunk=512)
        o_slots, alt, alt_logits = three_steps(other)
        other.close()
        floor_rms, floor_max = errors(alt_logits, ref_logits, alt[0] == ref[0])
        n_seq = got.shape[1]
        same_hist = got[0] == ref[0]                         # rows whose first decode step saw the same history
        rms, rel = errors(logits, ref_logits, same_hist)
        identical = int((got == ref).all(axis=0).sum())
        tol_rms = max(2e-2, 2.0 * (floor_rms or 0.0))
        tol_max = min(max(2e-2, 2.0 * (floor_max or 0.0)), 8e-2)
        parity = {"against": "a TP=1 engine with the same weights and prompts on rank 0", "steps": 3, "sequences": n_seq,
                  "sequences_with_identical_tokens": identical,
                  "sequences_with_identical_tokens_tp1_vs_tp1_other_chunking": int((alt == ref).all(axis=0).sum()),
                  "decode_step_1_logits_rms_err_over_rms_logit": rms, "decode_step_1_logits_max_err_over_max_logit": rel,
                  "rows_compared": int(same_hist.sum()),
                  "noise_floor_tp1_vs_tp1_other_chunking": {"rms": floor_rms, "max": floor_max},
                  "tolerance": {"rms": tol_rms, "max": tol_max,
                                "rule": "max(2e-2, 2 x the floor): both sides compute in bf16 with different summation orders; the "
                                        "maximum runs over 64 x 128256 logits of a random-init model that does not damp rounding noise"},
                  "all_ranks_agree_bitwise": bool(ranks_agree)}
        parity["ok"] = bool(ranks_agree and rms is not None and rms <= tol_rms and rel <= tol_max and same_hist.sum() >= n_seq // 2)
        if not parity["ok"]:
            print(json.dumps({"parity_failed": parity}), file=sys.stderr)
    ok = torch.tensor([1 if (parity is None or parity["ok"]) else 0])    # host tensor over gloo: the waiting GPUs stay idle
    dist.broadcast(ok, 0, group=gloo)
    if int(ok) == 0:
        raise SystemExit("tensor-parallel parity check failed: nothing was timed (details on rank 0's stderr)")
    return parity, one, one_slots


def run_ours(args):
    import torch
    from distributed_inference_server_b200 import Engine, InferenceParams, LLAMA3_8B, ModelShape, WorkerConfig
    from distributed_inference_server_b200 import binding as B

    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    if world != args.gpus:
        if args.gpus != 1:
            raise SystemExit(f"--gpus {args.gpus} needs torch.distributed.run with {args.gpus} processes (WORLD_SIZE={world})")
    dist, gloo = None, None
    torch.cuda.set_device(local)
    nccl_id = None
    if world > 1:
        import torch.distributed as dist
        dist.init_process_group("nccl", device_id=torch.device("cuda", local))
        gloo = dist.new_group(backend="gloo")   # host-side waits: a GPU that waits in an NCCL barrier spins a kernel
        idt = torch.zeros(128, dtype=torch.uint8, device="cuda")
        if rank == 0:
            import ctypes as C
            buf = (C.c_uint8 * 128)()
            assert B.load().dis_tp_unique_id(buf) == 0, B.load().dis_last_error()
            idt.copy_(torch.tensor(list(buf), dtype=torch.uint8))
        dist.broadcast(idt, 0)
        nccl_id = bytes(idt.cpu().tolist())

    from distributed_inference_server_b200 import LLAMA3_70B
    shape = LLAMA3_70B if args.model == "70b" else LLAMA3_8B
    if args.layers:
        shape = ModelShape(shape.vocab, shape.hidden, args.layers, shape.n_heads, shape.n_kv_heads, shape.head_dim,
                           shape.inter, shape.max_ctx, shape.rope_theta, shape.rms_eps)
    if args.shard_of > 1:  # one rank's shard of a TP=N run, without the exchange: isolates compute from communication
        n = args.shard_of
        shape = ModelShape(shape.vocab // n, shape.hidden, shape.layers, shape.n_heads // n, shape.n_kv_heads // n, shape.head_dim,
                           shape.inter // n, shape.max_ctx, shape.rope_theta, shape.rms_eps)
    batch = args.batch or (32 if (world == 1 or args.model == "70b") else 64)
    n_ttft = max(1, args.ttft_runs)
    total_steps = args.settle_steps + args.warmup + args.steps + args.profile_steps + 8
    page = 64
    pages_per_seq = (args.prompt_len + total_steps + page) // page + 1
    stream = torch.cuda.current_stream()
    if args.tp_exchange != "auto":
        os.environ["DIS_TP_EXCHANGE"] = args.tp_exchange

    def make_engine(tp, tp_rank, chunk=None):
        e = Engine(WorkerConfig(shape, device=local, tp_rank=tp_rank, tp_size=tp, max_batch=batch,
                                max_step_tokens=chunk or args.prefill_chunk, page_size=page, kv_pages=batch * pages_per_seq,
                                context_size=args.prompt_len + total_steps + 8, stream=stream.cuda_stream,
                                nccl_id=nccl_id if tp > 1 else None,
                                flags=(B.FLAG_NO_GRAPH if args.no_graph else 0) | (B.FLAG_NO_PDL if args.no_pdl else 0)))
        e.load_synthetic(0)
        e.finalize()
        return e

    eng = make_engine(world, rank)
    if world > 1 and not args.nccl_allreduce:
        # peer-memory exchange: all-gather the CUDA IPC handles of the symmetric buffers
        mine = torch.tensor(list(eng.tp_export()), dtype=torch.uint8, device="cuda")
        every = [torch.zeros_like(mine) for _ in range(world)]
        dist.all_gather(every, mine)
        eng.tp_import([bytes(t.cpu().tolist()) for t in every])

    rng = np.random.default_rng(1)
    prompts = [rng.integers(0, shape.vocab, args.prompt_len).tolist() for _ in range(batch)]
    # greedy is the headline (BASELINE's metric); --temperature > 0 times sampled decoding (A/B of the samplers)
    P = InferenceParams(max_tokens=total_steps + 4, temperature=args.temperature, top_p=args.top_p, seed=7)

    def barrier():
        if dist is not None:
            dist.barrier()
        torch.cuda.synchronize()

    def timed_decode(e, slots, steps, sync_ranks):
        """K decode steps: CUDA events on the engine's stream (value) and the host clock around the same calls (e2e)."""
        ev0, ev1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        barrier() if sync_ranks else torch.cuda.synchronize()
        t0 = time.perf_counter()
        ev0.record(stream)
        for _ in range(steps):
            e.decode(slots)
        ev1.record(stream)
        barrier() if sync_ranks else torch.cuda.synchronize()
        wall = (time.perf_counter() - t0) * 1e3
        return ev0.elapsed_time(ev1), wall

    # ---- tensor parallel: parity against a TP=1 engine BEFORE anything is timed, and that engine's own decode time at
    # the same batch (the like-for-like denominator of the scaling curve)
    parity, tp1 = None, None
    if world > 1 and not args.no_parity and args.model == "8b" and args.shard_of <= 1 and args.temperature <= 0:
        parity, one, one_slots = tp_parity_check(eng, make_engine, prompts, P, rank, world, dist, torch, gloo)
        if rank == 0:
            for _ in range(args.settle_steps + args.warmup):
                one.decode(one_slots)
            dev1, _ = timed_decode(one, one_slots, args.steps, False)
            tp1 = {"value": batch * args.steps / (dev1 / 1e3), "unit": "tokens/s", "ms_per_step": dev1 / args.steps, "batch": batch,
                   "what": "the same batch on ONE GPU (TP=1 engine on rank 0, same build, timed the same way while the other GPUs idle)"}
            one.close()
        dist.barrier(group=gloo)

    # ---- the batch is prefilled once (TTFT sample 0, which includes one-time set-up costs)
    ttft = []
    slots = [eng.seq_add(p, P) for p in prompts]
    barrier()
    t0 = time.perf_counter()
    eng.prefill(slots)
    torch.cuda.synchronize()
    ttft_first = (time.perf_counter() - t0) * 1e3
    sampler = ClockSampler(local)
    if rank == 0:
        sampler.start()
    # ---- settle: untimed decode steps that take the board from the prefill's power state to the decode's (a B200 sits at
    # its power cap after a tensor-bound prefill and needs some 100 ms of the lighter decode load to clock back up);
    # then the W warm-up steps the caller asked for, then EXACTLY K timed steps
    for _ in range(args.settle_steps + args.warmup):
        eng.decode(slots)
    launches0 = eng.status().kernel_launches
    sampler.mark_start()
    dev_ms, wall_ms = timed_decode(eng, slots, args.steps, True)
    sampler.mark_end()
    clocks = sampler.stop() if rank == 0 else None
    launches = eng.status().kernel_launches - launches0
    if dist is not None:
        t = torch.tensor([dev_ms, wall_ms], device="cuda", dtype=torch.float64)
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        dev_ms, wall_ms = float(t[0]), float(t[1])
    mean_ctx = args.prompt_len + args.settle_steps + args.warmup + (args.steps + 1) / 2.0

    # ---- the LAST step's CUDA graph (longest context of the run) replayed back to back: what the GPU needs without
    # the host between steps; compare with the last steps' times, not with the run's average
    graph_only_ms = None
    if not args.no_graph and world == 1:
        try:
            graph_only_ms = eng.replay_last_step(32)
        except Exception:
            graph_only_ms = None
    # ---- per-kernel-class device times (eager pass with CUDA events on the engine's stream)
    prof = {}
    for _ in range(args.profile_steps):
        _, _, cls = eng.profile_decode(slots)
        for k, (ms, n) in cls.items():
            a = prof.setdefault(k, [0.0, 0])
            a[0] += ms; a[1] += n
    # ---- TTFT, after the decode measurement so that the prefills' power state does not leak into it
    for s in slots:
        eng.seq_remove(s)
    for i in range(n_ttft):
        slots = [eng.seq_add(p, P) for p in prompts]
        barrier()
        t0 = time.perf_counter()
        eng.prefill(slots)
        torch.cuda.synchronize()
        ttft.append((time.perf_counter() - t0) * 1e3)
        for s in slots:
            eng.seq_remove(s)
    ttft1 = []
    for i in range(5):     # a single request on an idle engine (p50 of 5)
        s1 = eng.seq_add(prompts[i % batch], P)
        barrier()
        t0 = time.perf_counter()
        eng.prefill([s1])
        torch.cuda.synchronize()
        ttft1.append((time.perf_counter() - t0) * 1e3)
        eng.seq_remove(s1)
    st = eng.status()
    w_bytes, kv_per_tok = model_bytes(shape, world)
    step_bytes = w_bytes + batch * mean_ctx * kv_per_tok
    peak, peak_src = peaks()
    line = None
    if rank == 0:
        ms_per_step = dev_ms / args.steps
        value = batch * args.steps / (dev_ms / 1e3)
        # dominant kernel: the fused gate-up GEMM + SiLU (54 % of the streamed bytes)
        I_l, H = shape.inter // world, shape.hidden
        gu_bytes = 2 * (2 * I_l * H) + 2 * batch * H + 2 * batch * I_l
        roof = {"bound": "hbm", "kernel": "gemm_tcgen05_kernel<BN,dual=true> (gate-up projection + SiLU-gate)",
                "achieved": None, "peak": peak, "unit": "GB/s", "frac": None, "traffic": None, "peak_source": peak_src,
                "algorithmic_bytes_per_launch": gu_bytes}
        if world == 1 and args.model == "8b" and args.shard_of <= 1:
            roof["traffic"], roof["traffic_source"] = roofline_traffic(f"gateup_8b_batch{batch}")
        if prof.get("gateup", [0, 0])[1]:
            avg_ms = prof["gateup"][0] / prof["gateup"][1]
            roof["achieved"] = gu_bytes / (avg_ms * 1e-3) / 1e9
            roof["frac"] = roof["achieved"] / peak
            roof["avg_launch_ms"] = avg_ms
        step_gbs = step_bytes / (ms_per_step * 1e-3) / 1e9
        cfg = workload_config(world, batch, args.prompt_len, args.model)
        cfg["settle_steps"] = args.settle_steps
        if args.temperature > 0:
            cfg["sampling"] = {"temperature": args.temperature, "top_p": args.top_p, "tp_sampler": os.environ.get("DIS_TP_SAMPLE", "fused")}
        line = {"metric": "decode tokens/s Llama-3-" + args.model.upper(), "value": value, "unit": "tokens/s", "n_gpus": world,
                "steps": args.steps, "warmup": args.warmup, "ms_per_step": ms_per_step, "higher_is_better": True,
                "scaling": "strong" if world > 1 else "weak", "vs_baseline": None, "dtype": "bf16", "data": "synthetic",
                "config": cfg,
                "roofline": roof,
                "step_roofline": {"bound": "hbm", "algorithmic_bytes_per_step_per_gpu": step_bytes, "achieved": step_gbs,
                                  "peak": peak, "unit": "GB/s", "frac": step_gbs / peak, "frac_of_nominal_8TBs": step_gbs / 8000.0,
                                  "mean_ctx": mean_ctx},
                "e2e": {"value": batch * args.steps / (wall_ms / 1e3), "unit": "tokens/s",
                        "h2d_bytes_per_step": int(60 * batch), "d2h_bytes_per_step": 4 * (batch + batch),
                        "how": "host clock around K dis_engine_decode calls: slot list in, token ids + log-probabilities out, per-step metadata H2D and result D2H inside"},
                "ttft_ms_p50": statistics.median(ttft), "ttft_ms_all": ttft, "ttft_ms_first_call": ttft_first,
                "ttft_note": f"ttft_ms_p50: prefill of all {batch} x {args.prompt_len}-token prompts in one call (every request waits for it); "
                             f"ttft_single_ms_p50: one {args.prompt_len}-token request on an idle engine; measured after the decode run",
                "ttft_single_ms_p50": statistics.median(ttft1),
                "gpu_launches": int(launches), "clocks": clocks, "graph_only_ms_last_step": graph_only_ms,
                "kernel_class_ms_per_step": {k: v[0] / max(args.profile_steps, 1) for k, v in prof.items()},
                "kv_pages": st.kv_pages_total, "device_mem_gb": st.memory_used / 1e9}
        # prefill of the whole batch against the tensor roofline: 2 flop per weight and token + causal attention
        H_, d_, L_ = shape.hidden, shape.head_dim, shape.layers
        qd_, kd_ = shape.n_heads * d_, shape.n_kv_heads * d_
        p_mm = L_ * (H_ * qd_ + 2 * H_ * kd_ + qd_ * H_ + 3 * H_ * shape.inter) + shape.vocab * H_ * batch / (batch * args.prompt_len)
        pf_flops = (2.0 * p_mm * batch * args.prompt_len + 4.0 * L_ * qd_ * batch * args.prompt_len * (args.prompt_len + 1) / 2.0) / world
        tpk, tpk_src = tensor_peak()
        pf_tf = pf_flops / (statistics.median(ttft) * 1e-3) / 1e12
        line["prefill_roofline"] = {"bound": "tensor", "flops_per_gpu": pf_flops, "achieved": pf_tf, "peak": tpk, "unit": "TFLOP/s",
                                    "frac": pf_tf / tpk, "peak_source": tpk_src,
                                    "what": f"ttft_ms_p50: all {batch} x {args.prompt_len}-token prompts in one dis_engine_prefill call"}
        if parity is not None:
            line["parity"] = parity
        if tp1 is not None:
            line["tp1_same_batch"] = tp1
            line["scaling_vs_tp1_same_batch"] = {"speedup": value / tp1["value"], "of_linear": value / tp1["value"] / world}
        if args.layers:
            line["config"]["workload"] += f" [REDUCED DEPTH {args.layers} layers: not a valid headline number]"
        if world > 1:
            line["config"]["tp_exchange"] = "NCCL all-reduce" if args.nccl_allreduce else (
                "fused peer-memory exchange (k_tp.cu), " + (args.tp_exchange if args.tp_exchange != "auto" else ("wide (flag-in-data two-shot, one CTA per row and column block)" if shape.hidden > 4096 else "ll (flag-in-data two-shot, one CTA per row)")))
        if args.shard_of > 1:
            line["config"]["workload"] += f" [DIAGNOSTIC: one rank's shard of TP={args.shard_of}, no exchange: not a headline number]"
    eng.close()
    if dist is not None:
        dist.barrier(group=gloo)
        dist.destroy_process_group()
    # ---- CPU legs: rank 0, N = 1 only, OUTSIDE the process group (no GPU waits on them)
    if rank == 0 and world == 1 and not args.no_cpu:
        try:
            line["cpu_baseline"] = cpu_baseline_leg(batch, args.prompt_len)
        except Exception as ex:  # the CPU figure is a reported baseline; never lose the GPU line over it
            line["cpu_baseline"] = {"value": None, "unit": "tokens/s", "cores": host_cores(), "kind": "port", "sample": f"failed: {ex}"}
        try:
            line["cpu_config0"] = cpu_config0_leg()
        except Exception as ex:
            line["cpu_config0"] = {"failed": str(ex)}
    if rank == 0:
        print(json.dumps(line))
    return 0


def run_serving(args):
    """BASELINE configs[4]: continuous batching, 256 concurrent requests, prompt lengths log-uniform in [128, 8192]
    (seed 2), 128 new tokens each, through the admission queue + continuous batcher (dis_serving_*)."""
    import torch
    from distributed_inference_server_b200 import ContinuousBatcher, Engine, InferenceParams, LLAMA3_8B, ModelShape, WorkerConfig
    from distributed_inference_server_b200 import binding as B
    torch.cuda.set_device(0)
    shape = LLAMA3_8B
    if args.layers:
        shape = ModelShape(shape.vocab, shape.hidden, args.layers, shape.n_heads, shape.n_kv_heads, shape.head_dim,
                           shape.inter, shape.max_ctx, shape.rope_theta, shape.rms_eps)
    n_req, new = args.requests, args.new_tokens
    rng = np.random.default_rng(2)
    lens = np.exp(rng.uniform(np.log(128), np.log(8192 - new - 1), n_req)).astype(int)
    prompts = [rng.integers(0, shape.vocab, int(n)).tolist() for n in lens]
    page = 64
    pages = int(sum((n + new) // page + 2 for n in lens))
    eng = Engine(WorkerConfig(shape, device=0, max_batch=n_req, max_step_tokens=args.prefill_chunk, page_size=page,
                              kv_pages=pages, context_size=8192, stream=torch.cuda.current_stream().cuda_stream))
    eng.load_synthetic(0); eng.finalize()
    cb = ContinuousBatcher(eng, max_active=n_req, high_watermark=4 * n_req, low_watermark=2 * n_req, max_queue_size=8 * n_req,
                           request_timeout_s=3600.0)
    if args.prefill_budget:
        cb.set_prefill_budget(args.prefill_budget)
    torch.cuda.synchronize()
    t0 = time.perf_counter()
    for i, p in enumerate(prompts):
        cb.submit(i, p, InferenceParams(new, 0.0))
    # drive the scheduler; every token event is stamped with the host clock of the step that produced it
    tok_t = {i: [] for i in range(n_req)}
    finish = {}
    n_steps = 0
    while True:
        evs = cb.step()
        now = time.perf_counter()
        n_steps += 1
        for rid, kind, tok, idx, fin, ptoks, waited in evs:
            if kind == B.EVENT_TOKEN:
                tok_t[rid].append(now)
            elif kind == B.EVENT_DONE:
                finish[rid] = fin
        d = cb.depth()
        if d.total == 0 and d.active == 0 and not evs:
            break
    torch.cuda.synchronize()
    wall = time.perf_counter() - t0
    gen = sum(len(v) for v in tok_t.values())
    ttft = sorted((v[0] - t0) * 1e3 for v in tok_t.values() if v)
    itl = sorted((b - a) * 1e3 for v in tok_t.values() for a, b in zip(v, v[1:]))
    st = eng.status()
    pct = lambda xs, q: xs[min(len(xs) - 1, int(q * len(xs)))] if xs else None
    line = {"metric": "aggregate decode tokens/s, continuous batching", "value": gen / wall, "unit": "tokens/s", "n_gpus": 1,
            "higher_is_better": True, "dtype": "bf16", "data": "synthetic",
            "config": {"workload": f"BASELINE configs[4]: Llama-3-8B bf16, continuous batching, {n_req} concurrent requests (all submitted at t = 0), "
                                   f"prompt lengths log-uniform in [128, 8192] (mean {int(lens.mean())}), {new} new tokens each"
                                   + (f" [REDUCED DEPTH {args.layers} layers]" if args.layers else ""),
                       "scheduler": "one forward pass per step: every running sequence decodes a token, admitted prompts advance by at most "
                                    f"prefill_budget = {args.prefill_budget or 'library default (4096)'} tokens in admission order (dis_engine_step)"},
            "requests": n_req, "completed": sum(1 for f in finish.values() if f == B.FINISH_LENGTH),
            "generated_tokens": gen, "prompt_tokens": int(lens.sum()), "wall_s": wall, "scheduler_steps": n_steps,
            "ttft_ms_p50": pct(ttft, 0.5), "ttft_ms_max": ttft[-1] if ttft else None,
            "inter_token_ms_p50": pct(itl, 0.5), "inter_token_ms_p99": pct(itl, 0.99), "inter_token_ms_max": itl[-1] if itl else None,
            "total_tokens_per_s_incl_prompt": (gen + int(lens.sum())) / wall, "kv_pages": st.kv_pages_total,
            "gpu_launches": int(st.kernel_launches)}
    print(json.dumps(line))
    cb.close(); eng.close()
    return 0


def run_prefix(args):
    """Row f2 (prefix KV reuse): time to first token of one request on an idle engine, cold against a request whose
    leading tokens are already in the page pool (a shared system prompt).  Not a headline number: it shows what
    DIS_FLAG_PREFIX_CACHE buys; the decode bench runs with the cache off so that nothing is ever skipped there."""
    import torch
    from distributed_inference_server_b200 import Engine, InferenceParams, LLAMA3_8B, ModelShape, WorkerConfig
    from distributed_inference_server_b200 import binding as B
    torch.cuda.set_device(0)
    shape = LLAMA3_8B
    if args.layers:
        shape = ModelShape(shape.vocab, shape.hidden, args.layers, shape.n_heads, shape.n_kv_heads, shape.head_dim,
                           shape.inter, shape.max_ctx, shape.rope_theta, shape.rms_eps)
    page, runs = 64, 7
    rng = np.random.default_rng(3)
    eng = Engine(WorkerConfig(shape, device=0, max_batch=4, max_step_tokens=8192, page_size=page, kv_pages=1024,
                              context_size=8192, stream=torch.cuda.current_stream().cuda_stream, flags=B.FLAG_PREFIX_CACHE))
    eng.load_synthetic(0); eng.finalize()
    P = InferenceParams(max_tokens=4, temperature=0.0)
    out = {}
    for total, shared in ((512, 448), (2048, 1920), (8000, 7680)):
        def ttft(prompt):
            s = eng.seq_add(prompt, P)
            torch.cuda.synchronize()
            t0 = time.perf_counter()
            eng.prefill([s])
            torch.cuda.synchronize()
            dt = (time.perf_counter() - t0) * 1e3
            eng.seq_remove(s)
            return dt
        system = rng.integers(0, shape.vocab, shared).tolist()
        cold, hit = [], []
        for _ in range(runs):
            cold.append(ttft(rng.integers(0, shape.vocab, total).tolist()))          # nothing in common with anything cached
        ttft(system + rng.integers(0, shape.vocab, total - shared).tolist())         # registers the shared prefix
        before = eng.prefix_stats()["hit_tokens"]
        for _ in range(runs):
            hit.append(ttft(system + rng.integers(0, shape.vocab, total - shared).tolist()))
        reused = (eng.prefix_stats()["hit_tokens"] - before) // runs
        out[f"{total}"] = {"prompt_tokens": total, "shared_prefix_tokens": shared, "reused_tokens_per_request": int(reused),
                           "ttft_cold_ms_p50": statistics.median(cold), "ttft_prefix_hit_ms_p50": statistics.median(hit)}
    st = eng.prefix_stats()
    line = {"metric": "TTFT with prefix KV reuse", "value": out["2048"]["ttft_prefix_hit_ms_p50"], "unit": "ms", "n_gpus": 1,
            "higher_is_better": False, "dtype": "bf16", "data": "synthetic",
            "config": {"workload": "Llama-3-8B bf16, one request on an idle engine, DIS_FLAG_PREFIX_CACHE, page size 64"
                                   + (f" [REDUCED DEPTH {args.layers} layers]" if args.layers else "")},
            "by_prompt_length": out, "prefix_stats": st, "gpu_launches": int(eng.status().kernel_launches)}
    print(json.dumps(line))
    eng.close()
    return 0


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=128)
    ap.add_argument("--warmup", type=int, default=8)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--batch", type=int, default=0)
    ap.add_argument("--prompt-len", type=int, default=512)
    ap.add_argument("--prefill-chunk", type=int, default=8192)
    ap.add_argument("--profile-steps", type=int, default=3)
    ap.add_argument("--layers", type=int, default=0, help="debug only: reduced depth")
    ap.add_argument("--no-cpu", action="store_true", help="skip the cpu_baseline leg")
    ap.add_argument("--no-graph", action="store_true", help="run decode steps eagerly (profiling under ncu)")
    ap.add_argument("--ttft-runs", type=int, default=3)
    ap.add_argument("--no-pdl", action="store_true", help="disable programmatic dependent launch (A/B)")
    ap.add_argument("--nccl-allreduce", action="store_true", help="TP: use NCCL instead of the fused peer-memory exchange (A/B)")
    ap.add_argument("--tp-exchange", choices=["auto", "oneshot", "twoshot", "ll", "wide"], default="auto",
                    help="TP: peer-memory exchange kernel (auto = ll, the flag-in-data row kernel, up to hidden 4096; wide, the same protocol on a (row, column block) grid, above)")
    ap.add_argument("--workload", default="decode", choices=["decode", "serving", "prefix"], help="serving = BASELINE configs[4]; prefix = TTFT with prefix KV reuse (row f2)")
    ap.add_argument("--model", default="8b", choices=["8b", "70b"], help="70b = BASELINE configs[3] (TP=8, batch 32)")
    ap.add_argument("--requests", type=int, default=256)
    ap.add_argument("--prefill-budget", type=int, default=0, help="serving: prompt tokens read per scheduler step (0 = library default)")
    ap.add_argument("--shard-of", type=int, default=1, help="diagnostic: run one rank's shard of a TP=N model on one GPU, no exchange")
    ap.add_argument("--new-tokens", type=int, default=128)
    ap.add_argument("--settle-steps", type=int, default=48,
                    help="untimed decode steps before the W warm-up steps: lets the board leave the prefill's power state")
    ap.add_argument("--temperature", type=float, default=0.0, help="> 0: sampled decoding (not the headline; the TP parity gate is greedy-only and is skipped)")
    ap.add_argument("--top-p", type=float, default=0.9)
    ap.add_argument("--no-parity", action="store_true", help="TP: skip the parity check against a TP=1 engine on rank 0")
    args = ap.parse_args()
    if args.impl == "reference":
        return run_reference(args)
    if args.workload == "prefix":
        return run_prefix(args)
    if args.workload == "serving":
        return run_serving(args)
    return run_ours(args)


if __name__ == "__main__":
    sys.exit(main())
