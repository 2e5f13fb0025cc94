#!/bin/bash
# Round-2 GPU session E (1 GPU): does leaving shared memory free for the NEXT kernel's CTAs pay?  A/B of the ring depths.
set -u
mkdir -p gpurun_out
O=gpurun_out
timeout 900 python -m pytest tests -m gpu -q -x > $O/r2e_pytest.log 2>&1; echo "pytest rc=$?"; tail -3 $O/r2e_pytest.log
ab() { # tag env... -- args
  tag=$1; shift
  envs=(); while [ "$1" != "--" ]; do envs+=("$1"); shift; done; shift
  env "${envs[@]}" timeout 600 python bench.py --no-cpu --ttft-runs 1 --profile-steps 0 "$@" > $O/r2e_$tag.json 2> $O/r2e_$tag.err
  python - $O/r2e_$tag.json $tag <<'P'
import json,sys
try:
    d=json.loads(open(sys.argv[1]).read().strip().splitlines()[-1])
    print(f"{sys.argv[2]:34s} {d['value']:9.1f} tok/s  {d['ms_per_step']:.4f} ms/step  graph_only {d.get('graph_only_ms_last_step'):.4f}")
except Exception as ex:
    print(sys.argv[2], "failed", ex)
P
}
ab base X=1 -- --steps 40 --warmup 5
ab attn2 DIS_ATTN_STAGES=2 -- --steps 40 --warmup 5
ab gemm72 DIS_GEMM_SMEM_KB2=72 -- --steps 40 --warmup 5
ab gemm72_attn2 DIS_GEMM_SMEM_KB2=72 DIS_ATTN_STAGES=2 -- --steps 40 --warmup 5
ab gemm56_attn2 DIS_GEMM_SMEM_KB2=56 DIS_ATTN_STAGES=2 -- --steps 40 --warmup 5
ab gu150 DIS_GEMM_SMEM_KB1=150 -- --steps 40 --warmup 5
ab gu110_gemm72_attn2 DIS_GEMM_SMEM_KB1=110 DIS_GEMM_SMEM_KB2=72 DIS_ATTN_STAGES=2 -- --steps 40 --warmup 5
ab gu150_gemm72_attn2 DIS_GEMM_SMEM_KB1=150 DIS_GEMM_SMEM_KB2=72 DIS_ATTN_STAGES=2 -- --steps 40 --warmup 5
ab base_b X=1 -- --steps 40 --warmup 5
ab b1 X=1 -- --batch 1 --steps 64 --warmup 8
ab b64 X=1 -- --batch 64 --steps 40 --warmup 5
ab b64_gemm72_attn2 DIS_GEMM_SMEM_KB2=72 DIS_ATTN_STAGES=2 -- --batch 64 --steps 40 --warmup 5
