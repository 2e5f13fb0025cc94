#!/bin/bash
# Round-2 GPU session D (1 GPU): full suite; the 8B parity report; A/B of the early weight TMA issue and of the
# norm-fused batch-1 GEMV; serving with the graph-decode fall-through; ncu --set full of the tcgen05 prefill attention.
set -u
mkdir -p gpurun_out
O=gpurun_out
timeout 1500 python -m pytest tests -m gpu -q -x --durations=5 > $O/r2d_pytest.log 2>&1; echo "pytest rc=$?"
tail -6 $O/r2d_pytest.log
timeout 600 python -m pytest tests/test_baseline_configs_gpu.py -m gpu -q -s 2>&1 | grep -E "parity report|margin report|passed|failed" | cut -c1-420 > $O/r2d_parity_report.txt; cat $O/r2d_parity_report.txt | tail -8
ab() { # tag env... -- args
  tag=$1; shift
  envs=(); while [ "$1" != "--" ]; do envs+=("$1"); shift; done; shift
  env "${envs[@]}" timeout 600 python bench.py --no-cpu "$@" > $O/r2d_$tag.json 2> $O/r2d_$tag.err
  python - $O/r2d_$tag.json $tag <<'P'
import json,sys
try:
    d=json.loads(open(sys.argv[1]).read().strip().splitlines()[-1])
    print(f"{sys.argv[2]:28s} {d['value']:9.1f} tok/s  {d['ms_per_step']:.4f} ms/step  graph_only {d.get('graph_only_ms_last_step')}  ttft {d.get('ttft_ms_p50'):.1f} / {d.get('ttft_single_ms_p50'):.2f}")
except Exception as ex:
    print(sys.argv[2], "failed", ex)
P
}
ab early1 DIS_GEMM_EARLY_TMA=1 -- --steps 40 --warmup 5
ab early0 DIS_GEMM_EARLY_TMA=0 -- --steps 40 --warmup 5
ab early1b DIS_GEMM_EARLY_TMA=1 -- --steps 40 --warmup 5
ab early0b DIS_GEMM_EARLY_TMA=0 -- --steps 40 --warmup 5
ab b1_fused DIS_GEMV_FUSE_NORM=1 -- --batch 1 --steps 64 --warmup 8
ab b1_unfused DIS_GEMV_FUSE_NORM=0 -- --batch 1 --steps 64 --warmup 8
ab b1_tcgen05 DIS_GEMV_MAX_T=0 -- --batch 1 --steps 64 --warmup 8
ab b1_fused_b DIS_GEMV_FUSE_NORM=1 -- --batch 1 --steps 64 --warmup 8
timeout 600 python bench.py --workload serving > $O/r2d_serving.json 2> $O/r2d_serving.err; echo "serving rc=$?"
python - <<'P'
import json
d=json.loads(open("gpurun_out/r2d_serving.json").read().strip().splitlines()[-1])
print({k:d[k] for k in ("value","wall_s","ttft_ms_p50","ttft_ms_max","inter_token_ms_p50","inter_token_ms_p99","inter_token_ms_max","scheduler_steps")})
P
# ncu --set full of the tcgen05 prefill attention: one 2048-token request, 2 layers
CMD="timeout 300 python bench.py --batch 1 --prompt-len 2048 --steps 1 --warmup 1 --settle-steps 0 --layers 2 --no-cpu --profile-steps 0 --ttft-runs 1"
$CMD > $O/r2d_plain_pf.log 2>&1 && \
ncu --set full --clock-control none --import-source on --kernel-name-base demangled -k regex:attn_prefill_tc -c 2 -f -o $O/prof_r2d_pf $CMD > $O/r2d_ncu_pf.log 2>&1
echo "ncu prefill attention rc=$?"; ls -la $O/prof_r2d_pf.ncu-rep
