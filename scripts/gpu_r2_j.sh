#!/bin/bash
# Round-2 GPU session J (1 GPU): ncu --set full of the kernels no capture covers yet -- the prefill pass (persistent GEMMs,
# RoPE + KV write, prefill-sized norm, tcgen05 attention at batch prefill shape) and the batch-1 GEMV.
set -u
mkdir -p gpurun_out
O=gpurun_out
CMD="timeout 300 python bench.py --steps 1 --warmup 1 --settle-steps 0 --layers 2 --no-cpu --profile-steps 0 --ttft-runs 1"
$CMD > $O/r2j_plain.log 2>&1 && \
ncu --set full --clock-control none --kernel-name-base demangled -k regex:"gemm_persist|rope_kv_write|add_rmsnorm|attn_prefill_tc2|gemm_tcgen05" -c 24 -f -o $O/prof_r2j_prefill $CMD > $O/r2j_ncu_prefill.log 2>&1
echo "ncu prefill rc=$?"
CMD2="env DIS_GEMV_MAX_T=1 timeout 300 python bench.py --batch 1 --steps 1 --warmup 1 --settle-steps 0 --layers 2 --no-cpu --profile-steps 0 --ttft-runs 1 --no-graph"
$CMD2 > $O/r2j_plain2.log 2>&1 && \
ncu --set full --clock-control none --kernel-name-base demangled -k regex:gemv_kernel -c 10 -f -o $O/prof_r2j_gemv $CMD2 > $O/r2j_ncu_gemv.log 2>&1
echo "ncu gemv rc=$?"
ls -la $O/prof_r2j_*.ncu-rep
