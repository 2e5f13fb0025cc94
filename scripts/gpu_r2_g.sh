#!/bin/bash
# Round-2 GPU session G (1 GPU): the two-CTA-per-SM tcgen05 prefill attention (tc2) against the first variant (tc1).
set -u
mkdir -p gpurun_out
O=gpurun_out
for v in tc2 tc1; do
  DIS_ATTN_PREFILL=$v timeout 300 python -m pytest tests/test_kernels_gpu.py -m gpu -q -x -k "prefill" > $O/r2g_pf_$v.log 2>&1; echo "prefill kernel tests $v rc=$?"; tail -3 $O/r2g_pf_$v.log
done
timeout 1500 python -m pytest tests -m gpu -q -x > $O/r2g_pytest.log 2>&1; echo "pytest rc=$?"; tail -3 $O/r2g_pytest.log
for v in tc2 tc1 tc2 tc1; do
  DIS_ATTN_PREFILL=$v timeout 600 python bench.py --workload prefix > $O/r2g_prefix_$v.json 2> $O/r2g_prefix_$v.err; echo "prefix $v rc=$?"
  python - $O/r2g_prefix_$v.json <<'P'
import json,sys
d=json.loads(open(sys.argv[1]).read().strip().splitlines()[-1])
print({k:(round(v["ttft_cold_ms_p50"],2), round(v["ttft_prefix_hit_ms_p50"],2)) for k,v in d["by_prompt_length"].items()})
P
done
for v in tc2 tc1; do
  DIS_ATTN_PREFILL=$v timeout 600 python bench.py --workload serving > $O/r2g_serving_$v.json 2> $O/r2g_serving_$v.err; echo "serving $v rc=$?"
  python - $O/r2g_serving_$v.json <<'P'
import json,sys
d=json.loads(open(sys.argv[1]).read().strip().splitlines()[-1])
print({k:d[k] for k in ("value","wall_s","ttft_ms_p50","ttft_ms_max","inter_token_ms_p50","inter_token_ms_p99")})
P
done
DIS_ATTN_PREFILL=tc2 timeout 600 python bench.py --gpus 1 --steps 20 --warmup 5 --no-cpu > $O/r2g_bench.json 2> $O/r2g_bench.err
python - <<'P'
import json
d=json.loads(open("gpurun_out/r2g_bench.json").read().strip().splitlines()[-1])
print(d["value"], d["ms_per_step"], "ttft", d["ttft_ms_p50"], d["ttft_single_ms_p50"], d["prefill_roofline"]["frac"])
P
CMD="timeout 300 python bench.py --batch 1 --prompt-len 2048 --steps 1 --warmup 1 --settle-steps 0 --layers 2 --no-cpu --profile-steps 0 --ttft-runs 1"
$CMD > $O/r2g_plain_pf.log 2>&1 && \
ncu --set full --clock-control none --import-source on --kernel-name-base demangled -k regex:attn_prefill_tc -c 2 -f -o $O/prof_r2g_pf $CMD > $O/r2g_ncu_pf.log 2>&1
echo "ncu rc=$?"
