#!/bin/bash
# Round-2 GPU session C (1 GPU): the tcgen05 prefill attention -- kernel tests first (under a timeout: a wrong
# descriptor can hang), then the whole suite, then TTFT numbers against the mma.sync kernel.
set -u
mkdir -p gpurun_out
O=gpurun_out
timeout 300 python -m pytest tests/test_kernels_gpu.py -m gpu -q -x -k "prefill" > $O/r2c_pf.log 2>&1; echo "prefill kernel tests rc=$?"
tail -15 $O/r2c_pf.log
timeout 1500 python -m pytest tests -m gpu -q --durations=5 > $O/r2c_pytest.log 2>&1; echo "pytest rc=$?"
tail -12 $O/r2c_pytest.log
grep "parity report" $O/r2c_pytest.log | cut -c1-400
for mode in tc mma; do
  DIS_ATTN_PREFILL=$mode timeout 600 python bench.py --workload prefix > $O/r2c_prefix_$mode.json 2> $O/r2c_prefix_$mode.err; echo "prefix $mode rc=$?"
  python - $O/r2c_prefix_$mode.json <<'P'
import json,sys
d=json.loads(open(sys.argv[1]).read().strip().splitlines()[-1])
print({k:(v["ttft_cold_ms_p50"], v["ttft_prefix_hit_ms_p50"]) for k,v in d["by_prompt_length"].items()})
P
done
timeout 600 python bench.py --gpus 1 --steps 20 --warmup 5 --no-cpu > $O/r2c_bench.json 2> $O/r2c_bench.err; echo "bench rc=$?"
python - <<'P'
import json
d=json.loads(open("gpurun_out/r2c_bench.json").read().strip().splitlines()[-1])
print(d["value"], d["ms_per_step"], "ttft", d["ttft_ms_p50"], d["ttft_single_ms_p50"])
P
for b in 2048 4096 8192; do
  timeout 600 python bench.py --workload serving --prefill-budget $b > $O/r2c_serving_$b.json 2> $O/r2c_serving_$b.err; echo "serving $b rc=$?"
  cut -c1-200 $O/r2c_serving_$b.json; python - $O/r2c_serving_$b.json <<'P'
import json,sys
d=json.loads(open(sys.argv[1]).read().strip().splitlines()[-1])
print({k:d[k] for k in ("value","wall_s","ttft_ms_p50","ttft_ms_max","inter_token_ms_p50","inter_token_ms_p99","inter_token_ms_max","scheduler_steps")})
P
done
