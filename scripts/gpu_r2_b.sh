#!/bin/bash
# Round-2 GPU session B (1 GPU): all -m gpu tests; marginal in-graph cost of each kernel class (DIS_DEBUG_SKIP);
# the continuous-batching workload (BASELINE configs[4]) on the mixed-step batcher.
set -u
mkdir -p gpurun_out
O=gpurun_out
timeout 1500 python -m pytest tests -m gpu -q --durations=8 > $O/r2b_pytest.log 2>&1; echo "pytest rc=$?"
tail -15 $O/r2b_pytest.log
for k in none norm qkv attn oproj gateup down "norm,qkv,attn,oproj,down"; do
  DIS_DEBUG_SKIP=$k timeout 300 python bench.py --steps 40 --warmup 5 --no-cpu --profile-steps 0 --ttft-runs 1 > $O/r2b_skip_$k.json 2> $O/r2b_skip_$k.err
  python - "$k" <<'P'
import json,sys
k=sys.argv[1]
try:
    d=json.loads(open(f"gpurun_out/r2b_skip_{k}.json").read().strip().splitlines()[-1])
    print(f"skip {k:28s} ms/step {d['ms_per_step']:.4f} graph_only {d['graph_only_ms_last_step']:.4f}")
except Exception as ex:
    print("skip", k, "failed", ex)
P
done
timeout 600 python bench.py --workload serving > $O/r2b_serving.json 2> $O/r2b_serving.err; echo "serving rc=$?"
cat $O/r2b_serving.json | cut -c1-900
