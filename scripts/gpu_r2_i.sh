#!/bin/bash
# Round-2 GPU session I (1 GPU): suite on the final tree (graph-bucket padding), serving workload.
set -u
mkdir -p gpurun_out
O=gpurun_out
timeout 1500 python -m pytest tests -m gpu -q -x --durations=5 > $O/r2i_pytest.log 2>&1; echo "pytest rc=$?"; tail -4 $O/r2i_pytest.log
python -c "import __graft_entry__ as g; g.smoke()" > $O/r2i_smoke.log 2>&1; echo "smoke rc=$?"; tail -1 $O/r2i_smoke.log
timeout 600 python bench.py --workload serving > $O/r2i_serving.json 2> $O/r2i_serving.err; echo "serving rc=$?"
python - <<'P'
import json
d=json.loads(open("gpurun_out/r2i_serving.json").read().strip().splitlines()[-1])
print({k:d[k] for k in ("value","wall_s","ttft_ms_p50","ttft_ms_max","inter_token_ms_p50","inter_token_ms_p99","inter_token_ms_max","scheduler_steps")})
P
timeout 600 python bench.py --gpus 1 --steps 20 --warmup 5 --no-cpu > $O/r2i_bench.json 2> $O/r2i_bench.err; python - <<'P'
import json
d=json.loads(open("gpurun_out/r2i_bench.json").read().strip().splitlines()[-1]); print(d["value"], d["ms_per_step"])
P
