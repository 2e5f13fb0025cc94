#!/bin/bash
# Round-2 GPU session O (1 GPU): the suite on the final build, then BASELINE configs[4] with the default prompt order.
set -u
mkdir -p gpurun_out
O=gpurun_out
timeout 70 python -m pytest tests -m gpu -q -x > $O/r2o_pytest.log 2>&1; echo "pytest rc=$?"; tail -2 $O/r2o_pytest.log
timeout 60 python bench.py --workload serving > $O/r2o_serving.json 2> $O/r2o_serving.err; echo "serving rc=$?"
python - <<'P'
import json
d=json.loads(open("gpurun_out/r2o_serving.json").read().strip().splitlines()[-1])
print({k:d.get(k) for k in ("value","wall_s","ttft_ms_p50","ttft_ms_max","inter_token_ms_p50","inter_token_ms_p99","completed")})
P
