#!/bin/bash
# Round-2 GPU session N (1 GPU): suite + smoke on the final build; sampled decoding on one GPU, bucket search against bisection.
set -u
mkdir -p gpurun_out
O=gpurun_out
timeout 600 python -m pytest tests -m gpu -q --durations=3 > $O/r2n_pytest.log 2>&1; echo "pytest rc=$?"; tail -3 $O/r2n_pytest.log
python -c "import __graft_entry__ as g; g.smoke()" > $O/r2n_smoke.log 2>&1; echo "smoke rc=$?"; tail -1 $O/r2n_smoke.log
for mode in buckets bisect; do
  DIS_SAMPLER=$mode timeout 200 python bench.py --gpus 1 --steps 40 --warmup 5 --temperature 0.8 --top-p 0.9 --no-cpu --ttft-runs 1 --profile-steps 1 \
    > $O/r2n_sampled_$mode.json 2> $O/r2n_sampled_$mode.err
  python - "$O/r2n_sampled_$mode.json" "$mode" <<'P'
import json, sys
try:
    d=json.loads(open(sys.argv[1]).read().strip().splitlines()[-1]); print(sys.argv[2], round(d["value"]), round(d["ms_per_step"],4), d.get("kernel_class_ms_per_step",{}).get("sample"))
except Exception as e:
    print(sys.argv[2], "failed", e)
P
done
