#!/bin/bash
# Round-2 GPU session K (1 GPU): suite + smoke on the final tree, A/B of the norm kernel's early gain fetch
# (alternating runs on one box), then the driver's bench command.
set -u
mkdir -p gpurun_out
O=gpurun_out
timeout 1500 python -m pytest tests -m gpu -q -x --durations=5 > $O/r2k_pytest.log 2>&1; echo "pytest rc=$?"; tail -4 $O/r2k_pytest.log
python -c "import __graft_entry__ as g; g.smoke()" > $O/r2k_smoke.log 2>&1; echo "smoke rc=$?"; tail -1 $O/r2k_smoke.log
for i in 1 2; do for g in 0 1; do
  DIS_NORM_EARLY_GAIN=$g timeout 600 python bench.py --gpus 1 --steps 20 --warmup 5 --no-cpu > $O/r2k_gain${g}_$i.json 2> $O/r2k_gain${g}_$i.err
  python - "$O/r2k_gain${g}_$i.json" "gain=$g run $i" <<'P'
import json, sys
d=json.loads(open(sys.argv[1]).read().strip().splitlines()[-1]); print(sys.argv[2], d["value"], d["ms_per_step"], d.get("graph_only_ms_last_step"))
P
done; done
timeout 900 python bench.py --gpus 1 --steps 20 --warmup 5 > $O/r2k_bench.json 2> $O/r2k_bench.err; echo "bench rc=$?"; tail -c 1500 $O/r2k_bench.json
