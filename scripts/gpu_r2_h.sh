#!/bin/bash
# Round-2 GPU session H (1 GPU): final validation of the tree as committed + A/B of the full-ring K/V prefetch.
set -u
mkdir -p gpurun_out
O=gpurun_out
timeout 1500 python -m pytest tests -m gpu -q -x --durations=5 > $O/r2h_pytest.log 2>&1; echo "pytest rc=$?"; tail -4 $O/r2h_pytest.log
python -c "import __graft_entry__ as g; g.smoke()" > $O/r2h_smoke.log 2>&1; echo "smoke rc=$?"; tail -1 $O/r2h_smoke.log
ab() { tag=$1; shift; envs=(); while [ "$1" != "--" ]; do envs+=("$1"); shift; done; shift
  env "${envs[@]}" timeout 600 python bench.py --no-cpu --ttft-runs 1 --profile-steps 0 "$@" > $O/r2h_$tag.json 2> $O/r2h_$tag.err
  python - $O/r2h_$tag.json $tag <<'P'
import json,sys
try:
    d=json.loads(open(sys.argv[1]).read().strip().splitlines()[-1])
    print(f"{sys.argv[2]:30s} {d['value']:9.1f} tok/s {d['ms_per_step']:.4f} ms/step  graph_only {d.get('graph_only_ms_last_step'):.4f}")
except Exception as ex:
    print(sys.argv[2], "failed", ex)
P
}
ab ring1 DIS_ATTN_FULL_RING=1 -- --steps 40 --warmup 5
ab ring0 DIS_ATTN_FULL_RING=0 -- --steps 40 --warmup 5
ab ring1b DIS_ATTN_FULL_RING=1 -- --steps 40 --warmup 5
ab ring0b DIS_ATTN_FULL_RING=0 -- --steps 40 --warmup 5
ab ring1_b64 DIS_ATTN_FULL_RING=1 -- --batch 64 --steps 40 --warmup 5
ab ring0_b64 DIS_ATTN_FULL_RING=0 -- --batch 64 --steps 40 --warmup 5
ab ring1_b1 DIS_ATTN_FULL_RING=1 -- --batch 1 --steps 64 --warmup 8
ab ring0_b1 DIS_ATTN_FULL_RING=0 -- --batch 1 --steps 64 --warmup 8
timeout 900 python bench.py --gpus 1 --steps 20 --warmup 5 > $O/r2h_bench.json 2> $O/r2h_bench.err; echo "bench rc=$?"
python - <<'P'
import json
d=json.loads(open("gpurun_out/r2h_bench.json").read().strip().splitlines()[-1])
print(d["value"], d["ms_per_step"], d["roofline"]["frac"], d["step_roofline"]["frac"], d["ttft_ms_p50"], d["ttft_single_ms_p50"], d["cpu_baseline"]["value"])
P
