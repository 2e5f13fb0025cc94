#!/bin/bash
# Round-2 GPU session F (1 GPU): full suite on the latest build, the driver's two commands, and where a TP=8 rank's
# shard spends its step (bench.py --shard-of 8: one rank's weights on one GPU, no exchange) by leaving kernel classes out.
set -u
mkdir -p gpurun_out
O=gpurun_out
timeout 1500 python -m pytest tests -m gpu -q -x --durations=5 > $O/r2f_pytest.log 2>&1; echo "pytest rc=$?"; tail -4 $O/r2f_pytest.log
python -c "import __graft_entry__ as g; g.smoke()" > $O/r2f_smoke.log 2>&1; echo "smoke rc=$?"; tail -2 $O/r2f_smoke.log
timeout 900 python bench.py --impl reference --gpus 1 --steps 20 --warmup 5 > $O/r2f_ref.json 2> $O/r2f_ref.err; echo "ref rc=$?"
timeout 900 python bench.py --gpus 1 --steps 20 --warmup 5 > $O/r2f_bench.json 2> $O/r2f_bench.err; echo "bench rc=$?"
python - <<'P'
import json
d=json.loads(open("gpurun_out/r2f_bench.json").read().strip().splitlines()[-1])
print(d["value"], d["ms_per_step"], d["roofline"]["frac"], d["roofline"]["traffic"], d["step_roofline"]["frac"], d["clocks"], d["ttft_ms_p50"], d.get("prefill_roofline",{}).get("frac"))
print(d["cpu_baseline"]["value"], d["cpu_config0"])
P
ab() { tag=$1; shift; envs=(); while [ "$1" != "--" ]; do envs+=("$1"); shift; done; shift
  env "${envs[@]}" timeout 600 python bench.py --no-cpu --ttft-runs 1 --profile-steps 0 "$@" > $O/r2f_$tag.json 2> $O/r2f_$tag.err
  python - $O/r2f_$tag.json $tag <<'P'
import json,sys
try:
    d=json.loads(open(sys.argv[1]).read().strip().splitlines()[-1])
    print(f"{sys.argv[2]:30s} {d['ms_per_step']:.4f} ms/step  graph_only {d.get('graph_only_ms_last_step'):.4f}")
except Exception as ex:
    print(sys.argv[2], "failed", ex)
P
}
for k in none norm qkv attn oproj gateup down; do ab sh8_skip_$k DIS_DEBUG_SKIP=$k -- --shard-of 8 --batch 64 --steps 40 --warmup 5; done
ab sh8_splits1 DIS_ATTN_MAX_SPLITS=1 -- --shard-of 8 --batch 64 --steps 40 --warmup 5
ab sh8_splits2 DIS_ATTN_MAX_SPLITS=2 -- --shard-of 8 --batch 64 --steps 40 --warmup 5
ab sh4_none X=1 -- --shard-of 4 --batch 64 --steps 40 --warmup 5
ab sh4_splits1 DIS_ATTN_MAX_SPLITS=1 -- --shard-of 4 --batch 64 --steps 40 --warmup 5
timeout 300 python bench.py --workload prefix > $O/r2f_prefix.json 2> $O/r2f_prefix.err; python - <<'P'
import json
d=json.loads(open("gpurun_out/r2f_prefix.json").read().strip().splitlines()[-1])
print({k:(v["ttft_cold_ms_p50"], v["ttft_prefix_hit_ms_p50"]) for k,v in d["by_prompt_length"].items()})
P
