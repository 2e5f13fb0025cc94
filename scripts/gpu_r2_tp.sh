#!/bin/bash
# Round-2 tensor-parallel session: usage scripts/gpu_r2_tp.sh N [notest]  (run under gpurun --gpus N)
set -u
N=${1:-2}
mkdir -p gpurun_out
O=gpurun_out
nvidia-smi topo -m > $O/r2tp${N}_topo.txt 2>&1
[ "${2:-}" = notest ] || timeout 900 python -m pytest tests/test_tp_gpu.py -m gpu -q -x -k "tp$N or (not tp2 and not tp4 and not tp8)" --durations=5 > $O/r2tp${N}_pytest.log 2>&1; echo "pytest rc=$?"
tail -8 $O/r2tp${N}_pytest.log
run() { # tag, extra args
  tag=$1; shift
  timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29511 \
    bench.py --gpus $N --steps 40 --warmup 5 "$@" > $O/r2tp${N}_$tag.json 2> $O/r2tp${N}_$tag.err
  echo "bench $tag rc=$?"
  python - $O/r2tp${N}_$tag.json <<'P'
import json,sys
try:
    d=json.loads(open(sys.argv[1]).read().strip().splitlines()[-1])
    print(" ", d["value"], "tok/s", d["ms_per_step"], "ms/step", "parity", json.dumps(d.get("parity"))[:300], "tp1", d.get("tp1_same_batch",{}).get("value"), d.get("scaling_vs_tp1_same_batch"))
    print("  classes", d.get("kernel_class_ms_per_step"))
except Exception as ex:
    print("  failed:", ex); print(open(sys.argv[1].replace(".json",".err")).read()[-1500:])
P
}
run wide
run ll --tp-exchange ll --no-parity
if [ "$N" = 8 ]; then run 70b_ll --model 70b --no-parity --tp-exchange ll; run 70b_wide --model 70b --no-parity --tp-exchange wide; fi
