#!/bin/bash
# usage: scripts/gpu_r2_tpstream.sh N -- the streaming prefill exchange: its test at world N (2 or 4) and TTFT against the NCCL path
set -u
N=${1:-2}
mkdir -p gpurun_out
O=gpurun_out
if [ "$N" -le 4 ]; then timeout 600 python -m pytest tests/test_tp_gpu.py -m gpu -q -x -k "prefill_pass and tp$N" > $O/r2tps${N}_pytest.log 2>&1; echo "pytest rc=$?"; tail -4 $O/r2tps${N}_pytest.log; fi
L="python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29544"
for mode in fused nccl; do
  if [ $mode = fused ]; then export DIS_TP_PREFILL_FUSED=1; else unset DIS_TP_PREFILL_FUSED; fi
  timeout 900 $L bench.py --gpus $N --steps 20 --warmup 5 > $O/r2tps${N}_$mode.json 2> $O/r2tps${N}_$mode.err; echo "bench $mode rc=$?"
  python - $O/r2tps${N}_$mode.json <<'P'
import json,sys
try:
    d=json.loads(open(sys.argv[1]).read().strip().splitlines()[-1])
    print("  ttft", d.get("ttft_ms_p50"), d.get("ttft_ms_all"), round(d["value"]), "tok/s parity", (d.get("parity") or {}).get("ok"))
except Exception as ex:
    print("  failed:", ex); print(open(sys.argv[1].replace(".json",".err")).read()[-2000:])
P
done
