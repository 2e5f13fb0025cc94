#!/bin/bash
# Round-2 final tensor-parallel session: usage scripts/gpu_r2_tpfinal.sh N [notest] -- the world-N tests, then the driver's
# two commands (GPU arm and reference arm) launched the way the driver launches them.
set -u
N=${1:-2}
mkdir -p gpurun_out
O=gpurun_out
[ "${2:-}" = notest ] || { timeout 1500 python -m pytest tests/test_tp_gpu.py -m gpu -q -x -k "tp$N or (not tp2 and not tp4 and not tp8)" --durations=5 > $O/r2tpf${N}_pytest.log 2>&1; echo "pytest rc=$?"; tail -6 $O/r2tpf${N}_pytest.log; }
L="python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29533"
timeout 900 $L bench.py --gpus $N --steps 20 --warmup 5 > $O/r2tpf${N}_bench.json 2> $O/r2tpf${N}_bench.err; echo "bench rc=$?"
python - $O/r2tpf${N}_bench.json <<'P'
import json,sys
try:
    d=json.loads(open(sys.argv[1]).read().strip().splitlines()[-1])
    p=d.get("parity") or {}
    print(" ", round(d["value"]), "tok/s", round(d["ms_per_step"],4), "ms/step; parity ok", p.get("ok"), "rms", p.get("decode_step_1_logits_rms_err_over_rms_logit"), "tp1", d.get("tp1_same_batch",{}).get("value"), d.get("scaling_vs_tp1_same_batch"), "ttft", d.get("ttft_ms_p50"), d["config"]["tp_exchange"])
except Exception as ex:
    print("  failed:", ex); print(open(sys.argv[1].replace(".json",".err")).read()[-2500:])
P
timeout 900 $L bench.py --impl reference --gpus $N --steps 5 --warmup 1 > $O/r2tpf${N}_ref.json 2> $O/r2tpf${N}_ref.err; echo "ref rc=$?"; cut -c1-300 $O/r2tpf${N}_ref.json
