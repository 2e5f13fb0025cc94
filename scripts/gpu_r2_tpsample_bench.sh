#!/bin/bash
# Round-2 GPU session (2 GPUs): sampled decoding under tensor parallelism, the vocab-parallel sampler (default) against
# the gathered one (DIS_TP_SAMPLE=gather), alternating short runs on one box.
set -u
mkdir -p gpurun_out
O=gpurun_out
N=$(python -c "import torch; print(torch.cuda.device_count())")
for mode in fused gather; do
  i=$(ls $O/r2_tpsample_bench_${mode}_*.json 2>/dev/null | wc -l)
  DIS_TP_SAMPLE=$mode timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29513 \
    bench.py --gpus $N --steps 40 --warmup 5 --temperature 0.8 --top-p 0.9 --no-cpu --ttft-runs 1 --profile-steps 1 \
    > $O/r2_tpsample_bench_${mode}_$i.json 2> $O/r2_tpsample_bench_${mode}_$i.err
  python - "$O/r2_tpsample_bench_${mode}_$i.json" "$mode" <<'P'
import json, sys
try:
    d=json.loads(open(sys.argv[1]).read().strip().splitlines()[-1]); print(sys.argv[2], round(d["value"]), round(d["ms_per_step"],4), d.get("kernel_class_ms_per_step",{}).get("sample"))
except Exception as e:
    print(sys.argv[2], "failed", e)
P
done
