#!/bin/bash
# Round-2 GPU session (N GPUs): the tensor-parallel tests that sample (fused vocab-parallel sampler is the default).
set -u
mkdir -p gpurun_out
O=gpurun_out
N=$(python -c "import torch; print(torch.cuda.device_count())")
timeout 900 python -m pytest tests/test_tp_gpu.py -m gpu -q -x -k "(${1:-sampled or vocab_parallel}) and tp$N" --durations=3 > $O/r2_tpsample_w$N.log 2>&1; echo "world $N rc=$?"; tail -8 $O/r2_tpsample_w$N.log
