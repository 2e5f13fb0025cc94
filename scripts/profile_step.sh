#!/bin/bash
# One decode step under ncu (run on the GPU box through gpurun), one ncu pass per call; the same command runs
# plainly first and ncu only follows if that exited 0.
# usage: scripts/profile_step.sh launches <tag>
#        scripts/profile_step.sh full <tag> [kernel-regex] [skip] [count]
set -u
MODE=${1:-launches}
TAG=${2:-r1}
KRE=${3:-gemm_tcgen05_kernel}
SKIP=${4:-0}
CNT=${5:-2}
CMD="timeout 200 python bench.py --steps 1 --warmup 1 --layers 4 --no-cpu --profile-steps 0 --no-graph --ttft-runs 1 --prompt-len 512"
if [ "$MODE" = launches ]; then
  $CMD > gpurun_out/plain_$TAG.log 2>&1 &&
  ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/launches_$TAG.csv $CMD > gpurun_out/ncu_launch_$TAG.log 2>&1
else
  $CMD > gpurun_out/plain2_$TAG.log 2>&1 &&
  ncu --set full --clock-control none --import-source on --kernel-name-base demangled -k regex:$KRE -s $SKIP -c $CNT -f -o gpurun_out/prof_$TAG $CMD > gpurun_out/ncu_full_$TAG.log 2>&1
fi
echo done
