#!/bin/bash
# One decode step under ncu (run on the GPU box through gpurun).  The same command runs plainly first.
# usage: scripts/profile_step.sh <tag> [kernel-regex for the --set full capture] [skip] [count]
set -u
TAG=${1:-r1}
KRE=${2:-gemm_tcgen05_kernel}
SKIP=${3:-0}
CNT=${4:-2}
CMD="python bench.py --steps 1 --warmup 1 --layers 4 --no-cpu --profile-steps 0 --no-graph --ttft-runs 1 --prompt-len 512"
$CMD > gpurun_out/plain_$TAG.log 2>&1 &&
ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/launches_$TAG.csv $CMD > gpurun_out/ncu_launch_$TAG.log 2>&1
$CMD > gpurun_out/plain2_$TAG.log 2>&1 &&
ncu --set full --clock-control none --import-source on --kernel-name-base demangled -k regex:$KRE -s $SKIP -c $CNT -f -o gpurun_out/prof_$TAG $CMD > gpurun_out/ncu_full_$TAG.log 2>&1
echo done
