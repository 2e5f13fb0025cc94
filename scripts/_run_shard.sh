CMD="python bench.py --steps 1 --warmup 1 --layers 4 --no-cpu --profile-steps 0 --no-graph --ttft-runs 1 --prompt-len 512 --shard-of 8 --batch 64"
$CMD > gpurun_out/plain_shard8.log 2>&1 &&
ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/launches_shard8.csv $CMD > gpurun_out/ncu_launch_shard8.log 2>&1
echo done
