python bench.py --batch 64 --steps 64 --warmup 8 --no-cpu --ttft-runs 0 > gpurun_out/b64_tp1.json 2> gpurun_out/b64_tp1.err
for n in 2 4 8; do
python bench.py --batch 64 --shard-of $n --steps 64 --warmup 8 --no-cpu --ttft-runs 0 > gpurun_out/b64_shard$n.json 2> gpurun_out/b64_shard$n.err
done
