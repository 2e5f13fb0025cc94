for m in ll oneshot; do
timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node 4 --master-addr 127.0.0.1 --master-port 29511 bench.py --gpus 4 --steps 64 --warmup 8 --no-cpu --ttft-runs 0 --tp-exchange $m > gpurun_out/tp4_$m.json 2> gpurun_out/tp4_$m.err
done
