timeout 600 python -m pytest tests/test_tp_gpu.py -x -q 2>&1 | tail -15 > gpurun_out/tp_test.log; tail -3 gpurun_out/tp_test.log
for m in ll oneshot; do
timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29511 bench.py --gpus 2 --steps 64 --warmup 8 --no-cpu --ttft-runs 0 --tp-exchange $m > gpurun_out/tp2_$m.json 2> gpurun_out/tp2_$m.err
done
