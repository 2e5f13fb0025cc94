for m in ll twoshot; do
timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port 29511 bench.py --gpus 8 --steps 64 --warmup 8 --no-cpu --ttft-runs 0 --tp-exchange $m > gpurun_out/tp8_$m.json 2> gpurun_out/tp8_$m.err
done
timeout 400 python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port 29512 bench.py --gpus 8 --model 70b --steps 32 --warmup 4 --no-cpu --ttft-runs 0 --tp-exchange ll > gpurun_out/tp8_70b_ll.json 2> gpurun_out/tp8_70b_ll.err
