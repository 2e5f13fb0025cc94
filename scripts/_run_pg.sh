timeout 300 python -m pytest tests/test_kernels_gpu.py -x -q -k "gemm" 2>&1 | tail -15
timeout 300 python -m pytest tests/test_engine_gpu.py -x -q 2>&1 | tail -3
for p in 1 0; do
DIS_GEMM_PERSIST=$p timeout 300 python bench.py --steps 8 --warmup 3 --no-cpu --ttft-runs 3 --profile-steps 0 > gpurun_out/pg_$p.json 2> gpurun_out/pg_$p.err
done
