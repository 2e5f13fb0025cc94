set -e
python -m pytest tests/test_kernels_gpu.py -x -q -k "gemv" 2>&1 | tail -3
for v in 0 1 2; do
DIS_GEMV_VARIANT=$v python bench.py --batch 1 --steps 64 --warmup 8 --no-cpu --ttft-runs 0 > gpurun_out/b1_v$v.json 2> gpurun_out/b1_v$v.err
done
