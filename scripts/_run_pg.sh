timeout 300 python -m pytest tests/test_engine_gpu.py tests/test_kernels_gpu.py -x -q 2>&1 | tail -3
bash scripts/profile_step.sh launches r1e
