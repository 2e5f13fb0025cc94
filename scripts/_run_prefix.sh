python -m pytest tests/test_prefix_gpu.py -x -q 2>&1 | tail -40 > gpurun_out/prefix_test.log; tail -5 gpurun_out/prefix_test.log
python -m pytest tests/test_engine_gpu.py tests/test_serving_gpu.py -x -q 2>&1 | tail -5
