python -m pytest tests -m gpu -x -q 2>&1 | tail -4
python bench.py --workload prefix > gpurun_out/prefix_bench.json 2> gpurun_out/prefix_bench.err; tail -c 1500 gpurun_out/prefix_bench.json
