#!/bin/bash
# Round-2 GPU session M (1 GPU): the vocab-parallel sampler with its ranks emulated on one GPU, then the whole suite.
set -u
mkdir -p gpurun_out
O=gpurun_out
timeout 300 python -m pytest tests/test_kernels_gpu.py -m gpu -q -x -k "vocab_parallel" > $O/r2m_vp.log 2>&1; echo "vocab-parallel rc=$?"; tail -15 $O/r2m_vp.log
timeout 1200 python -m pytest tests -m gpu -q --durations=5 > $O/r2m_pytest.log 2>&1; echo "pytest rc=$?"; tail -4 $O/r2m_pytest.log
