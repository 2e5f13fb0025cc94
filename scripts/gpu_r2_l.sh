#!/bin/bash
# Round-2 GPU session L (1 GPU): the suite with guard bands around every buffer of the kernel tests, smoke.
set -u
mkdir -p gpurun_out
O=gpurun_out
timeout 1500 python -m pytest tests -m gpu -q --durations=5 > $O/r2l_pytest.log 2>&1; echo "pytest rc=$?"; tail -6 $O/r2l_pytest.log
python -c "import __graft_entry__ as g; g.smoke()" > $O/r2l_smoke.log 2>&1; echo "smoke rc=$?"; tail -1 $O/r2l_smoke.log
