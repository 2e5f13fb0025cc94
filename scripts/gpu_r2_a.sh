#!/bin/bash
# Round-2 GPU session A (1 GPU): parity tests, the driver's bench command, longer/other-batch bench lines,
# the reference arm, then the ncu passes (launch list; --set full of one whole decode step's kernels).
set -u
mkdir -p gpurun_out
O=gpurun_out
nvidia-smi --query-gpu=name,clocks.sm,clocks.max.sm,power.draw --format=csv > $O/r2a_smi.txt 2>&1
free -g > $O/r2a_host.txt; nproc >> $O/r2a_host.txt
timeout 1500 python -m pytest tests -m gpu -q -x --durations=8 > $O/r2a_pytest.log 2>&1; echo "pytest rc=$?"
tail -5 $O/r2a_pytest.log
timeout 600 python bench.py --gpus 1 --steps 20 --warmup 5 > $O/r2a_bench.json 2> $O/r2a_bench.err; echo "bench rc=$?"
timeout 600 python bench.py --gpus 1 --steps 128 --warmup 8 --no-cpu > $O/r2a_bench128.json 2> $O/r2a_bench128.err; echo "bench128 rc=$?"
timeout 600 python bench.py --gpus 1 --steps 20 --warmup 5 --settle-steps 0 --no-cpu > $O/r2a_bench_nosettle.json 2> $O/r2a_bench_nosettle.err; echo "bench nosettle rc=$?"
timeout 600 python bench.py --gpus 1 --batch 1 --steps 64 --warmup 8 --no-cpu > $O/r2a_bench_b1.json 2> $O/r2a_bench_b1.err; echo "bench b1 rc=$?"
timeout 900 python bench.py --impl reference --gpus 1 --steps 20 --warmup 5 > $O/r2a_ref.json 2> $O/r2a_ref.err; echo "ref rc=$?"
CMD="timeout 300 python bench.py --steps 1 --warmup 1 --settle-steps 0 --layers 4 --no-cpu --profile-steps 0 --no-graph --ttft-runs 1"
$CMD > $O/r2a_plain.log 2>&1 && \
ncu --metrics gpu__time_duration.sum --clock-control none -c 1200 --csv --log-file $O/launches_r2a.csv $CMD > $O/r2a_ncu_launch.log 2>&1
echo "launch list rc=$?"
# the second eager decode step = the launches between the 2nd and the 3rd sample_kernel (1st = the batch prefill)
read SKIP CNT < <(python - <<'P'
import csv
rows=[r for r in csv.reader(l for l in open("gpurun_out/launches_r2a.csv") if not l.startswith("=="))]
hdr=rows[0]; ix=hdr.index("Kernel Name"); data=[r for r in rows[1:] if len(r)==len(hdr)]
s=[i for i,r in enumerate(data) if r[ix].startswith("sample_kernel")]
print(s[1]+1, s[2]-s[1])
P
)
echo "decode step: skip $SKIP count $CNT"
$CMD > $O/r2a_plain2.log 2>&1 && \
ncu --set full --clock-control none --kernel-name-base demangled -s $SKIP -c $CNT -f -o $O/prof_r2a_step $CMD > $O/r2a_ncu_full.log 2>&1
echo "ncu full rc=$?"
ls -la $O/prof_r2a_step.ncu-rep
