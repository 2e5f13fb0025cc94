#!/usr/bin/env python
"""Any ncu --set full capture (gpurun_out/prof_<tag>.ncu-rep) -> profiles/<out>.txt: one block per distinct (kernel, grid),
metrics averaged over its launches.

    python scripts/summarize_full.py <tag> <out-name> "<one-line description of the command that was profiled>"
"""
import collections
import csv
import os
import subprocess
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
tag, name, what = sys.argv[1], sys.argv[2], sys.argv[3]
rep = os.path.join(ROOT, "gpurun_out", f"prof_{tag}.ncu-rep")
raw = subprocess.run(["ncu", "-i", rep, "--page", "raw", "--csv"], capture_output=True, text=True).stdout
r = list(csv.reader(raw.splitlines()))
h, u = r[0], r[1]
want = ["gpu__time_duration.sum", "dram__bytes_read.sum", "dram__bytes_write.sum",
        "gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed", "lts__t_sector_hit_rate.pct",
        "sm__pipe_tensor_cycles_active.avg.pct_of_peak_sustained_active", "sm__warps_active.avg.pct_of_peak_sustained_active",
        "sm__throughput.avg.pct_of_peak_sustained_elapsed", "sm__inst_executed_pipe_xu.avg.pct_of_peak_sustained_active",
        "launch__registers_per_thread", "launch__shared_mem_per_block_dynamic", "launch__occupancy_limit_shared_mem",
        "launch__occupancy_limit_registers", "launch__waves_per_multiprocessor", "smsp__inst_executed.sum"]
want = [w for w in want if w in h]
groups = collections.OrderedDict()
for row in r[2:]:
    key = (row[h.index("Kernel Name")].split("(")[0], row[h.index("Grid Size")], row[h.index("Block Size")])
    groups.setdefault(key, []).append(row)
out = os.path.join(ROOT, "profiles", name + ".txt")
with open(out, "w") as f:
    f.write(f"# ncu --set full --clock-control none: {what}\n# {len(r) - 2} launches, mean per distinct (kernel, grid); isolated, "
            f"serialised launches -- compare counters and shares, not sums\n")
    for (kname, grid, block), rows in groups.items():
        f.write(f"\n== {kname}  grid {grid} block {block}  launches {len(rows)}\n")
        for w in want:
            i = h.index(w)
            vals = [float(x[i].replace(",", "")) for x in rows if x[i] not in ("", "n/a")]
            if vals:
                f.write(f"{w} = {sum(vals) / len(vals):.6g} {u[i]}\n")
print("ok", out, len(groups), "kernels")
