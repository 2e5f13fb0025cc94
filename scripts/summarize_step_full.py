#!/usr/bin/env python
"""ncu --set full capture of ONE eager decode step (gpurun_out/prof_<tag>.ncu-rep, every kernel of the step) ->
profiles/<prefix>_decode_step_kernels_full.txt (one block per distinct kernel/grid: mean over its launches) and
profiles/roofline_traffic.json (DRAM bytes per launch of the dominant kernel, read by bench.py's `roofline.traffic`).

    python scripts/summarize_step_full.py <tag> <prefix> <batch>
"""
import collections
import csv
import json
import os
import subprocess
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
tag, pref, batch = sys.argv[1], sys.argv[2], int(sys.argv[3])
rep = os.path.join(ROOT, "gpurun_out", f"prof_{tag}.ncu-rep")
raw = subprocess.run(["ncu", "-i", rep, "--page", "raw", "--csv"], capture_output=True, text=True).stdout
r = list(csv.reader(raw.splitlines()))
h, u = r[0], r[1]
want = ["gpu__time_duration.sum", "dram__bytes_read.sum", "dram__bytes_write.sum",
        "gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed", "lts__t_sector_hit_rate.pct",
        "sm__pipe_tensor_cycles_active.avg.pct_of_peak_sustained_active", "sm__warps_active.avg.pct_of_peak_sustained_active",
        "sm__throughput.avg.pct_of_peak_sustained_elapsed", "launch__registers_per_thread",
        "launch__shared_mem_per_block_dynamic", "launch__occupancy_limit_shared_mem", "launch__waves_per_multiprocessor",
        "smsp__inst_executed.sum", "l1tex__data_pipe_lsu_wavefronts_mem_shared.sum", "lts__t_bytes.sum"]
want = [w for w in want if w in h]
scale = {"Gbyte": 1e9, "Mbyte": 1e6, "Kbyte": 1e3, "byte": 1.0}
groups = collections.OrderedDict()
for row in r[2:]:
    key = (row[h.index("Kernel Name")].split("(")[0], row[h.index("Grid Size")], row[h.index("launch__block_size")])
    groups.setdefault(key, []).append(row)
out = os.path.join(ROOT, "profiles", f"{pref}_decode_step_kernels_full.txt")
traffic = {}
with open(out, "w") as f:
    f.write(f"# ncu --set full --clock-control none of one eager decode step (Llama-3-8B shapes, batch {batch}, 4-layer run: per-kernel\n"
            f"# numbers do not depend on depth); {len(r) - 2} launches, mean per distinct (kernel, grid).  Isolated, serialised, cold-cache\n"
            f"# launches: inside the CUDA graph programmatic dependent launch overlaps neighbours, so compare shares and counters, not sums.\n")
    for (name, grid, block), rows in groups.items():
        f.write(f"\n== {name}  grid {grid} block {block}  launches {len(rows)}\n")
        for w in want:
            i = h.index(w)
            vals = [float(x[i].replace(",", "")) for x in rows if x[i] not in ("", "n/a")]
            if vals:
                f.write(f"{w} = {sum(vals) / len(vals):.6g} {u[i]}\n")
        if name.startswith("void gemm_tcgen05_kernel<32, 1>"):
            ir, iw = h.index("dram__bytes_read.sum"), h.index("dram__bytes_write.sum")
            rd = sum(float(x[ir]) * scale[u[ir]] for x in rows) / len(rows)
            wr = sum(float(x[iw]) * scale[u[iw]] for x in rows) / len(rows)
            traffic[f"gateup_8b_batch{batch}"] = {"dram_bytes_read": int(rd), "dram_bytes_write": int(wr), "launches": len(rows),
                                                  "source": f"profiles/{pref}_decode_step_kernels_full.txt (ncu --set full, gemm_tcgen05_kernel<32,1> grid {grid})"}
tp = os.path.join(ROOT, "profiles", "roofline_traffic.json")
old = json.load(open(tp)) if os.path.exists(tp) else {}
old.update(traffic)
json.dump(old, open(tp, "w"), indent=1)
print("ok", out, traffic)
