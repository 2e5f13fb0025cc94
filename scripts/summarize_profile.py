#!/usr/bin/env python
"""Turn gpurun_out/ ncu artefacts into small, tracked summaries under profiles/.

    python scripts/summarize_profile.py <tag> <round-prefix> [index of the sample_kernel launch that ends the decode step]

reads gpurun_out/launches_<tag>.csv (ncu --metrics gpu__time_duration.sum launch list) and
gpurun_out/prof_<tag>.ncu-rep (ncu --set full of the dominant kernel), writes
profiles/<round-prefix>_launches.csv, profiles/<round-prefix>_step_shares.txt, profiles/<round-prefix>_<tag>_full.txt."""
import collections
import csv
import os
import subprocess
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
tag, pref = sys.argv[1], sys.argv[2]
out = os.path.join(ROOT, "profiles")
os.makedirs(out, exist_ok=True)

lines = [l for l in open(os.path.join(ROOT, "gpurun_out", f"launches_{tag}.csv")) if not l.startswith("==")]
rows = list(csv.reader(lines))
hdr, data = rows[0], [r for r in rows[1:] if len(r) == len(rows[0])]
ix = {h: i for i, h in enumerate(hdr)}
with open(os.path.join(out, f"{pref}_launches.csv"), "w", newline="") as f:
    w = csv.writer(f)
    w.writerow(["id", "kernel", "grid", "block", "duration_ns"])
    for r in data:
        w.writerow([r[ix["ID"]], r[ix["Kernel Name"]].split("(")[0], r[ix["Grid Size"]], r[ix["Block Size"]], r[ix["Metric Value"]]])

# the last decode step = everything after the last-but-one sample_kernel
names = [r[ix["Kernel Name"]] for r in data]
samples = [i for i, n in enumerate(names) if n.startswith("sample_kernel")]
# which sample_kernel ends the decode step to summarise (bench.py's order: batch prefill, W+K decode steps, then TTFT runs)
k = int(sys.argv[3]) if len(sys.argv) > 3 else len(samples) - 1
step = data[samples[k - 1] + 1: samples[k] + 1] if len(samples) >= 2 else data
tot = sum(float(r[ix["Metric Value"]]) for r in step)
agg = collections.OrderedDict()
for r in step:
    k = r[ix["Kernel Name"]].split("(")[0] + " grid" + r[ix["Grid Size"]]
    a = agg.setdefault(k, [0, 0.0])
    a[0] += 1; a[1] += float(r[ix["Metric Value"]])
with open(os.path.join(out, f"{pref}_step_shares.txt"), "w") as f:
    f.write(f"# one decode step under ncu (cold cache, serialised launches: compare SHARES, not absolutes); total {tot / 1e3:.1f} us\n")
    f.write(f"{'kernel':70s} {'launches':>8s} {'us':>10s} {'share':>7s}\n")
    for k, (n, ns) in sorted(agg.items(), key=lambda kv: -kv[1][1]):
        f.write(f"{k[:70]:70s} {n:8d} {ns / 1e3:10.1f} {ns / tot:7.1%}\n")

rep = os.path.join(ROOT, "gpurun_out", f"prof_{tag}.ncu-rep")
if os.path.exists(rep):
    raw = subprocess.run(["ncu", "-i", rep, "--page", "raw", "--csv"], capture_output=True, text=True).stdout
    r = list(csv.reader(raw.splitlines()))
    h, u = r[0], r[1]
    want = ["Kernel Name", "launch__grid_size", "launch__block_size", "launch__registers_per_thread",
            "launch__shared_mem_per_block_dynamic", "gpu__time_duration.sum", "dram__bytes_read.sum", "dram__bytes_write.sum",
            "dram__bytes_read.sum.per_second", "gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed",
            "dram__cycles_active.avg.pct_of_peak_sustained_elapsed", "lts__t_sector_hit_rate.pct",
            "sm__pipe_tensor_cycles_active.avg.pct_of_peak_sustained_active", "sm__warps_active.avg.pct_of_peak_sustained_active",
            "sm__throughput.avg.pct_of_peak_sustained_elapsed", "sm__cycles_active.avg", "sm__cycles_elapsed.max",
            "smsp__inst_executed.sum"]
    with open(os.path.join(out, f"{pref}_{tag}_full.txt"), "w") as f:
        f.write(f"# ncu --set full --clock-control none, kernel launches captured: {len(r) - 2}\n")
        for row in r[2:]:
            for wname in want:
                if wname in h:
                    i = h.index(wname)
                    f.write(f"{wname} = {row[i]} {u[i]}\n")
            f.write("--\n")
print("ok")
