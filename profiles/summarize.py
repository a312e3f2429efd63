#!/usr/bin/env python
"""Turn gpurun_out/ ncu artefacts into the tracked summaries under profiles/.

    python profiles/summarize.py r01            # reads gpurun_out/launches.csv and gpurun_out/prof_*.ncu-rep
"""
import collections
import csv
import re
import subprocess
import sys

METRICS = ["gpu__time_duration.sum", "dram__bytes_read.sum", "dram__bytes_write.sum",
           "gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed",
           "sm__pipe_tensor_cycles_active.avg.pct_of_peak_sustained_active",
           "sm__warps_active.avg.pct_of_peak_sustained_active", "launch__registers_per_thread", "launch__grid_size",
           "launch__block_size", "launch__waves_per_multiprocessor", "launch__shared_mem_per_block_dynamic",
           "launch__occupancy_limit_shared_mem", "lts__t_sector_hit_rate.pct", "smsp__cycles_active.avg"]


def launches(tag, src="launches", suffix="launches_summary"):
    with open(f"gpurun_out/{src}.csv") as f:
        lines = [l for l in f if not l.startswith("==")]
    agg, n = collections.OrderedDict(), 0
    for row in csv.DictReader(lines):
        if row.get("Metric Name") != "gpu__time_duration.sum":
            continue
        v = float(row["Metric Value"].replace(",", ""))
        v = v / 1e3 if row["Metric Unit"] == "ns" else v * 1e3 if row["Metric Unit"] == "ms" else v
        key = re.sub(r"\(.*", "", row["Kernel Name"]).strip()
        a = agg.setdefault(key, [0, 0.0])
        a[0] += 1
        a[1] += v
        n += 1
    tot = sum(a[1] for a in agg.values())
    with open(f"profiles/{tag}_{suffix}.txt", "w") as out:
        out.write(f"# ncu --metrics gpu__time_duration.sum --clock-control none (cold-cache, serialised): {n} launches, "
                  f"{tot:.1f} us total\n# share   launches   avg_us   kernel\n")
        for k, (c, t) in sorted(agg.items(), key=lambda x: -x[1][1]):
            out.write(f"{t / tot * 100:6.2f}%  {c:5d}  {t / c:8.2f}  {k}\n")
    print(open(f"profiles/{tag}_{suffix}.txt").read())


def full(tag, rep):
    raw = subprocess.run(["ncu", "-i", f"gpurun_out/{rep}.ncu-rep", "--page", "raw", "--csv"], capture_output=True,
                         text=True).stdout
    r = list(csv.reader(raw.splitlines()))
    hdr, units = r[0], r[1]
    with open(f"profiles/{tag}_{rep}.txt", "w") as out:
        out.write(f"# ncu --set full --clock-control none, report {rep}.ncu-rep; one block per captured launch\n")
        for row in r[2:]:
            out.write(f"\n{row[hdr.index('Kernel Name')]}\n")
            for m in METRICS:
                if m in hdr:
                    out.write(f"  {m:70s} {row[hdr.index(m)]} {units[hdr.index(m)]}\n")
    print(open(f"profiles/{tag}_{rep}.txt").read())


if __name__ == "__main__":
    tag = sys.argv[1]
    launches(tag)
    import os
    if os.path.exists("gpurun_out/launches_prefill.csv"):
        launches(tag, "launches_prefill", "launches_prefill_summary")
    for rep in sys.argv[2:] or ["prof_gemm_decode", "prof_attn_decode"]:
        full(tag, rep)
