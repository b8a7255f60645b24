#!/usr/bin/env python
"""Turns gpurun_out/launches.csv (ncu launch list of `bench.py`) and a full ncu capture of the map
kernel into the text summaries committed under profiles/.
usage: summarize_profiles.py <launches.csv> <capture.ncu-rep> <out_prefix> [input_bytes_of_capture]"""
import collections
import csv
import json
import re
import subprocess
import sys

launches, rep, prefix = sys.argv[1], sys.argv[2], sys.argv[3]
cap_bytes = int(sys.argv[4]) if len(sys.argv) > 4 else None

rows = [r for r in csv.reader(open(launches)) if len(r) > 5]
hdr, agg, n = None, collections.OrderedDict(), 0
for r in rows:
    if r[0] == "ID":
        hdr = r
        continue
    if hdr is None or not r[0].isdigit():
        continue
    d = dict(zip(hdr, r))
    name = re.sub(r"\(.*", "", d["Kernel Name"])
    v = float(d["Metric Value"].replace(",", ""))
    v = v * 1000 if d["Metric Unit"] == "ms" else v / 1000 if d["Metric Unit"] == "ns" else v
    a = agg.setdefault(name, [0, 0.0])
    a[0] += 1
    a[1] += v
    n += 1
tot = sum(a[1] for k, a in agg.items() if "gen_kernel" not in k)
with open(prefix + "_bench_launches.txt", "w") as f:
    f.write("ncu --metrics gpu__time_duration.sum --clock-control none, python bench.py --steps 2 --warmup 1 --no-cpu-baseline\n")
    f.write("(cold-cache, serialised launches: compare shares, not absolutes; gen_kernel is the text generator, outside the timed region)\n")
    f.write("launches %d, kernel time without the generator %.1f us\n" % (n, tot))
    for k, (c, t) in sorted(agg.items(), key=lambda x: -x[1][1]):
        f.write("%6.2f%%  %10.1f us  x%4d  %s\n" % (100 * t / tot if "gen_kernel" not in k else 0.0, t, c, k))

raw = subprocess.run(["ncu", "-i", rep, "--page", "raw", "--csv"], capture_output=True, text=True).stdout
rr = list(csv.reader(raw.splitlines()))
d = dict(zip(rr[0], zip(rr[1], rr[2])))
keys = ["gpu__time_duration.sum", "dram__bytes_read.sum", "dram__bytes_write.sum", "smsp__inst_executed.sum",
        "smsp__issue_active.avg.pct_of_peak_sustained_active", "sm__warps_active.avg.pct_of_peak_sustained_active",
        "smsp__thread_inst_executed_per_inst_executed.ratio", "launch__registers_per_thread",
        "launch__occupancy_limit_shared_mem", "sm__inst_executed_pipe_alu.avg.pct_of_peak_sustained_active",
        "sm__inst_executed_pipe_fma.avg.pct_of_peak_sustained_active",
        "sm__inst_executed_pipe_lsu.avg.pct_of_peak_sustained_active",
        "sm__inst_executed_pipe_xu.avg.pct_of_peak_sustained_active", "gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed",
        "l1tex__data_bank_conflicts_pipe_lsu_mem_shared.sum", "l1tex__data_pipe_lsu_wavefronts_mem_shared.sum"]
with open(prefix + "_map_kernel_ncu.txt", "w") as f:
    f.write("ncu --set full --clock-control none, tokenize_combine_kernel, %s\n" % rep)
    for k in keys + sorted(x for x in d if "issue_stalled" in x and "per_issue_active" in x):
        if k in d:
            f.write("%-90s %-10s %s\n" % (k, d[k][0], d[k][1]))


def num(k):
    u, v = d[k]
    v = float(v.replace(",", ""))
    return v * {"byte": 1, "Kbyte": 1e3, "Mbyte": 1e6, "Gbyte": 1e9}.get(u, 1)


traffic = num("dram__bytes_read.sum") + num("dram__bytes_write.sum")
out = {"kernel": "tokenize_combine_kernel", "capture": rep, "dram_bytes_per_launch": traffic, "input_bytes_of_capture": cap_bytes}
with open(prefix + "_map_kernel_traffic.json", "w") as f:
    json.dump(out, f)
print(open(prefix + "_bench_launches.txt").read())
print(out)
