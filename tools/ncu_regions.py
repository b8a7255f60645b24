#!/usr/bin/env python
"""Sums executed warp instructions of an ncu source dump (cuda,sass) by function region of map_kernel.cu/common.cuh."""
import csv, sys, re
src = {}
def load(path):
    try:
        return open(path).read().split("\n")
    except Exception:
        return []
files = {"map_kernel.cu": load("locust_b200/csrc/map_kernel.cu"), "common.cuh": load("locust_b200/csrc/common.cuh")}
def region(f, ln):
    lines = files.get(f)
    if not lines: return f
    # nearest preceding function-ish header
    for i in range(ln - 1, -1, -1):
        m = re.match(r"^(?:template.*)?(?:LC_HD|LC_D|__device__|__global__|static|inline).*?(\w+)\s*\(", lines[i])
        if m: return f + ":" + m.group(1)
    return f
tot = 0; agg = {}; lanes = {}
cur = None
for r in csv.reader(open(sys.argv[1])):
    if len(r) >= 2 and r[0] == "File Path": cur = r[1].split("/")[-1]; continue
    if len(r) > 10 and r[0].strip().isdigit() and r[2] == "-":
        try: ie, te = int(r[7]), int(r[8])
        except ValueError: continue
        k = region(cur, int(r[0]))
        agg[k] = agg.get(k, 0) + ie; lanes[k] = lanes.get(k, 0) + te; tot += ie
rows = float(sys.argv[2]) if len(sys.argv) > 2 else 1.0
for k, v in sorted(agg.items(), key=lambda x: -x[1]):
    print("%5.1f%%  %7.1f instr/row  lanes=%4.1f  %s" % (100.0 * v / tot, v / rows, lanes[k] / max(v, 1), k))
print("total", tot, "per row", tot / rows)
