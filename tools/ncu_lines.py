#!/usr/bin/env python
"""Ranks source lines of an `ncu --page source --csv --print-source cuda,sass` dump by executed instructions."""
import csv
import sys

tot = 0
rows = []
cur = None
for r in csv.reader(open(sys.argv[1])):
    if len(r) >= 2 and r[0] == "File Path":
        cur = r[1].split("/")[-1]
        continue
    if len(r) > 10 and r[0].strip().isdigit() and r[2] == "-":
        try:
            samp, ie, te = int(r[6]), int(r[7]), int(r[8])
        except ValueError:
            continue
        rows.append((ie, te, samp, cur, int(r[0]), r[1].strip()[:100]))
        tot += ie
rows.sort(reverse=True)
print("total warp instructions", tot)
for ie, te, samp, f, ln, src in rows[: int(sys.argv[2]) if len(sys.argv) > 2 else 30]:
    print("%5.1f%% lanes=%5.1f samp=%6d %s:%d %s" % (100.0 * ie / tot, te / max(ie, 1), samp, f, ln, src))
