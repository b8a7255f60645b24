#!/usr/bin/env python
"""Lists the SASS of an `ncu --page source --csv --print-source cuda,sass` dump in address order with
executed warp instructions per unit (argv[2], e.g. rows of input) and the source line of each."""
import csv
import sys

unit = float(sys.argv[2]) if len(sys.argv) > 2 else 1.0
out, cur, line = [], None, None
for r in csv.reader(open(sys.argv[1])):
    if len(r) >= 2 and r[0] == "File Path":
        cur = r[1].split("/")[-1]
        continue
    if len(r) > 10 and r[0].strip().isdigit() and r[2] == "-":
        line = "%s:%s" % (cur, r[0].strip())
        continue
    if len(r) > 10 and r[0] == "" and r[2].startswith("0x"):
        try:
            out.append((int(r[2], 16), r[3].strip(), int(r[7]), int(r[8]), int(r[6]), line))
        except ValueError:
            pass
out.sort()
base = out[0][0] if out else 0
for a, s, ie, te, samp, ln in out:
    print("%05x %7.2f l=%4.1f s=%5d  %-60s %s" % (a - base, ie / unit, te / max(ie, 1), samp, s[:60], ln))
