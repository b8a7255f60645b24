#!/usr/bin/env python
"""Static view of a kernel's loops from its SASS (no GPU needed): every backward branch closes a loop; prints each
loop's address range and instruction count, innermost first, so the effect of a source change on the hot loops of
the MAP kernel can be read off before any GPU time is spent.

  python tools/sass_loops.py locust_b200/csrc/build/map_kernel.o 'tokenize_combine_kernelILb1'"""
import re, subprocess, sys
obj, pat = sys.argv[1], sys.argv[2]
txt = subprocess.run(["cuobjdump", "-sass", obj], capture_output=True, text=True).stdout
ins, on = [], False
for line in txt.split("\n"):
    if "Function :" in line:
        on = pat in line
        continue
    if not on: continue
    m = re.match(r"\s+/\*([0-9a-f]{4,})\*/\s+(.*?);", line)
    if m: ins.append((int(m.group(1), 16), m.group(2).strip()))
addr_ix = {a: i for i, (a, _) in enumerate(ins)}
print("instructions:", len(ins))
loops = []
for i, (a, s) in enumerate(ins):
    m = re.search(r"\bBRA(?:\.U)?(?:\.DIV)?\s+(?:[!\w]+,\s*)?`?\(?\.?L?_?x?_?\d*\)?\s*(0x[0-9a-f]+)", s)
    if "BRA" in s:
        m2 = re.search(r"0x([0-9a-f]+)\s*$", s)
        if m2:
            t = int(m2.group(1), 16)
            if t <= a and t in addr_ix and "BRA.DIV" not in s:
                loops.append((addr_ix[t], i))
loops.sort(key=lambda x: x[1] - x[0])
for lo, hi in loops:
    body = [s for _, s in ins[lo:hi + 1]]
    ops = {}
    for s in body:
        op = s.split()[1] if s.startswith("@") else s.split()[0]
        op = op.split(".")[0]
        ops[op] = ops.get(op, 0) + 1
    top = sorted(ops.items(), key=lambda kv: -kv[1])[:8]
    print("loop %5d..%5d  %4d instr  %s" % (lo, hi, hi - lo + 1, " ".join("%s:%d" % kv for kv in top)))
if len(sys.argv) > 3:
    lo, hi = int(sys.argv[3]), int(sys.argv[4])
    for k in range(lo, hi + 1): print("%5d  %s" % (k, ins[k][1]))
