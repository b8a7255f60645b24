#!/usr/bin/env python
"""Development aid: runs the MAP variants (LOCUST_MAP_VARIANT) over a ladder of inputs, each case in its own
process under a timeout, and compares with the oracle.  Prints one line per case."""
import os, subprocess, sys, time
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
CHILD = r'''
import sys, os, time
sys.path.insert(0, %r)
import numpy as np, torch, locust_b200
from oracle import parity
from locust_b200.textgen import TextGen
kind, n = sys.argv[1], int(sys.argv[2])
if kind == "word":
    data = np.frombuffer(b"word", dtype=np.uint8)
elif kind == "hamlet":
    data = np.frombuffer(open(os.path.join(%r, "tests/golden/hamlet.txt"), "rb").read() * max(1, n), dtype=np.uint8)
else:
    g = TextGen.config(kind, total_bytes=max(n, 4096))
    data = g.host(max(n, 4096))[:n]
    k = len(data)
    while k > 0 and chr(data[k-1]) not in " ,.\n": k -= 1
    data = data[:k] if k else data
dev = torch.from_numpy(data.copy()).cuda()
wc = locust_b200.WordCount(device=0, table_log2=int(sys.argv[3]))
t0 = time.time()
for _ in range(2): wc.run(dev)
ms = wc.stage_ms()
why = parity.same_result(wc.result_arrays(), parity.oracle_arrays(data))
print("ok" if why is None else "MISMATCH: " + why, "map ms %%.3f" %% ms[0], "GB/s %%.1f" %% (len(data) / ms[0] / 1e6), flush=True)
''' % (ROOT, ROOT)
cases = [("word", 4, 14), ("C2", 16, 14), ("C2", 1000, 14), ("C2", 1024, 14), ("C2", 1025, 14), ("C2", 4096, 14), ("hamlet", 1, 16), ("C2", 1 << 20, 17),
         ("C2", 16 << 20, 20), ("C3", 64 << 20, 21), ("C2", 256 << 20, 20), ("C5", 64 << 20, 21), ("C4", 32 << 20, 22), ("C2", 1 << 30, 20), ("C3", 1 << 30, 21)]
variants = sys.argv[1].split(",") if len(sys.argv) > 1 else ["3", "0"]
limit = int(sys.argv[2]) if len(sys.argv) > 2 else 40
if len(sys.argv) > 3 and sys.argv[3] == "big":
    cases = [("C2", 16 << 20, 20), ("C2", 256 << 20, 20), ("C2", 1 << 30, 20), ("C3", 1 << 30, 21)]
if len(sys.argv) > 3 and sys.argv[3] == "small":
    cases = [c for c in cases if c[1] <= (64 << 20)][:11]
for v in variants:
    bad = 0
    for kind, n, lg in cases:
        env = dict(os.environ, LOCUST_MAP_VARIANT=v)
        t0 = time.time()
        try:
            r = subprocess.run([sys.executable, "-c", CHILD, kind, str(n), str(lg)], env=env, capture_output=True, text=True, timeout=limit)
            out = (r.stdout.strip().splitlines() or ["rc=%d %s" % (r.returncode, r.stderr.strip()[-300:])])[-1]
        except subprocess.TimeoutExpired:
            out = "TIMEOUT"
            bad += 1
        print("variant %s %-6s %10d : %s (%.1f s)" % (v, kind, n, out, time.time() - t0), flush=True)
        if bad >= 2:
            print("variant %s: giving up after two timeouts" % v, flush=True)
            break
