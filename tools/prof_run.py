#!/usr/bin/env python
"""Small driver for ncu captures: a few word-count jobs over C2 text of --mib MiB."""
import argparse
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import locust_b200
from locust_b200.textgen import TextGen

ap = argparse.ArgumentParser()
ap.add_argument("--mib", type=int, default=256)
ap.add_argument("--runs", type=int, default=2)
ap.add_argument("--config", default="C2")
ap.add_argument("--table-log2", type=int, default=17)
ap.add_argument("--flags", type=int, default=0)
a = ap.parse_args()
n = a.mib << 20
g = TextGen.config(a.config, total_bytes=n, device=0)
d = torch.empty(n, dtype=torch.uint8, device="cuda")
g.fill_device(d)
torch.cuda.synchronize()
wc = locust_b200.WordCount(device=0, table_log2=a.table_log2, flags=a.flags)
for _ in range(a.runs):
    wc.run(d)
    print("stage ms", wc.stage_ms(), wc.stats(), "GB/s map", n / wc.stage_ms()[0] / 1e6)
wc.close()
