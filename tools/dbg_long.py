#!/usr/bin/env python
"""Development aid: MAP rate on text whose tokens all have 17..64 bytes, for two distributions of the words."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np, torch, locust_b200
rng = np.random.default_rng(11)
nwords = 20_000
lens = rng.integers(17, 65, size=nwords)
vocab = [bytes(rng.integers(97, 123, size=int(n), dtype=np.uint8)) for n in lens]
for kind in ("zipf1.3", "uniform"):
    picks = (rng.zipf(1.3, size=1_800_000) % nwords) if kind == "zipf1.3" else rng.integers(0, nwords, size=1_800_000)
    text = b" ".join(vocab[int(i)] for i in picks)[: 64 << 20]
    data = np.frombuffer(text[: text.rfind(b" ")] + b"\n", dtype=np.uint8)
    dev = torch.from_numpy(data.copy()).cuda()
    wc = locust_b200.WordCount(device=0)
    for _ in range(3):
        wc.run(dev)
    ms = wc.stage_ms()
    print(kind, "bytes", len(data), "stage ms", ms, "MAP GB/s %.1f" % (len(data) / ms[0] / 1e6), wc.stats(), flush=True)
    wc.close()
