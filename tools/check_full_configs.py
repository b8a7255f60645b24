#!/usr/bin/env python
"""BASELINE.json configs 3-5 at full size on N GPUs (run under torchrun): the distributed job is
checked through size-independent properties --
  * keys on rank 0 strictly ascending (sorted, unique),
  * sum of counts == sum over ranks of the tokens each rank counts in its own shard alone,
  * vals == exclusive prefix sums of counts,
  * for a sample of keys: global count == sum over ranks of the single-GPU counts of that key,
and timed.  Prints one JSON line per config on rank 0."""
import argparse, json, os, sys, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
import torch, torch.distributed as dist
import locust_b200
from locust_b200.textgen import TextGen, CONFIGS
from locust_b200.shuffle import DistributedWordCount, GpuBackend, PeerExchange

ap = argparse.ArgumentParser()
ap.add_argument("--configs", default="C3,C5,C4")
ap.add_argument("--gib", type=float, default=0.0, help="override the total size (GiB)")
ap.add_argument("--final", default="ordered", choices=["ordered", "gather"])
a = ap.parse_args()
rank, world, lr = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"]), int(os.environ["LOCAL_RANK"])
torch.cuda.set_device(lr); dev = torch.device("cuda", lr)
dist.init_process_group("nccl", device_id=dev)
st = torch.cuda.Stream(device=dev)


def packed(blob, offs):
    offs = offs.astype(np.int64); lens = np.diff(offs)
    keys = np.zeros((len(lens), 16), dtype=np.uint8)
    idx = np.repeat(np.arange(len(lens)), lens)
    keys[idx, np.arange(len(blob)) - np.repeat(offs[:-1], lens)] = blob
    return keys.view(">u8").reshape(-1, 2).astype(np.uint64)


for name in a.configs.split(","):
    total = int(a.gib * (1 << 30)) if a.gib else CONFIGS[name][5]
    per = total // world // 4096 * 4096
    log2 = {"C3": 22, "C5": 22, "C4": 27}[name]
    g = TextGen.config(name, total_bytes=per * world, device=lr)
    text = torch.empty(per, dtype=torch.uint8, device=dev); g.fill_device(text, byte_offset=rank * per); torch.cuda.synchronize()
    wc = locust_b200.WordCount(device=lr, table_log2=log2, stream=st.cuda_stream)
    with torch.cuda.stream(st):
        # each rank alone on its shard: token count and a sample of its keys
        wc.run(text)
        local_tokens = wc.stats()["tokens"]
        lb, lo, lc, _ = wc.result_arrays()
        ordered = a.final == "ordered"
        peers = PeerExchange(wc, dist, dev, 1 << log2, ordered=ordered) if name != "C4" else None     # C4: slots of 4 GB each, use NCCL
        job = DistributedWordCount(GpuBackend(wc, dev), dist, peers=peers, ordered=ordered)
        job.run(text)                                                                 # warm-up
        dist.barrier(); torch.cuda.synchronize(); t0 = time.perf_counter()
        job.run(text)
        torch.cuda.synchronize(); dist.barrier(); dt = time.perf_counter() - t0
        blob, offs, counts, vals = wc.result_arrays()
        stats = wc.stats()
    tok = torch.tensor([local_tokens], dtype=torch.int64, device=dev); dist.all_reduce(tok)
    # sample: every rank contributes counts for rank 0's sample keys (short keys only: <= 16 bytes)
    ok = True
    if rank == 0:
        k = packed(blob, offs) if (np.diff(offs.astype(np.int64)).max() <= 16) else None
        sorted_unique = bool(((k[:-1, 0] < k[1:, 0]) | ((k[:-1, 0] == k[1:, 0]) & (k[:-1, 1] < k[1:, 1]))).all()) if k is not None else None
        sums_ok = int(counts.sum()) == int(tok.item()) == stats["tokens"]
        vals_ok = bool((vals == np.concatenate([[0], np.cumsum(counts)[:-1]]).astype(np.uint64)).all())
        pick = np.linspace(0, len(counts) - 1, 4096).astype(np.int64)
        sample = torch.from_numpy(k[pick].astype(np.int64)).to(dev) if k is not None else torch.zeros((4096, 2), dtype=torch.int64, device=dev)
    else:
        sample = torch.zeros((4096, 2), dtype=torch.int64, device=dev)
    dist.broadcast(sample, 0)
    lk = packed(lb, lo) if len(lc) else np.zeros((0, 2), dtype=np.uint64)
    s = sample.cpu().numpy().astype(np.uint64)
    # local lookup by binary search on the (sorted) local keys
    order = np.lexsort((lk[:, 1], lk[:, 0]))          # already sorted; keeps the code honest
    lk, lcs = lk[order], lc[order]
    comb = np.rec.fromarrays([lk[:, 0], lk[:, 1]], names="a,b")
    want = np.rec.fromarrays([s[:, 0], s[:, 1]], names="a,b")
    pos = np.searchsorted(comb, want)
    pos = np.minimum(pos, max(len(comb) - 1, 0))
    found = (comb[pos] == want) if len(comb) else np.zeros(len(want), dtype=bool)
    mine = np.where(found, lcs[pos] if len(comb) else 0, 0).astype(np.int64)
    t = torch.from_numpy(mine).to(dev); dist.all_reduce(t)
    if rank == 0:
        sample_ok = bool((t.cpu().numpy().astype(np.uint64) == counts[pick]).all()) if k is not None else None
        print(json.dumps({"config": name, "n_gpus": world, "bytes": per * world, "tokens": int(tok.item()),
                          "unique_keys": int(len(counts)), "seconds": dt, "gbs": per * world / dt / 1e9,
                          "sorted_unique": sorted_unique, "sum_counts_equals_tokens": sums_ok, "vals_are_prefix_sums": vals_ok,
                          "sampled_counts_add_up_over_ranks": sample_ok, "exchange": "peers" if peers else "nccl", "final_stage": a.final}), flush=True)
    if peers:
        peers.close()
    wc.close(); del text
    torch.cuda.empty_cache()
dist.destroy_process_group()
