#!/usr/bin/env python
"""Wall-clock breakdown (synchronised after every phase, so sums exceed the pipelined step) of one
ordered distributed word-count step over peer memory (run under torchrun)."""
import os, sys, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch, torch.distributed as dist
import locust_b200
from locust_b200.textgen import TextGen
from locust_b200.shuffle import GpuBackend, PeerExchange

rank, world, lr = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"]), int(os.environ["LOCAL_RANK"])
torch.cuda.set_device(lr); dev = torch.device("cuda", lr)
dist.init_process_group("nccl", device_id=dev)
per = 2 << 30
g = TextGen.config("C3", total_bytes=per * world, device=lr)
text = torch.empty(per, dtype=torch.uint8, device=dev); g.fill_device(text, byte_offset=rank * per); torch.cuda.synchronize()
st = torch.cuda.Stream(device=dev)
wc = locust_b200.WordCount(device=lr, table_log2=22, stream=st.cuda_stream)
b = GpuBackend(wc, dev)
T = {}
def tick(name, t0):
    torch.cuda.synchronize(); t = time.perf_counter(); T[name] = T.get(name, 0.0) + (t - t0); return t
with torch.cuda.stream(st):
    px = PeerExchange(wc, dist, dev, 1 << 22)
    for it in range(6):
        if it == 1: T.clear()
        dist.barrier(); torch.cuda.synchronize(); t = time.perf_counter()
        b.reset(); b.map(text); t = tick("map", t)
        hist = b.histogram(); t = tick("histogram", t)
        px.share_histogram(hist); t = tick("put histogram", t)
        px.barrier(); t = tick("barrier0", t)
        wc.partition_by_range_sum(px.hist_slots(), world, world); t = tick("range map", t)
        dst = [px.bufs[d][0] + rank * px.slot_bytes for d in range(world)]
        wc.shuffle_pack_peers(dst, px.slot_bytes, px.short_cap); t = tick("pack_peers", t)
        px.barrier(); t = tick("barrier", t)
        wc.shuffle_begin_merge()
        for s in range(world): wc.shuffle_merge(px.mine[0] + s * px.slot_bytes, px.slot_bytes)
        t = tick("merge", t)
        b.process(); t = tick("process", t)
        b.reduce(); t = tick("reduce", t)
        wc.result_part_pack(px.bufs[0][2] + rank * px.result_slot); t = tick("pack result", t)
        px.barrier(); t = tick("barrier2", t)
        if rank == 0: wc.result_adopt_slots(px.mine[2], px.result_slot, world)
        t = tick("adopt", t)
if rank == 0:
    tot = sum(T.values())
    for k, v in T.items(): print("%-20s %8.3f ms" % (k, v / 5 * 1e3))
    print("total %.3f ms" % (tot / 5 * 1e3))
px.close()
dist.destroy_process_group()
