#!/usr/bin/env python
"""Wall-clock breakdown of one distributed word-count step (run under torchrun)."""
import os, sys, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch, torch.distributed as dist
import locust_b200
from locust_b200.textgen import TextGen
from locust_b200.shuffle import GpuBackend

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
    for it in range(6):
        if it == 1: T.clear()
        dist.barrier(); torch.cuda.synchronize(); t = time.perf_counter()
        b.reset(); b.map(text); t = tick("map", t)
        for phase, (nparts, dests) in enumerate(((world, list(range(world))), (1, [0]))):
            sizes = b.shuffle_sizes(nparts); t = tick("sizes%d" % phase, t)
            send = [0] * world
            for i, s in enumerate(sizes): send[dests[i]] += s
            ts, tr = b.sizes_tensor(send), b.sizes_tensor([0] * world)
            dist.all_to_all_single(tr, ts); recv = [int(x) for x in tr.tolist()]; t = tick("xsizes%d" % phase, t)
            offs, acc = [], 0
            for s in sizes: offs.append(acc); acc += s
            sb = b.alloc(acc); b.shuffle_pack(nparts, sb, offs); t = tick("pack%d" % phase, t)
            rb = b.alloc(sum(recv))
            dist.all_to_all_single(rb[:sum(recv)], sb[:acc], output_split_sizes=recv, input_split_sizes=send); t = tick("a2a%d" % phase, t)
            b.begin_merge(); off = 0
            for s in recv:
                if s: b.merge(rb, off, s)
                off += s
            t = tick("merge%d" % phase, t)
        b.process(); b.reduce(); t = tick("process+reduce", t)
if rank == 0:
    tot = sum(T.values())
    for k, v in T.items(): print("%-16s %8.3f ms" % (k, v / 5 * 1e3))
    print("total %.3f ms" % (tot / 5 * 1e3))
dist.destroy_process_group()
