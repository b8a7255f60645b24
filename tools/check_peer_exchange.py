#!/usr/bin/env python
"""Run under torchrun on >= 2 GPUs: every way of running a distributed job -- hash parts gathered
on rank 0 or key-range parts sorted in place ("ordered"), over NCCL or over IPC-mapped peer
buffers (NVLink), with the flag barrier or the all-reduce barrier -- must give the oracle's result
on rank 0."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch, torch.distributed as dist
import locust_b200
from locust_b200.textgen import TextGen
from locust_b200.shuffle import DistributedWordCount, GpuBackend, PeerExchange

rank, world, lr = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"]), int(os.environ["LOCAL_RANK"])
torch.cuda.set_device(lr); dev = torch.device("cuda", lr)
dist.init_process_group("nccl", device_id=dev)
per = 32 << 20
g = TextGen.config("C3", total_bytes=per * world, device=lr, vocab=300000)
text = torch.empty(per, dtype=torch.uint8, device=dev); g.fill_device(text, byte_offset=rank * per); torch.cuda.synchronize()
# a few long tokens so the long-record area is exercised too
extra = ("\n" + " ".join("longtoken%02d%s" % (i % 7, "x" * (20 + i % 5)) for i in range(50 * (rank + 1))) + "\n").encode()
text = torch.cat([text, torch.frombuffer(bytearray(extra), dtype=torch.uint8).to(dev)])
st = torch.cuda.Stream(device=dev)
log2 = 20
wc = locust_b200.WordCount(device=lr, table_log2=log2, stream=st.cuda_stream)
results = {}
with torch.cuda.stream(st):
    px = PeerExchange(wc, dist, dev, 1 << log2)
    px_nccl_barrier = PeerExchange(wc, dist, dev, 1 << log2, flag_barrier=False)
    px_counted = PeerExchange(wc, dist, dev, 1 << log2, counted=True)        # slots sized from a count pass, grown on demand
    for name, peers, ordered in (("gather nccl", None, False), ("gather peers", px, False), ("gather peers again", px, False),
                                 ("ordered nccl", None, True), ("ordered peers", px, True), ("ordered peers again", px, True),
                                 ("ordered peers, all-reduce barrier", px_nccl_barrier, True),
                                 ("ordered peers, counted slots", px_counted, True), ("ordered peers, counted slots again", px_counted, True),
                                 ("gather peers, counted slots", px_counted, False)):
        job = DistributedWordCount(GpuBackend(wc, dev), dist, peers=peers, ordered=ordered)
        job.run(text)
        results[name] = (wc.result()[:2], ordered)
    px.close()
    px_nccl_barrier.close()
    px_counted.close()
gathered = [None] * world
dist.all_gather_object(gathered, bytes(text.cpu().numpy()))
if rank == 0:
    from oracle import pyoracle
    ref = pyoracle.wordcount(b"".join(gathered), mode="fast", formats=False)
    for name, ((keys, counts), _) in results.items():
        assert keys == ref.keys and counts == ref.counts, name
        print("%s: %d keys match the oracle" % (name, len(keys)))
else:
    for name, ((keys, counts), ordered) in results.items():
        if ordered:          # this rank kept the slice of its key range
            assert keys == sorted(keys) and len(keys) > 0, name
        else:
            assert keys == [] and counts == [], name
dist.destroy_process_group()
print("rank %d ok" % rank)
