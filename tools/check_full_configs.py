#!/usr/bin/env python
"""BASELINE.json configs 3-5 at full size on N GPUs (run under torchrun), checked against the CPU oracle
record for record (oracle/parity.py):
  * every rank alone on its shard == wc_oracle_fast on the same bytes (keys, counts, order, tokens),
  * the distributed result == the key-wise sum of the oracle's per-shard results, compared range by range
    on the rank that owns the range; global order strictly ascending; vals == prefix sums,
and timed.  Prints one JSON line per config on rank 0.  TEST/BENCH TOOL: the product never imports the oracle."""
import argparse, json, os, sys, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
import torch, torch.distributed as dist
import locust_b200
from locust_b200.textgen import TextGen, CONFIGS
from locust_b200.shuffle import DistributedWordCount, GpuBackend, PeerExchange
from oracle import parity


def run_config(name, total, world, rank, dev, stream, final="ordered", shuffle="peers", steps=2, check=True):
    """one BASELINE config over the process group: timing of `steps` jobs (device events, max over ranks) and the
    parity block; returns a dict on rank 0"""
    per = total // world // 4096 * 4096
    log2 = {"C2": 20, "C3": 22, "C5": 22, "C4": 27}[name]
    g = TextGen.config(name, total_bytes=per * world, device=dev.index)
    text = torch.empty(per, dtype=torch.uint8, device=dev)
    g.fill_device(text, byte_offset=rank * per)
    torch.cuda.synchronize()
    wc = locust_b200.WordCount(device=dev.index, table_log2=log2, stream=stream.cuda_stream)
    peers, out = None, None
    try:
        with torch.cuda.stream(stream):
            ordered = final == "ordered"
            if shuffle == "peers" and world > 1:
                peers = PeerExchange(wc, dist, dev, 1 << log2, ordered=ordered)
            job = DistributedWordCount(GpuBackend(wc, dev), dist, peers=peers, ordered=ordered)
            job.run(text)                                              # warm-up (sizes the slots in counted mode)
            torch.cuda.synchronize(); dist.barrier(); torch.cuda.synchronize()
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record(stream)
            shuf_ms = []
            for _ in range(steps):
                job.run(text)
                shuf_ms.append(wc.shuffle_ms())
            e1.record(stream)
            torch.cuda.synchronize(); dist.barrier()
            t = torch.tensor([e0.elapsed_time(e1) / steps, sum(shuf_ms) / len(shuf_ms), float(job.last_shuffle_bytes)], device=dev)
            tmax = t.clone(); dist.all_reduce(tmax, op=dist.ReduceOp.MAX)
            res = wc.result_arrays()
            stats = wc.stats()
            sent = float(tmax[2])
            if peers is not None and not peers.counted:
                # slots sized by the table: nothing counts the records on the timed path, so count them once here
                b = job.b
                b.reset(); b.map(text)
                if ordered:
                    hist = b.histogram(); dist.all_reduce(hist); b.partition_by_range(hist, world)
                sizes = b.shuffle_sizes(world)
                ts = torch.tensor([float(sum(x for r, x in enumerate(sizes) if r != rank))], device=dev)
                dist.all_reduce(ts, op=dist.ReduceOp.MAX)
                sent = float(ts.item())
        out = {"config": name, "n_gpus": world, "bytes": per * world, "ms_per_job": float(tmax[0]), "gbs": per * world / float(tmax[0]) / 1e6,
               "unique_keys": int(len(res[2])) if rank == 0 else None, "exchange": "peers" if peers else "nccl",
               "slots": ("counted" if peers and peers.counted else "by table size") if peers else None, "final_stage": final,
               "shuffle_ms": float(tmax[1]), "bytes_sent_max_rank": sent,
               "shuffle_gbs_per_gpu": sent / float(tmax[1]) / 1e6 if float(tmax[1]) > 0 else None,
               "shuffle_frac_of_770_gbs": sent / float(tmax[1]) / 1e6 / 770.0 if float(tmax[1]) > 0 else None}
        if check:
            t0 = time.perf_counter()
            host = text.cpu().numpy()
            ref = parity.oracle_arrays(host)
            del host
            with torch.cuda.stream(stream):
                wc.run(text)
                single = wc.result_arrays()
                why = parity.same_result(single, ref)
                if why is None and wc.stats()["tokens"] != ref[3]:
                    why = "tokens differ"
                del single
                torch.cuda.synchronize()
                dist.barrier()                       # the oracle took a different time on every rank; the flag barrier of the
                                                     # exchange gives up after five seconds, so the ranks enter the job together
                job.run(text)
                res = wc.result_arrays()
            okt = torch.tensor([1 if why is None else 0], dtype=torch.int64, device=dev)
            dist.all_reduce(okt, op=dist.ReduceOp.MIN)
            tok = torch.tensor([ref[3]], dtype=torch.int64, device=dev)
            dist.all_reduce(tok)
            if why:
                sys.stderr.write("parity (rank %d alone on its shard): %s\n" % (rank, why))
            chk = parity.check_distributed(dist, dev, rank, world, ref[:3], res[:3], res if rank == 0 else None)
            if rank == 0:
                out["parity"] = dict(chk, every_rank_alone_on_its_shard_equals_oracle=bool(okt.item()), bytes_checked=per * world,
                                     oracle_tokens=int(tok.item()), ok=bool(chk["ok"] and okt.item() and chk.get("tokens") == int(tok.item())),
                                     seconds=time.perf_counter() - t0)
            del ref, res
    finally:
        if peers:
            peers.close()
        wc.close()
        del text
        torch.cuda.empty_cache()
    return out if rank == 0 else None


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--configs", default="C3,C5,C4")
    ap.add_argument("--gib", type=float, default=0.0, help="override the total size (GiB)")
    ap.add_argument("--final", default="ordered", choices=["ordered", "gather"])
    ap.add_argument("--shuffle", default="peers", choices=["peers", "nccl"])
    ap.add_argument("--no-check", action="store_true")
    a = ap.parse_args()
    rank, world, lr = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"]), int(os.environ["LOCAL_RANK"])
    torch.cuda.set_device(lr)
    dev = torch.device("cuda", lr)
    dist.init_process_group("nccl", device_id=dev)
    st = torch.cuda.Stream(device=dev)
    bad = False
    for name in a.configs.split(","):
        total = int(a.gib * (1 << 30)) if a.gib else CONFIGS[name][5]
        out = run_config(name, total, world, rank, dev, st, a.final, a.shuffle, check=not a.no_check)
        if rank == 0:
            print(json.dumps(out), flush=True)
            bad = bad or (out.get("parity") is not None and not out["parity"]["ok"])
    dist.destroy_process_group()
    if bad:
        sys.exit(1)


if __name__ == "__main__":
    main()
