#!/usr/bin/env python
"""bench.py -- word-count input throughput (BASELINE.json metric) on N B200s of one node.

  python bench.py --gpus 1 --steps 5 --warmup 3
  python -m torch.distributed.run --nnodes=1 --nproc-per-node N --master-addr 127.0.0.1 \
         --master-port P bench.py --gpus N --steps K --warmup W
  python bench.py --impl reference ...      (the reference's CPU algorithm on host cores)

A step is one complete word-count job (map -> [shuffle] -> process -> reduce) over the
rank's shard of synthetic text.  N=1 runs BASELINE.json configs[1] (1 GiB Zipf(1.1), 50 K
vocabulary); N>1 runs the configs[2] text (Zipf, 1 M vocabulary) at 2 GiB per rank, cut on
delimiters, so N=8 is the full 16 GiB configuration ("weak" scaling).  One JSON line on rank 0.
"""
import argparse
import json
import os
import subprocess
import sys
import threading
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)

METRIC = "word-count input GB/s (device-timed, max over ranks)"
GB = 1e9
FALLBACK_HBM_GBS = 6650.0        # /opt/skills/guides/B200_PROFILING.md fallback


def measured_peak():
    p = os.path.join(ROOT, "MEASURED_PEAKS.json")
    try:
        with open(p) as f:
            return float(json.load(f)["hbm_gbs"]), "measured (MEASURED_PEAKS.json hbm_gbs)"
    except Exception:
        return FALLBACK_HBM_GBS, "fallback (B200_PROFILING.md)"


class ClockSampler(threading.Thread):
    """nvidia-smi clocks / throttle reasons sampled while the timed region runs."""
    Q = ("index,clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.active,clocks_event_reasons.hw_slowdown,"
         "clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,"
         "clocks_event_reasons.sw_power_cap")

    def __init__(self, gpu_index):
        super().__init__(daemon=True)
        self.gpu, self.rows, self.proc = gpu_index, [], None

    def run(self):
        try:
            self.proc = subprocess.Popen(["nvidia-smi", "-i", str(self.gpu), "--query-gpu=" + self.Q,
                                          "--format=csv,noheader,nounits", "-lms", "20"],
                                         stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
            for line in self.proc.stdout:
                self.rows.append([x.strip() for x in line.split(",")])
        except Exception:
            pass

    def wait_first_sample(self, timeout=3.0):
        t0 = time.time()
        while not self.rows and time.time() - t0 < timeout:
            time.sleep(0.01)

    def mark(self):
        return len(self.rows)

    def stop(self, lo=0, hi=None):
        if self.proc:
            self.proc.terminate()
        self.join(timeout=2)
        window = self.rows[lo:hi] if len(self.rows[lo:hi]) >= 1 else self.rows
        scope = "timed region" if window is not self.rows else "whole run (timed region shorter than one sample)"
        sm = sorted(int(r[1]) for r in window if len(r) > 8 and r[1].isdigit())
        mx = [int(r[2]) for r in window if len(r) > 8 and r[2].isdigit()]
        reasons = set()
        for r in window:
            if len(r) > 8:
                for name, v in zip(("hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"), r[5:9]):
                    if v.lower().startswith("active"):
                        reasons.add(name)
        return {"sm_mhz": sm[len(sm) // 2] if sm else None, "sm_max_mhz": max(mx) if mx else None,
                "reasons": sorted(reasons), "samples": len(window), "scope": scope}


def workload(n_gpus):
    if n_gpus == 1:
        return "C2", 1 << 30, "C2: 1 GiB synthetic Zipf(s=1.1) English-like text, 50K vocab, word count on 1xB200"
    return "C3", (16 << 30) // n_gpus, ("C3: 16 GiB synthetic Zipf(s=1.1) text, 1M-word vocab, sharded by byte range cut on "
                                        "delimiters across %d B200 (%d GiB per GPU), shuffle over NVLink" % (n_gpus, 16 // n_gpus))


def bind_to_gpu_cpus(index):
    """Runs this rank on the CPUs next to its GPU (NVML's affinity mask), so that the pinned staging
    buffers of the end-to-end leg are allocated in that socket's memory and eight ranks do not pull
    their input through one socket.  Returns the number of CPUs bound to, or None if it could not."""
    try:
        import pynvml
        pynvml.nvmlInit()
        try:                                            # CUDA_VISIBLE_DEVICES may renumber: go by UUID
            import torch
            h = pynvml.nvmlDeviceGetHandleByUUID(("GPU-" + str(torch.cuda.get_device_properties(index).uuid)).encode())
        except Exception:                               # noqa: BLE001
            h = pynvml.nvmlDeviceGetHandleByIndex(index)
        words = (os.cpu_count() + 63) // 64
        mask = pynvml.nvmlDeviceGetCpuAffinity(h, words)
        cpus = {64 * w + b for w, m in enumerate(mask) for b in range(64) if (m >> b) & 1}
        cpus &= os.sched_getaffinity(0)
        if cpus:
            os.sched_setaffinity(0, cpus)
            return len(cpus)
    except Exception:                                   # noqa: BLE001 - an optimisation only
        pass
    return None


def cpu_port_rate(sample, threads=1):
    """GB/s of the oracle's map -> sort -> run-length reduce (the reference's CPU stages) on `sample`."""
    from oracle import pyoracle
    t0 = time.perf_counter()
    r = pyoracle.wordcount(sample, mode="sort", formats=False)
    dt = time.perf_counter() - t0
    return len(sample) / dt / GB, dt, r


def c1_reference(wc):
    """BASELINE config 1: the reference's own CPU branch (oracle/_ref/MapReduce_cpu, verbatim source with
    GPU_IMPLEMENTATION 0) on hamlet.txt with its own three timers, next to this library on the same file."""
    ref_bin = os.path.join(ROOT, "oracle", "_ref", "MapReduce_cpu")
    hamlet = os.path.join(ROOT, "tests", "golden", "hamlet.txt")
    if not os.path.exists(ref_bin):
        return {"unavailable": "oracle/_ref/MapReduce_cpu was not built (needs /root/reference at build time)"}
    best = None
    for _ in range(3):
        t0 = time.perf_counter()
        out = subprocess.run([ref_bin, hamlet], capture_output=True, check=True).stdout
        wall = time.perf_counter() - t0
        ns = {}
        for line in out.split(b"\n")[:8]:
            for key in (b"CPU mapping", b"CPU sorting", b"CPU reducing"):
                if line.startswith(key):
                    ns[key.decode().split()[1]] = int(line.split()[2])
        tot = sum(ns.values())
        if best is None or tot < best[0]:
            best = (tot, ns, wall)
    with open(hamlet, "rb") as f:
        data = f.read()
    wc.run(data)
    wc.run(data)
    gpu_ms = sum(wc.stage_ms())
    return {"input": "hamlet.txt (191734 B)", "reference_cpu_stage_ns": best[1], "reference_cpu_stages_ms": best[0] / 1e6,
            "reference_cpu_process_wall_s": best[2], "cores": 1, "this_library_stages_ms": gpu_ms,
            "note": "0.19 MB is a parity case, not a throughput result"}


def run_reference(args, rank, world):
    """The reference's CPU implementation of the path (oracle port: oracle/_ref cannot take more than
    5 800 lines), one host thread like the reference, on a bounded sample of the same workload."""
    if rank != 0:
        return
    from locust_b200.textgen import TextGen
    name, nbytes, desc = workload(args.gpus)
    sample_bytes = args.cpu_sample_mib << 20
    g = TextGen.config(name, total_bytes=nbytes * (args.gpus if args.gpus > 1 else 1))
    sample = g.host(sample_bytes)
    times = []
    for i in range(args.warmup + args.steps):
        rate, dt, _ = cpu_port_rate(sample)
        if i >= args.warmup:
            times.append(dt)
    tot = sum(times)
    value = sample_bytes * len(times) / tot / GB
    line = {"impl": "reference", "metric": METRIC, "value": value, "unit": "GB/s", "n_gpus": args.gpus,
            "steps": args.steps, "warmup": args.warmup, "ms_per_step": 1e3 * tot / len(times),
            "higher_is_better": True, "scaling": "strong", "vs_baseline": None, "dtype": "u8", "data": "synthetic",
            "config": {"workload": desc + " [reference arm: timed on the first %d MiB of that text, one pass per step]"
                                   % args.cpu_sample_mib},
            "cpu_baseline": {"value": value, "unit": "GB/s", "cores": 1, "kind": "port",
                             "sample": "first %d MiB of the workload text; map -> comparison sort of every token "
                                       "-> run-length reduce (oracle/wordcount_oracle.c), single thread like the "
                                       "reference" % args.cpu_sample_mib},
            "e2e": {"value": value, "unit": "GB/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0}}
    print(json.dumps(line), flush=True)


def c3_on_one_gpu(wc, dev, stream, index, nbytes=2 << 30):
    """BASELINE config 3's text (1 M vocabulary) on ONE GPU, 2 GiB as at N = 8: the base the N >= 2 numbers compare
    with (N = 1 of the contract runs config 2, another text).  Timed like the headline, checked against the oracle."""
    import torch
    from locust_b200.textgen import TextGen
    from oracle import parity as par
    text = torch.empty(nbytes, dtype=torch.uint8, device=dev)
    TextGen.config("C3", total_bytes=16 << 30, device=index).fill_device(text)
    torch.cuda.synchronize()
    with torch.cuda.stream(stream):
        for _ in range(3):
            wc.run(text)
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record(stream)
        maps = []
        for _ in range(5):
            wc.run(text)
            maps.append(wc.stage_ms()[0])
        e1.record(stream)
    torch.cuda.synchronize()
    ms = e0.elapsed_time(e1) / 5
    got = wc.result_arrays()
    ref = par.oracle_arrays(text.cpu().numpy())
    why = par.same_result(got, ref)
    return {"workload": "first 2 GiB of the C3 text (1M vocab), one GPU, table 2^22 slots as at N > 1", "ms_per_step": ms,
            "value": nbytes / ms / 1e6, "unit": "GB/s", "map_ms": sum(maps) / len(maps), "unique_keys": int(len(got[2])),
            "parity": {"oracle": "wc_oracle_fast", "bytes_checked": nbytes, "ok": why is None, "why": why}}


def run_driver_c(args):
    """N GPUs driven from this one process through locust_multi_wordcount_device: the same workload and timing
    rules as the torchrun arm (device-resident shards, events on rank 0's stream), parity against the oracle."""
    import threading as th
    import torch
    import locust_b200
    from locust_b200.textgen import TextGen
    from oracle import parity as par
    n = args.gpus
    name, per_gpu, desc = workload(n)
    if args.mib:
        per_gpu = args.mib << 20
        desc += " [size overridden to %d MiB per GPU: NOT the benchmark configuration]" % args.mib
    total = per_gpu * n
    ndev = torch.cuda.device_count()
    devices = list(range(n)) if ndev >= n else [i % ndev for i in range(n)]
    shards = []
    for r, d in enumerate(devices):
        t = torch.empty(per_gpu, dtype=torch.uint8, device="cuda:%d" % d)
        TextGen.config(name, total_bytes=total, device=d).fill_device(t, byte_offset=r * per_gpu)
        shards.append(t)
    for d in set(devices):
        torch.cuda.synchronize(d)
    table_log2 = args.table_log2 or (20 if name == "C2" else 22)
    m = locust_b200.MultiWordCount(devices, table_log2=table_log2)
    for _ in range(args.warmup):
        m.run_shards(shards)
    ms, t0 = [], time.perf_counter()
    for _ in range(args.steps):
        wc0 = m.run_shards(shards)
        ms.append(m.stats()["job_ms"])
    wall = time.perf_counter() - t0
    res = wc0.result_arrays()
    refs = [None] * n

    def orc(r):
        refs[r] = par.oracle_arrays(shards[r].cpu().numpy())[:3]
    ts = [th.Thread(target=orc, args=(r,)) for r in range(n)]
    [t.start() for t in ts]
    [t.join() for t in ts]
    parity = dict(par.check_sum_of_shards(refs, res), bytes_checked=total)
    mean = sum(ms) / len(ms)
    line = {"metric": METRIC, "value": total / mean / 1e6, "unit": "GB/s", "n_gpus": n, "steps": args.steps, "warmup": args.warmup,
            "ms_per_step": mean, "ms_per_step_median": sorted(ms)[len(ms) // 2], "ms_per_step_wall": 1e3 * wall / args.steps,
            "higher_is_better": True, "scaling": "strong", "vs_baseline": None, "dtype": "u8", "data": "synthetic",
            "driver": "c (one process, locust_multi_wordcount_device; device time on rank 0's stream from its first kernel to "
                      "the result in place)",
            "config": {"workload": desc, "bytes_per_gpu": per_gpu, "devices": devices, "table_log2": table_log2,
                       "unique_keys": int(len(res[2]))},
            "shuffle": {"record_bytes_exchanged": m.stats()["record_bytes_exchanged"]}, "parity": parity}
    print(json.dumps(line), flush=True)
    m.close()
    if not parity["ok"]:
        sys.exit("bench.py --driver c: the result differs from the oracle: %s" % parity.get("why"))


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=10)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="locust_b200", choices=["locust_b200", "reference"])
    ap.add_argument("--mib", type=int, default=0, help="override per-GPU input size (MiB); not the benchmark config")
    ap.add_argument("--cpu-sample-mib", type=int, default=256,
                    help="size of the text sample the CPU port is timed on (about 10 s of one core per pass at 256)")
    ap.add_argument("--table-log2", type=int, default=0)
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--other-configs", default="auto", choices=["auto", "none", "all"],
                    help="after the headline measurement also run BASELINE configs 5 and 4 (8 GiB each, parity-checked) and "
                         "report them as other_configs; auto = at 8 GPUs")
    ap.add_argument("--driver", default="torchrun", choices=["torchrun", "c"],
                    help="N>1: 'torchrun' = one process per GPU (the contract's launch); 'c' = this one process drives all "
                         "GPUs through locust_multi_* (the C entry point a host program would call)")
    ap.add_argument("--time-budget", type=float, default=600.0, help="seconds after which optional extras are skipped")
    ap.add_argument("--other-gib", type=float, default=8.0, help="total size of each of the other_configs (8 = as BASELINE.json states)")
    ap.add_argument("--final", default="ordered", choices=["ordered", "gather"],
                    help="N>1: 'ordered' = parts are key ranges, every rank sorts its range, rank 0 concatenates the "
                         "slices; 'gather' = hash parts, then everything is sent to rank 0 and sorted there")
    ap.add_argument("--shuffle", default="peers", choices=["peers", "nccl"],
                    help="N>1: 'peers' = packing kernel writes into the destination GPU's memory over NVLink "
                         "(fused pack+send), 'nccl' = pack locally, all_to_all_single")
    args = ap.parse_args()
    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local_rank = int(os.environ.get("LOCAL_RANK", "0"))
    if args.impl == "reference":
        run_reference(args, rank, world)
        return
    if args.driver == "c":
        run_driver_c(args)
        return
    t_start = time.time()
    if world != args.gpus:
        if world == 1 and args.gpus > 1:
            sys.exit("bench.py --gpus %d must be launched with torch.distributed.run --nproc-per-node %d"
                     % (args.gpus, args.gpus))
        sys.exit("WORLD_SIZE=%d does not match --gpus %d" % (world, args.gpus))

    import torch
    import torch.distributed as dist
    import locust_b200
    from locust_b200.textgen import TextGen
    if not torch.cuda.is_available():
        sys.exit("bench.py needs a CUDA device: the word-count path has no CPU implementation")
    locust_b200.load()
    torch.cuda.set_device(local_rank)
    dev = torch.device("cuda", local_rank)
    numa = bind_to_gpu_cpus(local_rank) if world > 1 else None
    if world > 1:
        os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
        dist.init_process_group("nccl", device_id=dev)

    name, per_gpu, desc = workload(world)
    if args.mib:
        per_gpu = args.mib << 20
        desc += " [size overridden to %d MiB per GPU: NOT the benchmark configuration]" % args.mib
    total = per_gpu * world
    gen = TextGen.config(name, total_bytes=total, device=local_rank)
    text = torch.empty(per_gpu, dtype=torch.uint8, device=dev)
    gen.fill_device(text, byte_offset=rank * per_gpu)      # blocks end in a delimiter: shard cuts are token boundaries
    torch.cuda.synchronize()

    stream = torch.cuda.Stream(device=dev)
    table_log2 = args.table_log2 or (20 if name == "C2" else 22)      # 20 is the library default
    wc = locust_b200.WordCount(device=local_rank, table_log2=table_log2, stream=stream.cuda_stream)
    if world > 1:
        from locust_b200.shuffle import DistributedWordCount, GpuBackend, PeerExchange
        peers = None
        if args.shuffle == "peers":
            # every rank must end up on the same path: if mapping the peers' buffers fails anywhere
            # (IPC not permitted in some container set-ups), all ranks fall back to the NCCL exchange
            ok = torch.ones(1, dtype=torch.int32, device=dev)
            with torch.cuda.stream(stream):
                try:
                    peers = PeerExchange(wc, dist, dev, 1 << table_log2, ordered=args.final == "ordered")
                except Exception as e:                       # noqa: BLE001 - reported, then agreed on below
                    sys.stderr.write("rank %d: peer exchange unavailable (%s)\n" % (rank, e))
                    ok.zero_()
                dist.all_reduce(ok, op=dist.ReduceOp.MIN)
            torch.cuda.synchronize()
            if int(ok.item()) == 0:
                peers, args.shuffle = None, "nccl"
        job = DistributedWordCount(GpuBackend(wc, dev), dist, peers=peers, ordered=args.final == "ordered")
        step = lambda buf: job.run(buf)
    else:
        step = lambda buf: wc.run(buf)

    def barrier():
        torch.cuda.synchronize()
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    def timed(buf, k):
        """device time of k steps (ms, max over ranks), the per-step map-kernel times, and the
        per-step device times (events between steps; max over ranks per step)"""
        map_ms, shuf = [], []
        barrier()
        ev = [torch.cuda.Event(enable_timing=True) for _ in range(k + 1)]
        with torch.cuda.stream(stream):
            ev[0].record(stream)
            for i in range(k):
                step(buf)
                ev[i + 1].record(stream)
                map_ms.append(wc.stage_ms()[0])
                if world > 1:
                    shuf.append((wc.shuffle_ms(), job.last_shuffle_bytes))
        barrier()
        ms = ev[0].elapsed_time(ev[k])
        per = [ev[i].elapsed_time(ev[i + 1]) for i in range(k)]
        if world > 1:
            t = torch.tensor([ms] + per, device=dev)
            dist.all_reduce(t, op=dist.ReduceOp.MAX)
            ms, per = float(t[0].item()), [float(x) for x in t[1:].tolist()]
        return ms, map_ms, shuf, per

    # ---- device-resident input ("value") ----------------------------------------------
    sampler = ClockSampler(local_rank)
    sampler.start()
    sampler.wait_first_sample()
    with torch.cuda.stream(stream):
        for _ in range(args.warmup):
            step(text)
    m0 = sampler.mark()
    ms, map_ms, shuf, step_ms = timed(text, args.steps)
    m1 = sampler.mark()
    st = wc.stats()
    stage_last = dict(zip(("map", "process", "reduce"), wc.stage_ms()))
    launches_per_step = st["kernel_launches"]
    uniq, tokens = st["unique_keys"], st["tokens"]
    value = total * args.steps / (ms * 1e-3) / GB

    # roofline of the dominant kernel (tokenize_combine, the whole MAP stage of a step)
    blob, offs, counts, vals = wc.result_arrays()
    alg_bytes = per_gpu + int(len(blob)) + 8 * int(len(counts))     # N_in + sum(len_k + 8), SURVEY.md 8(d)
    avg_map_ms = sum(map_ms) / len(map_ms)
    peak, peak_src = measured_peak()
    achieved = alg_bytes / (avg_map_ms * 1e-3) / GB
    # DRAM bytes per launch from the committed `ncu --set full` capture of this kernel on this
    # workload (profiles/rNN_map_kernel_traffic.json, written by tools/summarize_profiles.py)
    traffic, traffic_src = None, None
    import glob
    for tp in sorted(glob.glob(os.path.join(ROOT, "profiles", "r*_map_kernel_traffic.json")), reverse=True):
        try:
            with open(tp) as f:
                t = json.load(f)
            if world == 1 and t.get("input_bytes_of_capture") == per_gpu:
                traffic, traffic_src = t.get("dram_bytes_per_launch"), os.path.relpath(tp, ROOT)
                break
        except Exception:
            pass
    roofline = {"bound": "hbm", "kernel": "tokenize_combine_kernel", "achieved": achieved, "peak": peak, "unit": "GB/s",
                "frac": achieved / peak, "frac_of_nominal_8000_gbs": achieved / 8000.0, "traffic": traffic, "traffic_source": traffic_src, "peak_source": peak_src,
                "algorithmic_bytes_per_launch": alg_bytes, "avg_kernel_ms": avg_map_ms,
                "map_share_of_step": avg_map_ms * args.steps / ms}

    shuffle = None
    if world > 1 and shuf:
        # both exchanges (hash shuffle + gather) with their host-side size round trips; bytes = what this
        # rank sent to other ranks in the hash shuffle.  770 GB/s = measured peer copy per direction
        # (B200_PROFILING.md); the shuffle here is latency-, not bandwidth-bound.
        sms_, sb_ = sum(x for x, _ in shuf) / len(shuf), sum(b for _, b in shuf) / len(shuf)
        if args.shuffle == "peers":
            # fused pack+send: the packing kernel stores every record in the destination GPU's memory and no
            # sizes are exchanged, so the bytes that leave this rank are counted once here, outside the timing
            with torch.cuda.stream(stream):
                b = job.b
                b.reset()
                b.map(text)
                if args.final == "ordered":
                    hist = b.histogram()
                    dist.all_reduce(hist)
                    b.partition_by_range(hist, world)
                sizes = b.shuffle_sizes(world)
            sb_ = float(sum(x for r, x in enumerate(sizes) if r != rank))
            shuffle = {"path": "peers (pack kernel writes over NVLink peer memory; the ranks' streams meet at flag "
                               "barriers in peer memory)",
                       "ms": sms_, "interval": "key histogram + its exchange + range map + pack/send + barrier + merge"
                       if args.final == "ordered" else "both exchanges (hash shuffle, gather) with their barriers and merges",
                       "bytes_sent_rank0": sb_, "achieved_gbs": sb_ / (sms_ * 1e-3) / GB if sms_ > 0 else None,
                       "peak_gbs": 770.0, "frac": (sb_ / (sms_ * 1e-3) / GB / 770.0) if sms_ > 0 else None,
                       "note": "latency bound: tens of MB per rank against a 770 GB/s link"}
        else:
            shuffle = {"path": "nccl (local pack, all_to_all_single of sizes and records)", "ms": sms_,
                       "bytes_sent_rank0": sb_, "achieved_gbs": sb_ / (sms_ * 1e-3) / GB if sms_ > 0 else None,
                       "peak_gbs": 770.0, "frac": (sb_ / (sms_ * 1e-3) / GB / 770.0) if sms_ > 0 else None}

    # ---- end to end: pinned host input in, host result out, every step ---------------------
    host = torch.empty(per_gpu, dtype=torch.uint8, pin_memory=True)
    host.copy_(text)
    torch.cuda.synchronize()
    d2h = int(len(blob)) + 8 * (int(len(offs)) + 2 * int(len(counts)))

    def e2e_step(buf):
        step(buf)
        wc.result_arrays()

    e2e_steps = max(2, min(args.steps, 5))
    with torch.cuda.stream(stream):
        e2e_step(host)
    barrier()
    t0 = time.perf_counter()
    with torch.cuda.stream(stream):
        for _ in range(e2e_steps):
            e2e_step(host)
    barrier()
    e2e_s = time.perf_counter() - t0
    if world > 1:
        t = torch.tensor([e2e_s], device=dev)
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        e2e_s = float(t.item())
    e2e = {"value": total * e2e_steps / e2e_s / GB, "unit": "GB/s", "h2d_bytes_per_step": per_gpu,
           "d2h_bytes_per_step": d2h, "steps": e2e_steps,
           "note": "locust_map(LOCUST_HOST) from pinned memory, chunked copy overlapped with map; result copied to host"}

    # what the host links deliver for the same bytes: a bare pinned host -> device copy on every rank at
    # the same time, no word count (whole-job GB/s, slowest rank)
    scratch = torch.empty(per_gpu, dtype=torch.uint8, device=dev)
    with torch.cuda.stream(stream):
        scratch.copy_(host, non_blocking=True)
    barrier()
    t0 = time.perf_counter()
    with torch.cuda.stream(stream):
        for _ in range(3):
            scratch.copy_(host, non_blocking=True)
    barrier()
    bare_s = time.perf_counter() - t0
    if world > 1:
        t = torch.tensor([bare_s], device=dev)
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        bare_s = float(t.item())
    e2e["bare_h2d_copy_gbs"] = 3 * total / bare_s / GB
    e2e["frac_of_bare_copy"] = e2e["value"] / e2e["bare_h2d_copy_gbs"]
    del scratch

    clocks = sampler.stop(m0, m1)

    # ---- parity of what was just timed, against the CPU oracle (outside every timed region) ----
    import hashlib
    import numpy as np
    from oracle import parity as par
    host_np = host.numpy()
    chunk_sha = [hashlib.sha256(host_np[o:o + (1 << 30)]).hexdigest() for o in range(0, per_gpu, 1 << 30)]
    t_par = time.perf_counter()
    ref = par.oracle_arrays(host_np)                                  # this rank's shard, wc_oracle_fast
    with torch.cuda.stream(stream):
        wc.run(text)
    single = wc.result_arrays()
    my_tokens = wc.stats()["tokens"]
    why = par.same_result(single, ref)
    if why is None and my_tokens != ref[3]:
        why = "tokens: got %d, oracle %d" % (my_tokens, ref[3])
    if why:
        sys.stderr.write("parity (rank %d, single GPU on its shard): %s\n" % (rank, why))
    parity = {"oracle": "wc_oracle_fast (oracle/wordcount_oracle.c) on the full input of the timed steps",
              "bytes_checked": total, "keys": int(len(ref[2])), "ok": why is None}
    tokens_per_gpu, sha_all = [my_tokens], [chunk_sha]
    if world > 1:
        okt = torch.tensor([1 if why is None else 0], dtype=torch.int64, device=dev)
        dist.all_reduce(okt, op=dist.ReduceOp.MIN)
        tk = torch.zeros(world, dtype=torch.int64, device=dev)
        dist.all_gather_into_tensor(tk, torch.tensor([my_tokens], dtype=torch.int64, device=dev))
        tokens_per_gpu = [int(x) for x in tk.tolist()]
        sha_all = [None] * world
        dist.all_gather_object(sha_all, chunk_sha)
        torch.cuda.synchronize()
        dist.barrier()                    # the oracle took a different time on every rank: enter the job together
        with torch.cuda.stream(stream):
            job.run(text)
            res = wc.result_arrays()
        chk = par.check_distributed(dist, dev, rank, world, ref[:3], res[:3], res if rank == 0 else None)
        if rank == 0:
            parity = dict(chk, bytes_checked=total, every_rank_alone_on_its_shard_equals_oracle=bool(okt.item()),
                          ok=bool(chk["ok"] and okt.item()))
    parity["seconds"] = time.perf_counter() - t_par
    del ref, single

    # ---- the reference's CPU algorithm on this box's host cores (rank 0, N=1) -----------------
    cpu = None
    input_id = None
    if rank == 0 and world == 1 and not args.no_cpu_baseline:
        sample = host[: args.cpu_sample_mib << 20].numpy()
        rate, dt, r = cpu_port_rate(sample)
        cpu = {"value": rate, "unit": "GB/s", "cores": 1, "kind": "port",
               "sample": "first %d MiB of the same text (%.1f s): map -> comparison sort of every token -> run-length "
                         "reduce, oracle/wordcount_oracle.c, single thread like the reference"
                         % (args.cpu_sample_mib, dt)}
        cpu["c1_reference"] = c1_reference(wc)
        input_id = hashlib.sha256(host_np).hexdigest()                 # identity of the generated input

    # ---- extras: the same-text base of the N > 1 series, and the configs the headline does not cover ----
    c3_single, other = None, None
    del host, host_np
    if world == 1 and not args.mib and args.other_configs != "none" and time.time() - t_start < args.time_budget:
        wc3 = locust_b200.WordCount(device=local_rank, table_log2=22, stream=stream.cuda_stream)     # the table the N > 1 runs use
        c3_single = c3_on_one_gpu(wc3, dev, stream, local_rank)
        wc3.close()
    if world > 1 and (args.other_configs == "all" or (args.other_configs == "auto" and world == 8)):
        from tools.check_full_configs import run_config
        if rank == 0:
            sys.stderr.write("headline done: value %.1f GB/s, %.3f ms per step, parity ok=%s; running other_configs\n"
                             % (value, ms / args.steps, parity["ok"]))
        del text
        torch.cuda.empty_cache()
        other = []
        for cname in ("C5", "C4"):
            if time.time() - t_start > args.time_budget:
                if rank == 0:
                    other.append({"config": cname, "skipped": "time budget of %.0f s used up" % args.time_budget})
                continue
            o = run_config(cname, int(args.other_gib * (1 << 30)), world, rank, dev, stream, args.final, args.shuffle)
            if rank == 0:
                other.append(o)

    if rank == 0:
        line = {"metric": METRIC, "value": value, "unit": "GB/s", "n_gpus": world, "steps": args.steps,
                "warmup": args.warmup, "ms_per_step": ms / args.steps,
                "ms_per_step_median": sorted(step_ms)[len(step_ms) // 2], "ms_per_step_all": [round(x, 4) for x in step_ms],
                "higher_is_better": True, "scaling": "strong",
                "vs_baseline": None, "dtype": "u8", "data": "synthetic",
                "config": {"workload": desc, "bytes_per_gpu": per_gpu, "tokens_per_gpu": tokens_per_gpu,
                           "unique_keys": uniq, "table_log2": table_log2, "input_sha256": input_id,
                           "input_sha256_per_gib_chunk_per_rank": sha_all,
                           "final_stage": (args.final if world > 1 else None),
                           "cpus_bound_per_rank": numa,
                           "series": "N=1 is BASELINE config 2 (1 GiB, 50K vocab); N=2/4/8 are config 3 (16 GiB in total, "
                                     "1M vocab): total work is fixed from N=2 on",
                           "l2": "input per step (%d MiB per GPU) is larger than the 126 MB L2" % (per_gpu >> 20)},
                "roofline": roofline, "shuffle": shuffle, "cpu_baseline": cpu, "e2e": e2e, "parity": parity, "clocks": clocks,
                "gpu_launches": launches_per_step * args.steps, "gpu_launches_per_step": launches_per_step,
                "stage_ms_last_step": stage_last, "c3_single_gpu": c3_single, "other_configs": other}
        print(json.dumps(line), flush=True)
    wc.close()
    if world > 1:
        dist.destroy_process_group()
    if rank == 0 and not parity["ok"]:
        sys.exit("bench.py: the timed result differs from the oracle: %s" % parity.get("why", "see stderr"))
    if rank == 0 and other and any(o.get("parity") and not o["parity"]["ok"] for o in other):
        sys.exit("bench.py: a result of other_configs differs from the oracle")


if __name__ == "__main__":
    main()
