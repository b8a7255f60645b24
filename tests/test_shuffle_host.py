"""Host side of the multi-GPU path on CPU: shard planning and the exchange logic of
locust_b200.shuffle.DistributedWordCount, run with world_size 2 over gloo against a
pure-Python backend that speaks the same wire format as liblocust's shuffle kernels
(32-byte header {n_short, long_bytes}, 32-byte short records {k0, k1, count, 0}, long
records {len, count, bytes padded to 16}).  The oracle is the checker."""
import os
import re
import struct
import zlib

import pytest

from conftest import golden_input
from locust_b200.shuffle import DELIMS, DistributedWordCount, plan_shards

torch = pytest.importorskip("torch")


def test_plan_shards_cuts_on_delimiters():
    data = golden_input("hamlet")
    for p in (1, 2, 3, 4, 8):
        sh = plan_shards(data, p)
        assert sh[0][0] == 0 and sh[-1][1] == len(data)
        for (a, b), (c, d) in zip(sh, sh[1:]):
            assert b == c and a <= b
            assert b == len(data) or data[b - 1] in DELIMS        # no token straddles a cut
    # a text with one newline only (config 5 style) still cuts on spaces
    text = b"aaa bbb " * 1000
    sh = plan_shards(text, 8)
    assert all(text[b - 1] in DELIMS for _, b in sh[:-1]) and len({a for a, _ in sh}) == 8
    # one giant token: every cut slides to the end
    assert plan_shards(b"x" * 100, 4) == [(0, 100), (100, 100), (100, 100), (100, 100)]
    assert plan_shards(b"", 3) == [(0, 0)] * 3


class PyBackend:
    """Same interface as locust_b200.shuffle.GpuBackend, pure Python, CPU tensors."""

    def __init__(self):
        self.counts = {}
        self.result = None
        self.partmap = None

    def alloc(self, nbytes):
        return torch.zeros(max(int(nbytes), 32), dtype=torch.uint8)

    def sizes_tensor(self, values):
        return torch.tensor(values, dtype=torch.int64)

    def reset(self):
        self.counts, self.result, self.partmap = {}, None, None

    def map(self, buf):
        for t in re.split(b"[" + re.escape(DELIMS) + b"]+", bytes(buf)):
            if t:
                self.counts[t] = self.counts.get(t, 0) + 1

    def _parts(self, nparts):
        parts = [([], []) for _ in range(nparts)]
        for k, c in self.counts.items():
            dest = self.partmap[self._bin(k)] if self.partmap is not None else zlib.crc32(k) % nparts
            parts[dest][0 if len(k) <= 16 else 1].append((k, c))
        return parts

    # ordered shuffle: the same arithmetic as partition_hist_kernel / partition_map_kernel
    @staticmethod
    def _bin(k):
        return (k[0] << 8) | (k[1] if len(k) > 1 else 0)

    def histogram(self):
        h = torch.zeros(65536, dtype=torch.int64)
        for k in self.counts:
            h[self._bin(k)] += 1
        return h

    def partition_by_range(self, hist, nparts):
        hist = hist.tolist()
        total, before, self.partmap = sum(hist), 0, []
        for v in hist:
            self.partmap.append(min(before * nparts // total, nparts - 1) if total else 0)
            before += v

    def result_part(self):
        keys = [k for k, _ in self.result]
        counts = [c for _, c in self.result]
        n, blob = len(keys), b"".join(keys)
        offs = [sum(len(k) for k in keys[:i]) for i in range(n + 1)]
        vals = [sum(counts[:i]) for i in range(n)]
        nbytes = (32 + 8 * (n + 1) + 16 * n + len(blob) + 31) // 32 * 32
        raw = struct.pack("<4Q", n, len(blob), sum(counts), nbytes) + struct.pack("<%dQ" % (3 * n + 1), *offs, *counts, *vals) + blob
        t = torch.tensor(list(raw.ljust(nbytes, b"\0")), dtype=torch.uint8)
        return t, t[:32].view(torch.int64)

    def result_concat(self, buf, offsets, headers):
        raw = bytes(buf.numpy())
        self.result, total_before = [], 0
        for off, (n, blob_bytes, total, nbytes) in zip(offsets, headers):
            assert struct.unpack_from("<4Q", raw, off) == (n, blob_bytes, total, nbytes)
            words = struct.unpack_from("<%dQ" % (3 * n + 1), raw, off + 32)
            offs, counts, vals = words[:n + 1], words[n + 1:2 * n + 1], words[2 * n + 1:]
            blob = raw[off + 32 + 8 * (3 * n + 1):][:blob_bytes]
            assert list(vals) == [sum(counts[:i]) for i in range(n)] and sum(counts) == total
            self.result += [(blob[offs[i]:offs[i + 1]], counts[i]) for i in range(n)]
            total_before += total

    @staticmethod
    def _long_bytes(longs):
        return sum(16 + (len(k) + 15) // 16 * 16 for k, _ in longs)

    def shuffle_sizes(self, nparts):
        self._packed = self._parts(nparts)
        return [(32 + 32 * len(s) + self._long_bytes(l) + 31) // 32 * 32 for s, l in self._packed]

    def shuffle_pack(self, nparts, out, offsets):
        view = out.numpy()
        for (shorts, longs), off in zip(self._packed, offsets):
            blob = struct.pack("<4Q", len(shorts), self._long_bytes(longs), 32 + 32 * len(shorts), 0)
            for k, c in shorts:
                blob += k.ljust(16, b"\0") + struct.pack("<2Q", c, 0)
            for k, c in longs:
                blob += struct.pack("<2Q", len(k), c) + k.ljust((len(k) + 15) // 16 * 16, b"\0")
            view[off:off + len(blob)] = list(blob)

    def begin_merge(self):
        self.counts = {}

    def merge(self, buf, offset, nbytes):
        raw = bytes(buf.numpy()[offset:offset + nbytes])
        n_short, long_bytes, long_off, _ = struct.unpack_from("<4Q", raw, 0)
        assert 32 + 32 * n_short <= long_off and long_off + long_bytes <= nbytes
        p = 32
        for _ in range(n_short):
            k = raw[p:p + 16].rstrip(b"\0")
            c = struct.unpack_from("<Q", raw, p + 16)[0]
            self.counts[k] = self.counts.get(k, 0) + c
            p += 32
        p = long_off
        end = p + long_bytes
        while p < end:
            ln, c = struct.unpack_from("<2Q", raw, p)
            k = raw[p + 16:p + 16 + ln]
            self.counts[k] = self.counts.get(k, 0) + c
            p += 16 + (ln + 15) // 16 * 16

    def process(self):
        pass

    def reduce(self):
        self.result = sorted(self.counts.items())


def _worker(rank, world, port, data, out_path, ordered=False):
    import torch.distributed as dist
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        lo, hi = plan_shards(data, world)[rank]
        job = DistributedWordCount(PyBackend(), dist, ordered=ordered)
        for _ in range(2):                       # twice: the job object is reusable
            job.run(data[lo:hi])
        if rank == 0:
            with open(out_path, "wb") as f:
                for k, c in job.b.result:
                    f.write(k + b"\t%d\n" % c)
        elif ordered:
            # this rank kept the slice of its key range: sorted, and a proper part of the whole
            keys = [k for k, _ in job.b.result]
            assert keys == sorted(keys) and len(set(keys)) == len(keys)
            with open(out_path + ".%d" % rank, "wb") as f:
                f.write(b"\n".join(keys))
        else:
            assert job.b.result == []            # everything was gathered on rank 0
        assert job.last_shuffle_bytes % 32 == 0
    finally:
        dist.destroy_process_group()


@pytest.mark.parametrize("world", [2, 3])
def test_two_rank_exchange_matches_the_oracle(tmp_path, world):
    import torch.multiprocessing as mp
    from oracle import pyoracle
    data = golden_input("hamlet") + golden_input("edge_long_tokens") + b"\n" + golden_input("rand_longtok_s5")
    out = str(tmp_path / "result.tsv")
    port = 29500 + (os.getpid() % 2000) + world
    mp.spawn(_worker, args=(world, port, data, out), nprocs=world, join=True)
    with open(out, "rb") as f:
        got = f.read()
    assert got == pyoracle.wordcount(data).tsv


@pytest.mark.parametrize("world", [2, 3])
def test_ordered_exchange_matches_the_oracle(tmp_path, world):
    """ordered=True: parts are key ranges, every rank sorts its own, rank 0 concatenates."""
    import torch.multiprocessing as mp
    from oracle import pyoracle
    data = golden_input("hamlet") + golden_input("edge_long_tokens") + b"\n" + golden_input("rand_longtok_s5")
    out = str(tmp_path / "result.tsv")
    port = 29500 + (os.getpid() % 2000) + 10 + world
    mp.spawn(_worker, args=(world, port, data, out, True), nprocs=world, join=True)
    ref = pyoracle.wordcount(data)
    with open(out, "rb") as f:
        assert f.read() == ref.tsv
    # the slices the other ranks kept are consecutive tails of the sorted key list
    tail = []
    for r in range(1, world):
        with open(out + ".%d" % r, "rb") as f:
            got = f.read()
        tail += got.split(b"\n") if got else []
    assert tail == ref.keys[len(ref.keys) - len(tail):] and 0 < len(tail) < len(ref.keys)
