"""Multi-GPU word count: one process per GPU, shards by byte range, partial (word, count)
records exchanged by hash(word) % P.

Replaces the reference's multi mode -- `[line_start, line_end)` windows per node
(MapReduce/src/main.cu:367-374), sorted partials written to /tmp/out.txt (:428-430), a TCP
command channel (Distributor/slave.py) -- with an all-to-all over NVLink:

    map (each rank, its shard)            locust_map
    pack partials by destination          locust_shuffle_sizes / locust_shuffle_pack
    all-to-all of sizes, then of records  torch.distributed (NCCL; gloo in the CPU tests)
      -- or, on one box, both at once:    locust_shuffle_pack_peers writes every part straight
                                          into the destination GPU's memory over NVLink (PeerExchange)
    merge what arrived                    locust_shuffle_begin_merge / locust_shuffle_merge
    gather the partitions on rank 0       the same pack / exchange / merge with one destination
    process + reduce on rank 0            locust_process / locust_reduce  -> sorted result

  ordered=True replaces the last three lines: the parts of the exchange are key ranges
  (locust_partition_histogram, all-reduce, locust_partition_by_range), every rank runs process +
  reduce on its own range, and rank 0 receives the sorted slices and puts them one after the
  other (locust_result_part_pack / locust_result_concat) -- nobody sorts the whole result.

torch.distributed is plumbing only; every byte of arithmetic happens in liblocust.so.  The class
is written against a small backend interface so the exchange logic can be tested on CPU with
gloo (tests/test_shuffle_host.py) without a GPU.
"""

DELIMS = b" ,.-;:'()\"\t\n"


def plan_shards(data, nparts):
    """Byte ranges [(lo, hi)] of `data` (bytes-like) for nparts ranks: nominal r*N/P, moved
    forward to the first position whose preceding byte is a delimiter, so no token straddles
    two ranks.  Works on any delimiter, not only newlines."""
    n = len(data)
    cuts = [0]
    for r in range(1, nparts):
        c = max(cuts[-1], (r * n) // nparts)
        while 0 < c < n and data[c - 1] not in DELIMS:
            c += 1
        cuts.append(min(c, n))
    cuts.append(n)
    return [(cuts[i], cuts[i + 1]) for i in range(nparts)]


class GpuBackend:
    """liblocust.so + torch CUDA tensors.

    liblocust enqueues on the context's own stream (locust_stream); torch's collectives and its
    allocator order only against torch's *current* stream.  on_stream() makes the context's stream
    torch's current one -- DistributedWordCount.run does everything inside it -- so that a
    collective never reads a buffer the packing kernel is still writing."""

    def __init__(self, wc, device):
        import torch
        self.wc, self.device, self.torch = wc, device, torch
        self.stream = torch.cuda.ExternalStream(wc.stream_ptr(), device=device)

    def on_stream(self):
        return self.torch.cuda.stream(self.stream)

    def alloc(self, nbytes):
        return self.torch.empty(max(int(nbytes), 32), dtype=self.torch.uint8, device=self.device)

    def sizes_tensor(self, values):
        return self.torch.tensor(values, dtype=self.torch.int64, device=self.device)

    def reset(self):
        self.wc.reset()

    def map(self, buf):
        self.wc.map(buf)

    def shuffle_sizes(self, nparts):
        return self.wc.shuffle_sizes(nparts)

    def shuffle_pack(self, nparts, out, offsets):
        self.wc.shuffle_pack(nparts, out.data_ptr(), offsets)

    def begin_merge(self):
        self.wc.shuffle_begin_merge()

    def merge(self, buf, offset, nbytes):
        self.wc.shuffle_merge(buf.data_ptr() + offset, nbytes, keepalive=buf)

    def process(self):
        self.wc.process()

    def reduce(self):
        self.wc.reduce()

    # ordered shuffle
    def histogram(self):
        h = self.torch.zeros(65536, dtype=self.torch.int64, device=self.device)
        self.wc.partition_histogram(h.data_ptr())
        return h

    def partition_by_range(self, hist, nparts):
        self.wc.partition_by_range(hist.data_ptr(), nparts)

    def result_part(self):
        """(packed result of this rank as a uint8 tensor, its 4-word header as an int64 tensor)"""
        nbytes = self.wc.result_part_bytes()
        buf = self.alloc(nbytes)
        self.wc.result_part_pack(buf.data_ptr())
        return buf[:nbytes], buf[:32].view(self.torch.int64)

    def result_concat(self, buf, offsets, headers):
        self.wc.result_concat([buf.data_ptr() + o for o in offsets], [h[0] for h in headers], [h[1] for h in headers],
                              [h[2] for h in headers], keepalive=buf)


class PeerExchange:
    """Receive buffers for the fused pack + send path, mapped into every rank of one box.

    Rank r owns two buffers of `world` slots each (hash shuffle, gather); slot s of a buffer is
    where rank s writes.  The buffers are cudaMalloc'ed by liblocust, their IPC handles are
    all-gathered once, and every rank maps every other rank's buffers (NVLink peer access).  A
    slot holds `short_cap` short records plus `long_room` bytes of long records.  By default short_cap
    is the sender's table size, so short records can never overflow.  With a very large table (BASELINE
    config 4: 2^27 slots = 4 GB per slot) that is not affordable: `counted` mode starts with small slots,
    DistributedWordCount counts the records of every part before packing (locust_shuffle_sizes /
    locust_shuffle_counts, which also lets a full table grow and redo MAP) and the slots are re-made
    when the largest part of any rank does not fit (ensure_shuffle / ensure_result, collective)."""

    HIST_BYTES = 65536 * 8
    COUNTED_ABOVE = 256 << 20            # slot size from which the slots are sized by counting

    def __init__(self, wc, dist, device, table_slots, long_room=16 << 20, group=None, result_room=None, ordered=None,
                 flag_barrier=True, counted=None):
        import torch
        self.wc, self.dist, self.group, self.device = wc, dist, group, device
        self.rank, self.world = dist.get_rank(group), dist.get_world_size(group)
        self.short_cap = int(table_slots)
        self.counted = (32 + 32 * self.short_cap + int(long_room) > self.COUNTED_ABOVE) if counted is None else bool(counted)
        if self.counted:
            self.short_cap, long_room = min(self.short_cap, 1 << 18), min(int(long_room), 1 << 20)
        self.long_room = int(long_room)
        self._need_gather = None
        self.slot_bytes = 32 + 32 * self.short_cap + int(long_room)
        self.slot_bytes = (self.slot_bytes + 31) // 32 * 32
        # ordered jobs: rank 0 owns one slot per rank for that rank's sorted slice (offsets, counts, vals
        # and key bytes of as many keys as a table has slots, 24 key bytes each on average)
        self.result_slot = (int(result_room) if result_room else 64 + 48 * self.short_cap + 31) // 32 * 32
        # only rank 0 receives the gather and the result; everyone receives the shuffle.  A rank that
        # cannot allocate still takes part in the all-gather (with an empty handle) so nobody hangs.
        # fourth buffer: `world` barrier flags (256 bytes), then one histogram slot per rank
        nbuf = 4
        self.flag_barrier, self._epoch = bool(flag_barrier), 0
        self.mine, self.bufs, raw = [], [], bytes(64 * nbuf)
        try:
            # ordered=True / False: only the buffers that kind of job uses (None: both)
            need_gather = self.rank == 0 and ordered is not True
            need_result = self.rank == 0 and ordered is not False
            self._need_gather = need_gather
            self.mine = [wc.exchange_alloc(self.slot_bytes * self.world),
                         wc.exchange_alloc(self.slot_bytes * self.world if need_gather else 64),
                         wc.exchange_alloc(self.result_slot * self.world if need_result else 64),
                         wc.exchange_alloc(256 + self.HIST_BYTES * self.world)]
            raw = b"".join(wc.ipc_export(m) for m in self.mine)
        except Exception:                                        # noqa: BLE001 - raised below, on every rank
            pass
        handles = torch.tensor(list(raw), dtype=torch.uint8, device=device)
        allh = torch.empty(self.world * 64 * nbuf, dtype=torch.uint8, device=device)
        dist.all_gather_into_tensor(allh, handles, group=group)
        allh = bytes(allh.cpu().numpy())
        per = 64 * nbuf
        if any(allh[r * per:(r + 1) * per] == bytes(per) for r in range(self.world)):
            raise RuntimeError("a rank could not allocate or export its exchange buffers")
        for r in range(self.world):     # bufs[r] = (shuffle, gather, result, sync buffer) of rank r, as seen from here
            if r == self.rank:
                self.bufs.append(tuple(self.mine))
            else:
                h = allh[r * per:(r + 1) * per]
                self.bufs.append(tuple(wc.ipc_open(h[64 * i:64 * (i + 1)]) for i in range(nbuf)))
        self._token = torch.zeros(1, dtype=torch.int32, device=device)

    def _remake(self, index, nbytes):
        """collective: this rank's buffer `index` is replaced by one of nbytes and every rank maps the new ones"""
        import torch
        wc = self.wc
        torch.cuda.synchronize(self.device)
        self.dist.barrier(group=self.group)                       # nobody is still writing into the old buffers
        for r, mapped in enumerate(self.bufs):
            if r != self.rank:
                wc.ipc_close(mapped[index])
        wc.exchange_free(self.mine[index])
        raw = bytes(64)
        try:
            self.mine[index] = wc.exchange_alloc(nbytes)
            raw = wc.ipc_export(self.mine[index])
        except Exception:                                        # noqa: BLE001 - raised below, on every rank
            pass
        handles = torch.tensor(list(raw), dtype=torch.uint8, device=self.device)
        allh = torch.empty(self.world * 64, dtype=torch.uint8, device=self.device)
        self.dist.all_gather_into_tensor(allh, handles, group=self.group)
        allh = bytes(allh.cpu().numpy())
        if any(allh[r * 64:(r + 1) * 64] == bytes(64) for r in range(self.world)):
            raise RuntimeError("a rank could not allocate or export a larger exchange buffer")
        for r in range(self.world):
            b = list(self.bufs[r])
            b[index] = self.mine[index] if r == self.rank else wc.ipc_open(allh[r * 64:(r + 1) * 64])
            self.bufs[r] = tuple(b)

    def ensure_shuffle(self, short_need, long_need):
        """collective (same arguments on every rank): slots that hold short_need short records and
        long_need bytes of long records"""
        if short_need <= self.short_cap and long_need <= self.long_room:
            return
        self.short_cap = int(short_need) + int(short_need) // 4 + 1024
        self.long_room = (int(long_need) + int(long_need) // 4 + 4096 + 31) // 32 * 32
        self.slot_bytes = 32 + 32 * self.short_cap + self.long_room
        self._remake(0, self.slot_bytes * self.world)

    def ensure_result(self, slice_need):
        """collective: result slots on rank 0 that hold a sorted slice of slice_need bytes"""
        if slice_need <= self.result_slot:
            return
        self.result_slot = (int(slice_need) + int(slice_need) // 4 + 4096 + 31) // 32 * 32
        self._remake(2, self.result_slot * self.world if self.rank == 0 else 64)

    def barrier(self):
        """every rank's stream has got this far (its packing kernel has finished, say) before any rank's
        stream goes on: flags written into each other's memory by a one-block kernel, or a one-element
        all-reduce; either way on the current stream, no host synchronisation"""
        if self.flag_barrier:
            self._epoch += 1
            self.wc.peer_barrier([b[3] for b in self.bufs], self.rank, self.world, self._epoch)
        else:
            self.dist.all_reduce(self._token, group=self.group)

    def share_histogram(self, hist):
        """puts this rank's histogram (int64[65536] on the device) into its slot on every rank; after
        barrier() the slots of all ranks lie one after the other at hist_slots()"""
        dsts = [b[3] + 256 + self.rank * self.HIST_BYTES for b in self.bufs]
        self.wc.peer_put(dsts, hist.data_ptr(), self.HIST_BYTES)

    def hist_slots(self):
        return self.mine[3] + 256

    def close(self):
        for r, mapped in enumerate(self.bufs):
            if r != self.rank:
                for m in mapped:
                    self.wc.ipc_close(m)
        self.bufs = []


class DistributedWordCount:
    """run(shard) on every rank; afterwards rank 0's backend holds the sorted result of the
    whole input and the other ranks hold an empty one (ordered=True: the slice of their key range)."""

    def __init__(self, backend, dist, group=None, peers=None, ordered=False):
        self.b, self.dist, self.group, self.peers, self.ordered = backend, dist, group, peers, ordered
        self.rank = dist.get_rank(group)
        self.world = dist.get_world_size(group)
        self.last_shuffle_bytes = 0          # bytes this rank sent to other ranks in the hash shuffle

    def _exchange(self, send_sizes):
        """all-to-all of per-destination byte counts; returns what each source sends to me"""
        t_send = self.b.sizes_tensor(send_sizes)
        t_recv = self.b.sizes_tensor([0] * self.world)
        self.dist.all_to_all_single(t_recv, t_send, group=self.group)
        return [int(x) for x in t_recv.tolist()]

    def _shuffle(self, nparts_local, dest_of_part):
        """pack the local records into nparts_local parts, send part i to rank dest_of_part[i],
        merge everything received"""
        b = self.b
        sizes = b.shuffle_sizes(nparts_local)
        send_sizes = [0] * self.world
        for i, s in enumerate(sizes):
            send_sizes[dest_of_part[i]] += s
        offsets, acc = [], 0
        for s in sizes:
            offsets.append(acc)
            acc += s
        sendbuf = b.alloc(acc)
        b.shuffle_pack(nparts_local, sendbuf, offsets)
        recv_sizes = self._exchange(send_sizes)
        recvbuf = b.alloc(sum(recv_sizes))
        self.dist.all_to_all_single(recvbuf[:sum(recv_sizes)], sendbuf[:acc], output_split_sizes=recv_sizes,
                                    input_split_sizes=send_sizes, group=self.group)
        b.begin_merge()
        off = 0
        for s in recv_sizes:
            if s:
                b.merge(recvbuf, off, s)
            off += s
        return send_sizes

    def _shuffle_peers(self, which, dests):
        """fused path: part i is written by the packing kernel into rank dests[i]'s buffer
        `which` (0 hash shuffle, 1 gather), slot = my rank; barrier; merge my own slots"""
        px, wc = self.peers, self.b.wc
        if px.counted:
            # sizes first (host round trip; a table that overflowed in MAP grows here), largest part of any rank
            sizes = wc.shuffle_sizes(len(dests))
            n_short, long_bytes = wc.shuffle_counts(len(dests))
            t = self.b.sizes_tensor([max(n_short), max(long_bytes)])
            self.dist.all_reduce(t, op=self.dist.ReduceOp.MAX, group=self.group)
            px.ensure_shuffle(*[int(x) for x in t.tolist()])
            self.last_shuffle_bytes = sum(x for d, x in zip(dests, sizes) if d != self.rank)
        dst = [px.bufs[d][which] + self.rank * px.slot_bytes for d in dests]
        wc.shuffle_pack_peers(dst, px.slot_bytes, px.short_cap)
        px.barrier()
        wc.shuffle_begin_merge()
        nslots = self.world if (which == 0 or self.rank == 0) else 0
        for s in range(nslots):
            wc.shuffle_merge(px.mine[which] + s * px.slot_bytes, px.slot_bytes)

    def _push_results(self):
        """the same over peer memory: every rank packs its slice straight into its slot of rank 0's
        result buffer; barrier; rank 0 strings the slots together"""
        wc, px = self.b.wc, self.peers
        nbytes = wc.result_part_bytes()
        if px.counted:
            t = self.b.sizes_tensor([nbytes])
            self.dist.all_reduce(t, op=self.dist.ReduceOp.MAX, group=self.group)
            px.ensure_result(int(t.item()))
        dst = px.bufs[0][2] + self.rank * px.result_slot
        if nbytes <= px.result_slot:
            wc.result_part_pack(dst)
        else:
            wc.result_header(dst)                 # rank 0 sees the size and refuses the job as well
        px.barrier()
        if self.rank == 0:
            wc.result_adopt_slots(px.mine[2], px.result_slot, self.world)
        if nbytes > px.result_slot:
            raise RuntimeError("sorted slice of %d bytes does not fit its %d-byte slot on rank 0" % (nbytes, px.result_slot))

    def _gather_results(self):
        """every rank's sorted slice goes to rank 0, which strings them together in rank order"""
        b = self.b
        part, header = b.result_part()
        headers = b.sizes_tensor([0] * (4 * self.world))
        self.dist.all_gather_into_tensor(headers, header, group=self.group)
        headers = headers.tolist()
        headers = [headers[4 * r:4 * r + 4] for r in range(self.world)]
        sizes = [h[3] for h in headers]
        recv_sizes = sizes if self.rank == 0 else [0] * self.world
        send_sizes = [sizes[self.rank]] + [0] * (self.world - 1)
        recvbuf = b.alloc(sum(recv_sizes))
        self.dist.all_to_all_single(recvbuf[:sum(recv_sizes)], part, output_split_sizes=recv_sizes,
                                    input_split_sizes=send_sizes, group=self.group)
        if self.rank == 0:
            offsets = [sum(sizes[:r]) for r in range(self.world)]
            b.result_concat(recvbuf, offsets, headers)

    def run(self, shard):
        if hasattr(self.b, "on_stream"):
            with self.b.on_stream():
                return self._run(shard)
        return self._run(shard)

    def _run(self, shard):
        b = self.b
        b.reset()
        b.map(shard)
        if self.world > 1 and self.ordered:
            hist = b.histogram()
            if self.peers is not None:
                # no collective call at all: histograms, records and slices travel as peer stores
                self.peers.share_histogram(hist)
                self.peers.barrier()
                b.wc.partition_by_range_sum(self.peers.hist_slots(), self.world, self.world)
                self._shuffle_peers(0, list(range(self.world)))
            else:
                self.dist.all_reduce(hist, group=self.group)
                b.partition_by_range(hist, self.world)
                sent = self._shuffle(self.world, list(range(self.world)))
                self.last_shuffle_bytes = sum(s for r, s in enumerate(sent) if r != self.rank)
            b.process()
            b.reduce()
            if self.peers is not None:
                self._push_results()
            else:
                self._gather_results()
            return self
        if self.world > 1 and self.peers is not None:
            self._shuffle_peers(0, list(range(self.world)))
            self._shuffle_peers(1, [0])
        elif self.world > 1:
            # hash-partitioned exchange: part i goes to rank i
            sent = self._shuffle(self.world, list(range(self.world)))
            self.last_shuffle_bytes = sum(s for r, s in enumerate(sent) if r != self.rank)
            # gather: every rank's merged partition goes to rank 0 as one part
            self._shuffle(1, [0])
        b.process()
        b.reduce()
        return self
