"""The CUDA side of the multi-GPU shuffle (locust_shuffle_* through the C ABI) on ONE GPU:
P ranks are emulated by P contexts on cuda:0 and the all-to-all is done by handing each
packed part to its destination context.  The exchange logic itself (torch.distributed) is
covered on CPU in tests/test_shuffle_host.py and on real GPUs by bench.py --gpus N."""
import random

import pytest

from conftest import golden_input
from locust_b200.shuffle import plan_shards

pytestmark = pytest.mark.gpu

torch = pytest.importorskip("torch")


def dev_zeros(*a, **k):
    """zeroed CUDA tensor, finished: the contexts enqueue on their own streams, not on torch's"""
    t = torch.zeros(*a, **k)
    torch.cuda.synchronize()
    return t


@pytest.fixture(scope="module")
def lb():
    if not torch.cuda.is_available():
        pytest.skip("no CUDA device")
    import locust_b200
    locust_b200.load()
    return locust_b200


def emulate(lb, data, P, **kw):
    """word count of `data` over P emulated ranks; returns (keys, counts) of rank 0 and the
    per-rank partition sizes after the hash shuffle"""
    dev = torch.frombuffer(bytearray(data) if data else bytearray(1), dtype=torch.uint8)[:len(data)].cuda()
    shards = plan_shards(data, P)
    ctx = [lb.WordCount(device=0, **kw) for _ in range(P)]
    keep = []

    def exchange(nparts, dest_of_part, senders):
        # pack on every sender, then deliver part i of sender s to context dest_of_part[i]
        inbox = [[] for _ in range(P)]
        for s in senders:
            sizes = ctx[s].shuffle_sizes(nparts)
            assert all(x % 32 == 0 and x >= 32 for x in sizes)
            offs, acc = [], 0
            for x in sizes:
                offs.append(acc)
                acc += x
            buf = dev_zeros(acc, dtype=torch.uint8, device="cuda")
            ctx[s].shuffle_pack(nparts, buf.data_ptr(), offs)
            keep.append(buf)
            for i, (o, x) in enumerate(zip(offs, sizes)):
                inbox[dest_of_part[i]].append((buf, o, x))
        torch.cuda.synchronize()                         # every part packed before anybody merges (the ranks' barrier)
        for r in range(P):
            ctx[r].shuffle_begin_merge()
            for buf, o, x in inbox[r]:
                ctx[r].shuffle_merge(buf.data_ptr() + o, x, keepalive=buf)

    try:
        for r, (lo, hi) in enumerate(shards):
            ctx[r].reset()
            ctx[r].map(dev[lo:hi])
        exchange(P, list(range(P)), range(P))
        # partition sizes: every key lives on exactly one rank now
        exchange(1, [0], range(P))
        sizes = []
        for r in range(P):
            ctx[r].process()
            ctx[r].reduce()
            sizes.append(ctx[r].stats()["unique_keys"])
        keys, counts, _ = ctx[0].result()
        return keys, counts, sizes
    finally:
        for c in ctx:
            c.close()


@pytest.mark.parametrize("P", [1, 2, 3, 8])
def test_emulated_ranks_match_the_oracle(lb, P):
    from oracle import pyoracle
    data = golden_input("hamlet") + golden_input("edge_long_tokens") + b" " + golden_input("rand_longtok_s5")
    keys, counts, sizes = emulate(lb, data, P, table_log2=14)
    ref = pyoracle.wordcount(data)
    assert keys == ref.keys and counts == ref.counts
    assert sizes[0] == len(ref.keys) and all(s == 0 for s in sizes[1:])


def test_hash_partitions_are_disjoint_and_balanced(lb):
    from oracle import pyoracle
    rng = random.Random(1)
    words = [bytes(rng.choice(b"abcdefghij") for _ in range(rng.randint(1, 20))) for _ in range(30000)]
    data = b" ".join(rng.choice(words) for _ in range(200000))
    P = 4
    dev = torch.frombuffer(bytearray(data), dtype=torch.uint8).cuda()
    wc = lb.WordCount(device=0, table_log2=16)
    wc.reset()
    wc.map(dev)
    sizes = wc.shuffle_sizes(P)
    buf = dev_zeros(sum(sizes), dtype=torch.uint8, device="cuda")
    offs = [sum(sizes[:i]) for i in range(P)]
    wc.shuffle_pack(P, buf.data_ptr(), offs)
    torch.cuda.synchronize()
    raw = bytes(buf.cpu().numpy())
    import struct
    seen, total = {}, 0
    for p_, (o, x) in enumerate(zip(offs, sizes)):
        n_short, long_bytes, long_off, _ = struct.unpack_from("<4Q", raw, o)
        assert long_off == 32 + 32 * n_short
        assert (32 + 32 * n_short + long_bytes + 31) // 32 * 32 == x
        q = o + 32
        for _ in range(n_short):
            k = raw[q:q + 16].rstrip(b"\0")
            c = struct.unpack_from("<Q", raw, q + 16)[0]
            assert k not in seen
            seen[k] = (p_, c)
            q += 32
        end = q + long_bytes
        while q < end:
            ln, c = struct.unpack_from("<2Q", raw, q)
            k = raw[q + 16:q + 16 + ln]
            seen[k] = (p_, seen.get(k, (p_, 0))[1] + c)      # long keys: one record per occurrence
            q += 16 + (ln + 15) // 16 * 16
        total += n_short
    ref = pyoracle.wordcount(data).as_dict()
    assert {k: c for k, (_, c) in seen.items()} == ref
    per_part = [sum(1 for _, (p_, _) in seen.items() if p_ == i) for i in range(P)]
    assert min(per_part) > 0.8 * len(seen) / P                   # hash partitions are balanced
    wc.close()


def test_synthetic_c3_style_two_ranks(lb):
    from oracle import pyoracle
    from locust_b200.textgen import TextGen
    n = 32 << 20
    g = TextGen.config("C3", total_bytes=n, device=0, vocab=200000)
    d = torch.empty(n, dtype=torch.uint8, device="cuda")
    g.fill_device(d)
    torch.cuda.synchronize()
    host = bytes(d.cpu().numpy())
    keys, counts, _ = emulate(lb, host, 2, table_log2=19)
    ref = pyoracle.wordcount(host, mode="fast", formats=False)
    assert keys == ref.keys and counts == ref.counts


@pytest.mark.parametrize("P", [2, 5])
def test_pack_straight_into_peer_slots(lb, P):
    """locust_shuffle_pack_peers: the packing kernel writes every part into the destination
    rank's receive buffer itself (on a multi-GPU box those are NVLink peer addresses; here the
    ranks are contexts on one GPU, so no IPC mapping is needed).  Two jobs in a row: the
    slots of the first job must not leak into the second."""
    from oracle import pyoracle
    texts = [golden_input("hamlet") + golden_input("edge_long_tokens") + b" " + golden_input("rand_longtok_s5"),
             golden_input("rand_heavy_s4")]
    log2 = 14
    short_cap = 1 << log2
    slot_bytes = 32 + 32 * short_cap + (1 << 20)
    ctx = [lb.WordCount(device=0, table_log2=log2) for _ in range(P)]
    shuffle_buf = [c.exchange_alloc(slot_bytes * P) for c in ctx]
    gather_buf = ctx[0].exchange_alloc(slot_bytes * P)
    try:
        for data in texts:
            dev = torch.frombuffer(bytearray(data), dtype=torch.uint8).cuda()
            for r, (lo, hi) in enumerate(plan_shards(data, P)):
                ctx[r].reset()
                ctx[r].map(dev[lo:hi])
            for r in range(P):                                    # hash shuffle: part p -> rank p, slot r
                ctx[r].shuffle_pack_peers([shuffle_buf[p] + r * slot_bytes for p in range(P)], slot_bytes, short_cap)
            torch.cuda.synchronize()                              # stands in for the barrier over all ranks
            for r in range(P):
                ctx[r].shuffle_begin_merge()
                for s in range(P):
                    ctx[r].shuffle_merge(shuffle_buf[r] + s * slot_bytes, slot_bytes)
            for r in range(P):                                    # gather on rank 0
                ctx[r].shuffle_pack_peers([gather_buf + r * slot_bytes], slot_bytes, short_cap)
            torch.cuda.synchronize()
            for r in range(P):
                ctx[r].shuffle_begin_merge()
            for s in range(P):
                ctx[0].shuffle_merge(gather_buf + s * slot_bytes, slot_bytes)
            for r in range(P):
                ctx[r].process()
                ctx[r].reduce()
            keys, counts, _ = ctx[0].result()
            ref = pyoracle.wordcount(data)
            assert keys == ref.keys and counts == ref.counts
            assert all(ctx[r].stats()["unique_keys"] == 0 for r in range(1, P))
        # a slot that is too small for the long records is reported, not silently truncated
        tiny = 32 + 32 * short_cap + 64
        ctx[0].reset()
        ctx[0].map(torch.frombuffer(bytearray(texts[0]), dtype=torch.uint8).cuda())
        small = ctx[0].exchange_alloc(tiny)
        ctx[0].shuffle_pack_peers([small], tiny, short_cap)
        ctx[0].shuffle_begin_merge()
        ctx[0].shuffle_merge(small, tiny)
        ctx[0].process()
        with pytest.raises(lb.LocustError, match="did not fit"):
            ctx[0].reduce()
    finally:
        for c in ctx:
            c.close()


def emulate_ordered(lb, data, P, via="concat", room=None, hist_via="shared", **kw):
    """the ordered shuffle over P emulated ranks: parts are key ranges, every rank sorts its own
    range, rank 0 strings the slices together.  Returns rank 0's (keys, counts, vals) and the
    per-rank slices [(keys, counts)]."""
    dev = torch.frombuffer(bytearray(data) if data else bytearray(1), dtype=torch.uint8)[:len(data)].cuda()
    shards = plan_shards(data, P)
    ctx = [lb.WordCount(device=0, **kw) for _ in range(P)]
    keep = []
    try:
        hist = dev_zeros(lb.PARTITION_BINS, dtype=torch.int64, device="cuda")
        slots = dev_zeros(P * lb.PARTITION_BINS, dtype=torch.int64, device="cuda")
        own = [dev_zeros(lb.PARTITION_BINS, dtype=torch.int64, device="cuda") for _ in range(P)]
        torch.cuda.synchronize()
        for r, (lo, hi) in enumerate(shards):
            ctx[r].reset()
            ctx[r].map(dev[lo:hi])
            if hist_via == "shared":
                ctx[r].partition_histogram(hist.data_ptr())    # every rank adds into the same array = the all-reduce
            else:
                # what ranks on one box do: own histogram put into slot r (of every rank; here one copy serves all)
                ctx[r].partition_histogram(own[r].data_ptr())
                ctx[r].peer_put([slots.data_ptr() + 8 * lb.PARTITION_BINS * r], own[r].data_ptr(), 8 * lb.PARTITION_BINS)
        torch.cuda.synchronize()
        for r in range(P):
            if hist_via == "shared":
                ctx[r].partition_by_range(hist.data_ptr(), P)
            else:
                ctx[r].partition_by_range_sum(slots.data_ptr(), P, P)
        inbox = [[] for _ in range(P)]
        for s in range(P):
            sizes = ctx[s].shuffle_sizes(P)
            offs = [sum(sizes[:i]) for i in range(P)]
            buf = dev_zeros(sum(sizes), dtype=torch.uint8, device="cuda")
            torch.cuda.synchronize()
            ctx[s].shuffle_pack(P, buf.data_ptr(), offs)
            keep.append(buf)
            for i in range(P):
                inbox[i].append((buf, offs[i], sizes[i]))
        slices, parts = [], []
        torch.cuda.synchronize()                         # every part packed before anybody merges (the ranks' barrier)
        for r in range(P):
            ctx[r].shuffle_begin_merge()
            for buf, o, x in inbox[r]:
                ctx[r].shuffle_merge(buf.data_ptr() + o, x, keepalive=buf)
            ctx[r].process()
            ctx[r].reduce()
            k, c, _ = ctx[r].result()
            slices.append((k, c))
            part = dev_zeros(ctx[r].result_part_bytes(), dtype=torch.uint8, device="cuda")
            torch.cuda.synchronize()
            ctx[r].result_part_pack(part.data_ptr())
            parts.append(part)
        torch.cuda.synchronize()
        headers = [p[:32].view(torch.int64).tolist() for p in parts]
        assert all(h[3] == p.numel() for h, p in zip(headers, parts))
        if via == "slots":
            # what ranks on one box do: every slice packed into its slot of one buffer, strung together from there
            slot = room if room is not None else (max(p.numel() for p in parts) + 64) // 32 * 32
            landing = dev_zeros(slot * P, dtype=torch.uint8, device="cuda")
            torch.cuda.synchronize()
            for r in range(P):
                if parts[r].numel() <= slot:
                    ctx[r].result_part_pack(landing.data_ptr() + r * slot)
                else:
                    ctx[r].result_header(landing.data_ptr() + r * slot)
            torch.cuda.synchronize()
            ctx[0].result_adopt_slots(landing.data_ptr(), slot, P)
            keys, counts, vals = ctx[0].result()
            return keys, counts, vals, slices, ctx[0].stats()
        ctx[0].result_concat([p.data_ptr() for p in parts], [h[0] for h in headers], [h[1] for h in headers],
                             [h[2] for h in headers], keepalive=parts)
        keys, counts, vals = ctx[0].result()
        return keys, counts, vals, slices, ctx[0].stats()
    finally:
        for c in ctx:
            c.close()


@pytest.mark.parametrize("P", [1, 2, 3, 8])
def test_ordered_shuffle_matches_the_oracle(lb, P):
    from oracle import pyoracle
    data = golden_input("hamlet") + golden_input("edge_long_tokens") + b" " + golden_input("rand_longtok_s5")
    keys, counts, vals, slices, stats = emulate_ordered(lb, data, P, table_log2=14)
    ref = pyoracle.wordcount(data)
    assert keys == ref.keys and counts == ref.counts
    assert vals == [sum(ref.counts[:i]) for i in range(len(ref.counts))]
    assert stats["unique_keys"] == len(ref.keys) and stats["tokens"] == sum(ref.counts)
    # slices are consecutive key ranges in rank order, and no rank got everything (hamlet has many first letters)
    flat = [k for ks, _ in slices for k in ks]
    assert flat == ref.keys
    if P > 1:
        assert max(len(ks) for ks, _ in slices) < len(ref.keys)


@pytest.mark.parametrize("P", [1, 2, 8])
def test_ordered_shuffle_slices_in_slots(lb, P):
    from oracle import pyoracle
    data = golden_input("hamlet") + golden_input("edge_long_tokens") + b" " + golden_input("rand_longtok_s5")
    keys, counts, vals, _, stats = emulate_ordered(lb, data, P, via="slots", table_log2=14)
    ref = pyoracle.wordcount(data)
    assert keys == ref.keys and counts == ref.counts
    assert vals == [sum(ref.counts[:i]) for i in range(len(ref.counts))]
    assert stats["unique_keys"] == len(ref.keys) and stats["tokens"] == sum(ref.counts)
    # slots that are too small are refused, not half filled
    with pytest.raises(lb.LocustError):
        emulate_ordered(lb, data, P, via="slots", room=4096, table_log2=14)
    keys, counts, _, _, _ = emulate_ordered(lb, b"", P, via="slots", table_log2=10)
    assert keys == [] and counts == []


def test_ordered_shuffle_every_first_byte_and_tiny_inputs(lb):
    from oracle import pyoracle
    rng = random.Random(11)
    alphabet = [bytes([b]) for b in range(1, 256) if bytes([b]) not in b" ,.-;:'()\"\t\n"]
    words = [b"".join(rng.choice(alphabet) for _ in range(rng.choice((1, 1, 2, 3, 9, 17, 40)))) for _ in range(3000)]
    data = b" ".join(rng.choice(words) for _ in range(20000))
    keys, counts, _, slices, _ = emulate_ordered(lb, data, 4, table_log2=13)
    ref = pyoracle.wordcount(data)
    assert keys == ref.keys and counts == ref.counts
    assert [k for ks, _ in slices for k in ks] == ref.keys
    # fewer keys than ranks, and no keys at all
    for tiny in (b"b a", b"", b"same same same"):
        keys, counts, _, _, _ = emulate_ordered(lb, tiny, 4, table_log2=10)
        ref = pyoracle.wordcount(tiny)
        assert keys == ref.keys and counts == ref.counts


@pytest.mark.parametrize("P", [2, 8])
def test_ordered_shuffle_histograms_in_slots(lb, P):
    from oracle import pyoracle
    data = golden_input("hamlet") + b" " + golden_input("rand_longtok_s5")
    keys, counts, _, slices, _ = emulate_ordered(lb, data, P, hist_via="slots", table_log2=14)
    ref = pyoracle.wordcount(data)
    assert keys == ref.keys and counts == ref.counts
    same, _, _, slices2, _ = emulate_ordered(lb, data, P, table_log2=14)
    assert [len(k) for k, _ in slices] == [len(k) for k, _ in slices2]      # the same ranges either way


def test_peer_barrier_alone_and_put_to_many(lb):
    """world = 1: the barrier only meets itself (on one GPU a kernel must not wait for another launch);
    the two-rank case runs on real GPUs in bench.py --gpus 2 and tools/check_peer_exchange.py"""
    c = lb.WordCount(device=0, table_log2=10)
    try:
        flags = dev_zeros(64, dtype=torch.int32, device="cuda")
        src = torch.arange(4096, dtype=torch.int32, device="cuda")
        dst = dev_zeros(3, 4096, dtype=torch.int32, device="cuda")
        torch.cuda.synchronize()
        c.reset()
        c.map(torch.frombuffer(bytearray(b"a b"), dtype=torch.uint8).cuda())
        for epoch in (1, 2, 3):
            c.peer_barrier([flags.data_ptr()], 0, 1, epoch)
        c.peer_put([dst[i].data_ptr() for i in range(3)], src.data_ptr(), 4096 * 4)
        c.process()
        c.reduce()                                   # would report a barrier that timed out
        assert int(flags[0].item()) == 3
        assert bool((dst == src.unsqueeze(0)).all())
        with pytest.raises(lb.LocustError):
            c.peer_put([dst.data_ptr() + 4], src.data_ptr(), 64)          # unaligned destination
    finally:
        c.close()


def test_range_parts_need_matching_nparts(lb):
    c = lb.WordCount(device=0, table_log2=10)
    try:
        d = torch.frombuffer(bytearray(b"a b c d"), dtype=torch.uint8).cuda()
        c.reset()
        c.map(d)
        hist = dev_zeros(lb.PARTITION_BINS, dtype=torch.int64, device="cuda")
        torch.cuda.synchronize()
        c.partition_histogram(hist.data_ptr())
        c.partition_by_range(hist.data_ptr(), 4)
        with pytest.raises(lb.LocustError):
            c.shuffle_sizes(2)
        assert len(c.shuffle_sizes(4)) == 4
        c.partition_by_range(0, 0)                 # back to hash parts
        assert len(c.shuffle_sizes(2)) == 2
    finally:
        c.close()


def test_distributed_wordcount_on_the_contexts_own_stream(lb):
    """DistributedWordCount + GpuBackend with a default-constructed WordCount (private stream): the
    collectives between pack and merge must be ordered against that stream, not torch's default one.
    One rank (one GPU here), so the exchange is driven by hand: _shuffle packs, runs the NCCL
    all-to-alls and merges, exactly as a multi-rank job does."""
    import torch.distributed as dist
    from oracle import pyoracle
    from locust_b200.shuffle import DistributedWordCount, GpuBackend
    made = False
    if not dist.is_initialized():
        dist.init_process_group("nccl", init_method="tcp://127.0.0.1:29591", rank=0, world_size=1,
                                device_id=torch.device("cuda:0"))
        made = True
    try:
        data = (golden_input("hamlet") + b" ") * 40
        ref = pyoracle.wordcount(data)
        dev = torch.frombuffer(bytearray(data), dtype=torch.uint8).cuda()
        torch.cuda.synchronize()
        for ordered in (False, True):
            wc = lb.WordCount(device=0, table_log2=14)           # no stream given
            b = GpuBackend(wc, torch.device("cuda:0"))
            assert b.stream.cuda_stream == wc.stream_ptr() != 0
            dwc = DistributedWordCount(b, dist, ordered=ordered)
            for _ in range(3):
                with b.on_stream():
                    b.reset()
                    b.map(dev)
                    if ordered:
                        hist = b.histogram()
                        dist.all_reduce(hist)
                        b.partition_by_range(hist, 1)
                    dwc._shuffle(1, [0])
                    b.process()
                    b.reduce()
                    if ordered:
                        dwc._gather_results()
                keys, counts, _ = wc.result()
                assert keys == ref.keys and counts == ref.counts
                dwc.run(dev)                                        # world 1: map, process, reduce
                keys, counts, _ = wc.result()
                assert keys == ref.keys and counts == ref.counts
            wc.close()
    finally:
        if made:
            dist.destroy_process_group()
