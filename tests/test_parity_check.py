"""The checker used by bench.py and tools/check_full_configs.py (oracle/parity.py) is itself tested here,
on CPU: array-level equality with the list-level oracle API, and the range-by-range distributed check
over gloo at world size 3, on a correct result and on results with one planted fault each."""
import os

import numpy as np
import pytest

from conftest import golden_input
from locust_b200.shuffle import plan_shards
from oracle import parity, pyoracle

torch = pytest.importorskip("torch")


def _arrays(res):
    blob = np.frombuffer(b"".join(res.keys), dtype=np.uint8)
    offs = np.cumsum([0] + [len(k) for k in res.keys]).astype(np.uint64)
    return blob, offs, np.array(res.counts, dtype=np.uint64)


def test_oracle_arrays_equal_the_list_api():
    data = golden_input("hamlet") + golden_input("edge_long_tokens")
    blob, offs, counts, tokens = parity.oracle_arrays(np.frombuffer(data, dtype=np.uint8))
    ref = pyoracle.wordcount(data)
    assert tokens == ref.tokens
    assert parity.same_result((blob, offs, counts), _arrays(ref)) is None
    rows = parity.packed(blob, offs)
    assert parity._strictly_ascending(rows)
    # a planted fault is found and named
    bad = counts.copy(); bad[7] += 1
    assert "key 7" in parity.same_result((blob, offs, bad), (blob, offs, counts))
    assert parity.same_result((blob[:-1], offs, counts), (blob, offs, counts)) is not None
    flipped = blob.copy(); flipped[100] ^= 1
    assert "blob byte 100" in parity.same_result((flipped, offs, counts), (blob, offs, counts))
    e = parity.oracle_arrays(b"")
    assert len(e[2]) == 0 and e[3] == 0


def _worker(rank, world, port, data, fault, out_path):
    import torch.distributed as dist
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        lo, hi = plan_shards(data, world)[rank]
        shard = parity.oracle_arrays(np.frombuffer(data[lo:hi], dtype=np.uint8))[:3]
        blob, offs, counts, _ = parity.oracle_arrays(np.frombuffer(data, dtype=np.uint8))
        n = len(counts)
        cuts = [0, n // 2, n // 2, n]            # rank 1's range is empty on purpose
        counts = counts.copy()
        if fault == "count":
            counts[n - 3] += 1                   # lies in the last rank's slice
        a, b = cuts[rank], cuts[rank + 1]
        sl = (blob[int(offs[a]):int(offs[b])], offs[a:b + 1] - offs[a], counts[a:b])
        glob = None
        if rank == 0:
            vals = np.concatenate([[0], np.cumsum(counts)[:-1]]).astype(np.uint64)
            gcounts = counts.copy()
            if fault == "global":
                gcounts[n - 1] += 5              # rank 0's copy of the last slice differs from the slice itself
            if fault == "vals":
                vals[3] += 1
            glob = (blob, offs, gcounts, vals)
        res = parity.check_distributed(dist, torch.device("cpu"), rank, world, shard, sl, glob)
        if rank == 0:
            with open(out_path, "w") as f:
                f.write(repr(res))
    finally:
        dist.destroy_process_group()


@pytest.mark.parametrize("fault", ["none", "count", "global", "vals"])
def test_distributed_check_over_gloo(tmp_path, fault):
    import torch.multiprocessing as mp
    data = golden_input("hamlet")
    out = str(tmp_path / "res.txt")
    port = 29650 + ["none", "count", "global", "vals"].index(fault)
    mp.spawn(_worker, args=(3, port, data, fault, out), nprocs=3, join=True)
    res = eval(open(out).read())
    ref = pyoracle.wordcount(data)
    assert res["keys"] == len(ref.keys)
    if fault == "none":
        assert res["ok"] is True and res["tokens"] == ref.tokens and res["slices"][1] == 0
    else:
        assert res["ok"] is False and res["why"]
