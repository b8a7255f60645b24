"""TEST INFRASTRUCTURE -- exact comparison of word-count results with the CPU oracle at any size.

Used by bench.py (after its timed region, as the checker of what was just measured),
tools/check_full_configs.py and the GPU tests.  The product package (locust_b200) never imports it.

Everything works on numpy arrays (key blob, offsets, counts), not Python lists, so that BASELINE
config 4 (about 50 M unique keys) can be compared record for record:

  oracle_arrays(data)                 wc_oracle_fast on a host buffer -> (blob, offsets, counts, tokens)
  same_result(a, b)                   bit-exact equality of two (blob, offsets, counts) results
  check_distributed(...)              the result of a multi-rank job == the key-wise sum of the oracle's
                                      per-shard results, exactly, checked range by range on all ranks
"""
import ctypes

import numpy as np

from . import pyoracle


def oracle_arrays(data):
    """(key_blob uint8[], key_offsets uint64[n+1], counts uint64[n], tokens) of the oracle on `data`
    (numpy uint8 array or bytes); keys in ascending byte order."""
    L = pyoracle.lib()
    ptr, n, keep = pyoracle._buf_ptr(data)
    r = pyoracle._WcResult()
    rc = L.wc_oracle_fast(ptr, n, ctypes.byref(r))
    if rc != 0:
        raise RuntimeError("oracle failed: %d" % rc)
    try:
        nu = int(r.n)
        if nu == 0:
            return np.zeros(0, np.uint8), np.zeros(1, np.uint64), np.zeros(0, np.uint64), int(r.tokens)
        offs = np.ctypeslib.as_array(r.key_offsets, shape=(nu + 1,)).copy()
        counts = np.ctypeslib.as_array(r.counts, shape=(nu,)).copy()
        blob = np.ctypeslib.as_array(r.key_blob, shape=(int(offs[nu]),)).copy()
        return blob, offs, counts, int(r.tokens)
    finally:
        L.wc_result_free(ctypes.byref(r))


def same_result(got, want):
    """None if (blob, offsets, counts) are identical, else a one-line description of the first difference."""
    gb, go, gc = (np.asarray(x) for x in got[:3])
    wb, wo, wc = (np.asarray(x) for x in want[:3])
    if len(gc) != len(wc):
        return "unique keys: got %d, oracle %d" % (len(gc), len(wc))
    go, wo = go.astype(np.uint64), wo.astype(np.uint64)
    if len(go) != len(wo) or not np.array_equal(go, wo):
        d = np.flatnonzero(go[: len(wo)] != wo[: len(go)])
        return "key lengths differ first at key %d" % (max(int(d[0]) - 1, 0) if len(d) else min(len(go), len(wo)))
    if len(gb) != len(wb):
        return "key bytes: got %d, oracle %d" % (len(gb), len(wb))
    if not np.array_equal(gb, wb):
        i = int(np.flatnonzero(gb != wb)[0])
        return "key bytes differ first at blob byte %d (key %d)" % (i, int(np.searchsorted(wo, i, side="right")) - 1)
    gc, wc = gc.astype(np.uint64), wc.astype(np.uint64)
    if not np.array_equal(gc, wc):
        i = int(np.flatnonzero(gc != wc)[0])
        return "counts differ first at key %d: got %d, oracle %d" % (i, int(gc[i]), int(wc[i]))
    return None


def packed(blob, offs, width=None):
    """Keys as rows of big-endian uint64 columns (zero padded) plus a length column: integer order of
    (col 0, col 1, ..., length) is the byte order of the keys (no key holds a NUL byte).
    Returns a uint64 array [n, W + 1]."""
    offs = np.asarray(offs).astype(np.int64)
    lens = np.diff(offs)
    n = len(lens)
    maxlen = int(lens.max()) if n else 0
    W = max(1, (maxlen + 7) // 8) if width is None else width
    if maxlen > 8 * W:
        raise ValueError("key of %d bytes does not fit %d columns" % (maxlen, W))
    out = np.zeros((n, W + 1), dtype=np.uint64)
    if n:
        raw = np.zeros((n, 8 * W), dtype=np.uint8)
        idx = np.repeat(np.arange(n), lens)
        raw[idx, np.arange(len(blob)) - np.repeat(offs[:-1], lens)] = np.asarray(blob)
        out[:, :W] = raw.view(">u8").astype(np.uint64)
        out[:, W] = lens.astype(np.uint64)
    return out


def _ge(rows, key):
    """rows[i] >= key, lexicographically over the columns"""
    ge = np.ones(len(rows), dtype=bool)
    for c in range(rows.shape[1] - 1, -1, -1):
        ge = (rows[:, c] > key[c]) | ((rows[:, c] == key[c]) & ge)
    return ge


def _strictly_ascending(rows):
    if len(rows) < 2:
        return True
    gt = np.zeros(len(rows) - 1, dtype=bool)
    for c in range(rows.shape[1] - 1, -1, -1):
        gt = (rows[1:, c] > rows[:-1, c]) | ((rows[1:, c] == rows[:-1, c]) & gt)
    return bool(gt.all())


def sum_by_key(rows, counts):
    """sorted unique rows and the sum of counts per row"""
    if len(rows) == 0:
        return rows, counts
    order = np.lexsort(tuple(rows[:, c] for c in range(rows.shape[1] - 1, -1, -1)))
    rows, counts = rows[order], counts[order]
    new = np.ones(len(rows), dtype=bool)
    new[1:] = (rows[1:] != rows[:-1]).any(axis=1)
    starts = np.flatnonzero(new)
    return rows[starts], np.add.reduceat(counts.astype(np.uint64), starts)


def _checksum(rows, counts, base):
    """order-sensitive 63-bit checksum of a run of result records that starts at global index `base`"""
    if len(counts) == 0:
        return 0
    with np.errstate(over="ignore"):
        i = (np.arange(len(counts), dtype=np.uint64) + np.uint64(base)) * np.uint64(0x9E3779B97F4A7C15)
        h = counts.astype(np.uint64) * np.uint64(0xC2B2AE3D27D4EB4F) + i
        for c in range(rows.shape[1]):
            h = (h ^ (h >> np.uint64(29))) * np.uint64(0xBF58476D1CE4E5B9) + rows[:, c]
        return int(np.bitwise_xor.reduce(h ^ (h >> np.uint64(32)))) & 0x7FFFFFFFFFFFFFFF


def check_sum_of_shards(shard_oracles, global_result):
    """One process holds everything: the global (blob, offsets, counts, vals) equals the key-wise sum of the
    per-shard oracle results, is strictly ascending, and vals are the prefix sums.  Returns a dict."""
    W = 2
    for s in shard_oracles:
        if len(s[2]):
            W = max(W, (int(np.diff(np.asarray(s[1]).astype(np.int64)).max()) + 7) // 8)
    if len(global_result[2]):
        W = max(W, (int(np.diff(np.asarray(global_result[1]).astype(np.int64)).max()) + 7) // 8)
    rows = np.concatenate([packed(s[0], s[1], W) for s in shard_oracles], axis=0)
    counts = np.concatenate([np.asarray(s[2]).astype(np.uint64) for s in shard_oracles])
    want_rows, want_counts = sum_by_key(rows, counts)
    g_rows = packed(global_result[0], global_result[1], W)
    g_counts = np.asarray(global_result[2]).astype(np.uint64)
    res = {"oracle": "wc_oracle_fast on every shard, records added up by key", "keys": int(len(g_counts)), "ok": True}
    if len(want_counts) != len(g_counts):
        res.update(ok=False, why="unique keys: got %d, oracle %d" % (len(g_counts), len(want_counts)))
    elif len(g_counts) and not np.array_equal(want_rows, g_rows):
        res.update(ok=False, why="keys differ from the oracle's")
    elif len(g_counts) and not np.array_equal(want_counts, g_counts):
        i = int(np.flatnonzero(want_counts != g_counts)[0])
        res.update(ok=False, why="count of key %d is %d, oracle %d" % (i, int(g_counts[i]), int(want_counts[i])))
    elif not _strictly_ascending(g_rows):
        res.update(ok=False, why="keys are not strictly ascending")
    else:
        vals = np.asarray(global_result[3]).astype(np.uint64)
        want_vals = np.concatenate([[0], np.cumsum(g_counts)[:-1]]).astype(np.uint64) if len(g_counts) else vals
        if not np.array_equal(vals, want_vals):
            res.update(ok=False, why="vals are not the exclusive prefix sums of counts")
    if res["ok"]:
        res["tokens"] = int(g_counts.sum())
    return res


def check_distributed(dist, device, rank, world, shard_oracle, my_slice, global_result):
    """Exact parity of a multi-rank ordered job, run on every rank:

      shard_oracle   (blob, offsets, counts) of the oracle on this rank's shard alone
      my_slice       (blob, offsets, counts) this rank's context holds after the job (ranks > 0: the sorted
                     slice of their key range; rank 0: ignored, its slice is the head of the global result)
      global_result  rank 0: (blob, offsets, counts, vals) of the whole job; other ranks: None

    Every rank sends the oracle records that fall into rank r's key range to rank r (all-to-all over the
    process group), rank r adds them up by key and compares with its slice record for record; rank 0
    also checks that the global result is the slices in rank order (checksums), strictly ascending, and
    that vals are the exclusive prefix sums of counts.  Returns a dict on rank 0 (None elsewhere)."""
    import torch
    lens_max = torch.tensor([int(np.diff(np.asarray(shard_oracle[1]).astype(np.int64)).max()) if len(shard_oracle[2]) else 0],
                            dtype=torch.int64, device=device)
    dist.all_reduce(lens_max, op=dist.ReduceOp.MAX)
    W = max(2, (int(lens_max.item()) + 7) // 8)
    ref_rows = packed(shard_oracle[0], shard_oracle[1], W)
    ref_counts = np.asarray(shard_oracle[2]).astype(np.uint64)
    if rank == 0:
        g_rows = packed(global_result[0], global_result[1], W)
        g_counts = np.asarray(global_result[2]).astype(np.uint64)
    else:
        s_rows = packed(my_slice[0], my_slice[1], W)
        s_counts = np.asarray(my_slice[2]).astype(np.uint64)
    # slice sizes and first keys (ranks > 0); rank 0's slice is what is left at the head of the global result
    mine = np.zeros(W + 2, dtype=np.int64)
    if rank > 0:
        mine[0] = len(s_counts)
        if len(s_counts):
            mine[1:] = s_rows[0].view(np.int64)
    allm = torch.zeros(world * (W + 2), dtype=torch.int64, device=device)
    dist.all_gather_into_tensor(allm, torch.from_numpy(mine).to(device))
    allm = allm.cpu().numpy().reshape(world, W + 2)
    n_slice = allm[:, 0].copy()
    n_glob = torch.tensor([len(g_counts) if rank == 0 else 0], dtype=torch.int64, device=device)
    dist.broadcast(n_glob, 0)
    n_slice[0] = int(n_glob.item()) - int(n_slice[1:].sum())
    impossible = n_slice[0] < 0
    if rank == 0:
        s_rows, s_counts = g_rows[: max(int(n_slice[0]), 0)], g_counts[: max(int(n_slice[0]), 0)]
    # lower bound of rank r's range = first key of the first non-empty slice at or after r (none: nothing goes there)
    first = np.ascontiguousarray(allm[:, 1:]).view(np.uint64)
    dest = np.zeros(len(ref_counts), dtype=np.int64)
    nxt = None
    bounds = [None] * world
    for r in range(world - 1, 0, -1):
        if n_slice[r] > 0:
            nxt = first[r]
        bounds[r] = nxt
    for r in range(1, world):
        if bounds[r] is not None:
            dest += _ge(ref_rows, bounds[r])
    send_n = np.bincount(dest, minlength=world).astype(np.int64)
    t_send = torch.from_numpy(send_n).to(device)
    t_recv = torch.zeros(world, dtype=torch.int64, device=device)
    dist.all_to_all_single(t_recv, t_send)
    recv_n = t_recv.cpu().numpy()
    cols = W + 2
    payload = np.ascontiguousarray(np.concatenate([ref_rows, ref_counts[:, None]], axis=1)).view(np.int64)   # dest is monotone in the key
    sendbuf = torch.from_numpy(payload).to(device).reshape(-1)
    recvbuf = torch.empty(int(recv_n.sum()) * cols, dtype=torch.int64, device=device)
    dist.all_to_all_single(recvbuf, sendbuf, output_split_sizes=[int(x) * cols for x in recv_n],
                           input_split_sizes=[int(x) * cols for x in send_n])
    got = recvbuf.cpu().numpy().reshape(-1, cols).view(np.uint64)
    del sendbuf, recvbuf
    want_rows, want_counts = sum_by_key(got[:, : W + 1], got[:, W + 1])
    why = None
    if impossible:
        why = "slices hold more keys than the global result"
    elif len(want_counts) != len(s_counts):
        why = "rank %d: slice holds %d keys, the oracle's records of its range add up to %d keys" % (rank, len(s_counts), len(want_counts))
    elif len(s_counts) and not np.array_equal(want_rows, s_rows):
        why = "rank %d: keys of the slice differ from the oracle's" % rank
    elif len(s_counts) and not np.array_equal(want_counts, s_counts):
        i = int(np.flatnonzero(want_counts != s_counts)[0])
        why = "rank %d: count of key %d of the slice is %d, oracle %d" % (rank, i, int(s_counts[i]), int(want_counts[i]))
    # the global result is the slices one after the other
    base = int(n_slice[:rank].sum())
    chk = torch.tensor([_checksum(s_rows, s_counts, base), 0 if why is None else 1, len(s_counts)], dtype=torch.int64, device=device)
    allc = torch.zeros(3 * world, dtype=torch.int64, device=device)
    dist.all_gather_into_tensor(allc, chk)
    allc = allc.cpu().numpy().reshape(world, 3)
    if rank != 0:
        if why:
            import sys
            sys.stderr.write("parity: %s\n" % why)
        return None
    bad = [int(r) for r in range(world) if allc[r, 1]]
    res = {"oracle": "wc_oracle_fast on every rank's shard; records added up by key on the rank that owns the key range, "
                     "compared record for record with that rank's slice; global result = slices in rank order",
           "keys": int(len(g_counts)), "slices": [int(x) for x in n_slice], "ok": True}
    if bad:
        res.update(ok=False, why=why or "slice of rank(s) %s differs from the oracle (see stderr of that rank)" % bad)
        return res
    off = 0
    for r in range(world):
        seg = slice(off, off + int(n_slice[r]))
        if _checksum(g_rows[seg], g_counts[seg], off) != int(allc[r, 0]):
            res.update(ok=False, why="global result differs from slice %d (checksum)" % r)
            return res
        off += int(n_slice[r])
    if not _strictly_ascending(g_rows):
        res.update(ok=False, why="global keys are not strictly ascending")
        return res
    vals = np.asarray(global_result[3]).astype(np.uint64)
    want_vals = np.concatenate([[0], np.cumsum(g_counts)[:-1]]).astype(np.uint64) if len(g_counts) else vals
    if not np.array_equal(vals, want_vals):
        res.update(ok=False, why="vals are not the exclusive prefix sums of counts")
        return res
    res["tokens"] = int(g_counts.sum())
    return res
