"""locust_b200 -- host-side mirror of Locust's MapReduce word-count operator surface.

The product is liblocust.so (hand-written sm_100a CUDA behind the C ABI of
include/locust.h); this module only binds it with ctypes and keeps the reference's
vocabulary: a job runs map -> process -> reduce (MapReduce/src/main.cu:397-468) and
yields (key, count) records in strcmp order.

There is no CPU path here: if the library or a CUDA device is missing every call
raises LocustError.
"""
import ctypes
import os

__all__ = ["LocustError", "WordCount", "lib_path", "load", "HOST", "DEVICE",
           "F_GENERIC_TOKENISER", "F_NO_SMEM_COMBINER", "F_WEAK_HASH", "F_SORT_PATH"]

_HERE = os.path.dirname(os.path.abspath(__file__))
HOST, DEVICE = 0, 1
STAGE_MAP, STAGE_PROCESS, STAGE_REDUCE = 0, 1, 2
F_GENERIC_TOKENISER, F_NO_SMEM_COMBINER, F_WEAK_HASH, F_SORT_PATH = 1, 2, 4, 8

EXPORTS = [
    "locust_version", "locust_create", "locust_destroy", "locust_reset", "locust_map",
    "locust_process", "locust_reduce", "locust_wordcount", "locust_pagerank", "locust_pagerank_stats", "locust_result_host",
    "locust_result_device", "locust_format_result", "locust_stage_ms", "locust_stats",
    "locust_last_error", "locust_stream", "locust_shuffle_sizes", "locust_shuffle_pack",
    "locust_shuffle_begin_merge", "locust_shuffle_merge", "locust_count_static",
    "locust_map_intermediate", "locust_exchange_alloc", "locust_ipc_export", "locust_ipc_open", "locust_ipc_close",
    "locust_shuffle_pack_peers", "locust_partition_histogram", "locust_partition_by_range",
    "locust_result_part_bytes", "locust_result_part_pack", "locust_result_concat",
    "locust_result_header", "locust_result_adopt_slots", "locust_peer_put", "locust_partition_by_range_sum",
    "locust_peer_barrier", "locust_shuffle_counts", "locust_exchange_free", "locust_wait_input",
    "locust_multi_create", "locust_multi_destroy", "locust_multi_wordcount", "locust_multi_wordcount_device",
    "locust_multi_context", "locust_multi_stats", "locust_multi_last_error",
]
PARTITION_BINS = 65536
STATIC_MAX_KEYS = 16384
STATIC_MAX_BYTES = 1 << 32


class LocustError(RuntimeError):
    pass


class _Config(ctypes.Structure):
    _fields_ = [("device", ctypes.c_int32), ("table_log2", ctypes.c_uint32), ("flags", ctypes.c_uint32),
                ("reserved", ctypes.c_uint32), ("long_capacity", ctypes.c_uint64), ("stream", ctypes.c_void_p)]


def lib_path():
    # LOCUST_LIB: development switch, another build of the same library (kernel experiments)
    return os.environ.get("LOCUST_LIB") or os.path.join(_HERE, "liblocust.so")


_lib = None


def load():
    """Loads liblocust.so (built in-tree by __graft_entry__.build()).  Fails loudly."""
    global _lib
    if _lib is not None:
        return _lib
    p = lib_path()
    if not os.path.exists(p):
        raise LocustError("liblocust.so is not built: run `python -c 'import __graft_entry__ as g; g.build()'`")
    L = ctypes.CDLL(p)
    vp, u64, u32 = ctypes.c_void_p, ctypes.c_uint64, ctypes.c_uint32
    pp = ctypes.POINTER
    L.locust_version.restype = ctypes.c_char_p
    L.locust_create.argtypes = [pp(_Config), pp(vp)]
    L.locust_destroy.argtypes = [vp]
    L.locust_reset.argtypes = [vp]
    L.locust_map.argtypes = [vp, vp, u64, ctypes.c_int]
    L.locust_process.argtypes = [vp]
    L.locust_wait_input.argtypes = [vp]
    L.locust_reduce.argtypes = [vp]
    L.locust_wordcount.argtypes = [vp, vp, u64, ctypes.c_int]
    L.locust_pagerank.argtypes = [vp, vp, u64, ctypes.c_int, ctypes.c_uint32, u64]
    L.locust_pagerank_stats.argtypes = [vp, pp(u64), pp(u64)]
    L.locust_result_host.argtypes = [vp, pp(vp), pp(vp), pp(vp), pp(vp), pp(u64)]
    L.locust_result_device.argtypes = [vp, pp(vp), pp(vp), pp(vp), pp(vp), pp(u64)]
    L.locust_format_result.argtypes = [vp, ctypes.c_int, pp(vp), pp(u64)]
    L.locust_stage_ms.argtypes = [vp, ctypes.c_int]
    L.locust_stage_ms.restype = ctypes.c_float
    L.locust_stats.argtypes = [vp, pp(u64), pp(u64), pp(u64)]
    L.locust_stream.argtypes = [vp]
    L.locust_stream.restype = ctypes.c_void_p
    L.locust_last_error.argtypes = [vp]
    L.locust_last_error.restype = ctypes.c_char_p
    L.locust_shuffle_sizes.argtypes = [vp, u32, pp(u64)]
    L.locust_shuffle_counts.argtypes = [vp, u32, pp(u64), pp(u64)]
    L.locust_multi_create.argtypes = [pp(ctypes.c_int), ctypes.c_int, vp, pp(vp)]
    L.locust_multi_destroy.argtypes = [vp]
    L.locust_multi_wordcount.argtypes = [vp, vp, u64]
    L.locust_multi_wordcount_device.argtypes = [vp, pp(vp), pp(u64)]
    L.locust_multi_context.argtypes = [vp, ctypes.c_int]
    L.locust_multi_context.restype = vp
    L.locust_multi_stats.argtypes = [vp, pp(ctypes.c_float), pp(u64)]
    L.locust_multi_last_error.argtypes = [vp]
    L.locust_multi_last_error.restype = ctypes.c_char_p
    L.locust_shuffle_pack.argtypes = [vp, u32, vp, pp(u64)]
    L.locust_shuffle_begin_merge.argtypes = [vp]
    L.locust_shuffle_merge.argtypes = [vp, vp, u64]
    L.locust_map_intermediate.argtypes = [vp, ctypes.c_char_p, u64]
    L.locust_exchange_alloc.argtypes = [vp, u64, pp(vp)]
    L.locust_exchange_free.argtypes = [vp, vp]
    L.locust_ipc_export.argtypes = [vp, vp, ctypes.c_char_p]
    L.locust_ipc_open.argtypes = [vp, ctypes.c_char_p, pp(vp)]
    L.locust_ipc_close.argtypes = [vp, vp]
    L.locust_shuffle_pack_peers.argtypes = [vp, u32, pp(vp), u64, u64]
    L.locust_count_static.argtypes = [vp, u64, vp, u64, pp(u64), vp, ctypes.c_char_p, ctypes.c_size_t]
    L.locust_partition_histogram.argtypes = [vp, vp]
    L.locust_partition_by_range.argtypes = [vp, vp, u32]
    L.locust_result_part_bytes.argtypes = [vp, pp(u64)]
    L.locust_result_part_pack.argtypes = [vp, vp]
    L.locust_result_concat.argtypes = [vp, u32, pp(vp), pp(u64), pp(u64), pp(u64)]
    L.locust_result_header.argtypes = [vp, vp]
    L.locust_result_adopt_slots.argtypes = [vp, vp, u64, u32]
    L.locust_peer_put.argtypes = [vp, u32, pp(vp), vp, u64]
    L.locust_partition_by_range_sum.argtypes = [vp, vp, u32, u32]
    L.locust_peer_barrier.argtypes = [vp, pp(vp), u32, u32, u32]
    _lib = L
    return L


def count_static(dev_data_ptr, nbytes, dev_records_ptr, max_records, stream=None):
    """locust_count_static: tokenise + combine without any allocation, for callers that cannot
    obtain device memory (see include/locust.h).  Works on raw device addresses of the current
    device; returns the number of 32-byte {k0, k1, count, 0} records written to dev_records."""
    L = load()
    n = ctypes.c_uint64()
    err = ctypes.create_string_buffer(256)
    rc = L.locust_count_static(dev_data_ptr, nbytes, dev_records_ptr, max_records, ctypes.byref(n), stream, err, 256)
    if rc != 0:
        e = LocustError("locust_count_static failed (%d): %s" % (rc, err.value.decode()))
        e.code, e.needed = rc, n.value
        raise e
    return n.value


def _ptr_of(data):
    """(address, nbytes, where, keepalive) for bytes-like, numpy arrays and torch tensors."""
    if hasattr(data, "data_ptr"):                       # torch tensor (uint8, contiguous)
        if not data.is_contiguous():
            raise LocustError("input tensor must be contiguous")
        n = data.numel() * data.element_size()
        return data.data_ptr(), n, (DEVICE if data.is_cuda else HOST), data
    if hasattr(data, "ctypes") and hasattr(data, "nbytes"):   # numpy
        return data.ctypes.data, data.nbytes, HOST, data
    b = bytes(data)
    return ctypes.cast(ctypes.c_char_p(b), ctypes.c_void_p).value or 0, len(b), HOST, b


class WordCount:
    """One word-count job context on one GPU: the three stages of the reference's
    main() as methods, results as (keys, counts)."""

    def __init__(self, device=0, table_log2=0, flags=0, long_capacity=0, stream=None):
        L = load()
        cfg = _Config(device, table_log2, flags, 0, long_capacity, stream)
        h = ctypes.c_void_p()
        rc = L.locust_create(ctypes.byref(cfg), ctypes.byref(h))
        self._L, self._h = L, h
        self._keep = []
        if rc != 0:
            msg = L.locust_last_error(h).decode() if h else "no usable CUDA device"
            if h:
                L.locust_destroy(h)
            self._h = None
            raise LocustError("locust_create failed (%d): %s" % (rc, msg))

    # -- plumbing -------------------------------------------------------------
    def _check(self, rc, what):
        if rc != 0:
            raise LocustError("%s failed (%d): %s" % (what, rc, self._L.locust_last_error(self._h).decode()))

    def close(self):
        if getattr(self, "_h", None):
            self._L.locust_destroy(self._h)
            self._h = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    def __enter__(self):
        return self

    def __exit__(self, *a):
        self.close()

    def stream_ptr(self):
        """the cudaStream_t (as an int) this context enqueues its work on"""
        return self._L.locust_stream(self._h) or 0

    # -- stages (main.cu:406, :411-415, :447-465) -------------------------------
    def reset(self):
        self._check(self._L.locust_reset(self._h), "locust_reset")
        self._keep = []

    def map(self, data):
        ptr, n, where, keep = _ptr_of(data)
        self._keep.append(keep)            # device inputs must outlive reduce()
        self._check(self._L.locust_map(self._h, ptr, n, where), "locust_map")

    def map_intermediate(self, text):
        """adds "key \\tvalue\\n" lines (the reference's /tmp/out.txt format, main.cu:121) to the job"""
        b = bytes(text)
        self._check(self._L.locust_map_intermediate(self._h, b, len(b)), "locust_map_intermediate")

    def process(self):
        self._check(self._L.locust_process(self._h), "locust_process")

    def reduce(self):
        self._check(self._L.locust_reduce(self._h), "locust_reduce")

    def run(self, data):
        """reset + map + process + reduce in one call (locust_wordcount)."""
        ptr, n, where, keep = _ptr_of(data)
        self._keep = [keep]                 # device inputs must outlive reduce()
        self._check(self._L.locust_wordcount(self._h, ptr, n, where), "locust_wordcount")
        return self

    def pagerank(self, edges, iterations=20, damping=0.85):
        """Second job (locust_pagerank): `edges` = text with one "src dst" pair per line.  Ranks are 64-bit
        fixed point with a total mass of 2^48; damping may be a float or an int already scaled by 2^32.
        Read the result with result() / result_arrays(): keys = node names, counts = ranks."""
        d = damping if isinstance(damping, int) else int(damping * 4294967296.0)
        ptr, n, where, keep = _ptr_of(edges)
        self._keep = [keep]
        self._check(self._L.locust_pagerank(self._h, ptr, n, where, int(iterations), d), "locust_pagerank")
        return self

    def pagerank_stats(self):
        u64 = ctypes.c_uint64
        nn, ne = u64(), u64()
        self._check(self._L.locust_pagerank_stats(self._h, ctypes.byref(nn), ctypes.byref(ne)), "locust_pagerank_stats")
        return {"nodes": nn.value, "edges": ne.value}

    # -- results ------------------------------------------------------------------
    def result(self):
        """(keys: list[bytes] in strcmp order, counts: list[int], vals: list[int])."""
        vp, u64 = ctypes.c_void_p, ctypes.c_uint64
        blob, offs, cnts, vals, n = vp(), vp(), vp(), vp(), u64()
        self._check(self._L.locust_result_host(self._h, ctypes.byref(blob), ctypes.byref(offs), ctypes.byref(cnts),
                                               ctypes.byref(vals), ctypes.byref(n)), "locust_result_host")
        n = n.value
        o = (u64 * (n + 1)).from_address(offs.value)
        offsets = list(o)
        b = ctypes.string_at(blob.value, offsets[n]) if n else b""
        keys = [b[offsets[i]:offsets[i + 1]] for i in range(n)]
        counts = list((u64 * n).from_address(cnts.value)) if n else []
        values = list((u64 * n).from_address(vals.value)) if n else []
        return keys, counts, values

    def result_arrays(self):
        """numpy views (copied) of blob, offsets, counts, vals -- for large results."""
        import numpy as np
        vp, u64 = ctypes.c_void_p, ctypes.c_uint64
        blob, offs, cnts, vals, n = vp(), vp(), vp(), vp(), u64()
        self._check(self._L.locust_result_host(self._h, ctypes.byref(blob), ctypes.byref(offs), ctypes.byref(cnts),
                                               ctypes.byref(vals), ctypes.byref(n)), "locust_result_host")
        n = n.value
        offsets = np.ctypeslib.as_array((u64 * (n + 1)).from_address(offs.value)).copy()
        nb = int(offsets[n])
        b = np.ctypeslib.as_array((ctypes.c_uint8 * max(nb, 1)).from_address(blob.value))[:nb].copy()
        c = np.ctypeslib.as_array((u64 * max(n, 1)).from_address(cnts.value))[:n].copy()
        v = np.ctypeslib.as_array((u64 * max(n, 1)).from_address(vals.value))[:n].copy()
        return b, offsets, c, v

    def device_result(self):
        """(blob_ptr, offsets_ptr, counts_ptr, vals_ptr, n) -- raw device addresses."""
        vp, u64 = ctypes.c_void_p, ctypes.c_uint64
        blob, offs, cnts, vals, n = vp(), vp(), vp(), vp(), u64()
        self._check(self._L.locust_result_device(self._h, ctypes.byref(blob), ctypes.byref(offs), ctypes.byref(cnts),
                                                 ctypes.byref(vals), ctypes.byref(n)), "locust_result_device")
        return blob.value, offs.value, cnts.value, vals.value, n.value

    def format(self, cpu_format=False, intermediate=False):
        """The reference's `print key:` lines (main.cu:132, or :286 with cpu_format), or the
        intermediate "%s \\t%d\\n" lines of multi mode (main.cu:121)."""
        t, nb = ctypes.c_void_p(), ctypes.c_uint64()
        fmt = 2 if intermediate else (1 if cpu_format else 0)
        self._check(self._L.locust_format_result(self._h, fmt, ctypes.byref(t), ctypes.byref(nb)),
                    "locust_format_result")
        return ctypes.string_at(t.value, nb.value) if nb.value else b""

    def stage_ms(self):
        return tuple(float(self._L.locust_stage_ms(self._h, s)) for s in (0, 1, 2))

    def shuffle_ms(self):
        """device time from the first shuffle_sizes to the last shuffle_merge of the job (-1 if none)"""
        return float(self._L.locust_stage_ms(self._h, 3))

    def stats(self):
        u64 = ctypes.c_uint64
        t, u, k = u64(), u64(), u64()
        self._check(self._L.locust_stats(self._h, ctypes.byref(t), ctypes.byref(u), ctypes.byref(k)), "locust_stats")
        return {"tokens": t.value, "unique_keys": u.value, "kernel_launches": k.value}

    # -- multi-GPU shuffle ---------------------------------------------------------
    def shuffle_sizes(self, nparts):
        arr = (ctypes.c_uint64 * nparts)()
        self._check(self._L.locust_shuffle_sizes(self._h, nparts, arr), "locust_shuffle_sizes")
        return list(arr)

    def shuffle_counts(self, nparts):
        """(short records per part, bytes of long records per part) of the last shuffle_sizes call"""
        a, b = (ctypes.c_uint64 * nparts)(), (ctypes.c_uint64 * nparts)()
        self._check(self._L.locust_shuffle_counts(self._h, nparts, a, b), "locust_shuffle_counts")
        return list(a), list(b)

    def shuffle_pack(self, nparts, dev_out_ptr, part_offsets):
        arr = (ctypes.c_uint64 * nparts)(*part_offsets)
        self._check(self._L.locust_shuffle_pack(self._h, nparts, dev_out_ptr, arr), "locust_shuffle_pack")

    # -- pack + send in one kernel over peer memory ---------------------------------------
    def exchange_alloc(self, nbytes):
        """zeroed receive buffer on this context's device (freed with the context); device address"""
        p = ctypes.c_void_p()
        self._check(self._L.locust_exchange_alloc(self._h, nbytes, ctypes.byref(p)), "locust_exchange_alloc")
        return p.value

    def exchange_free(self, dev_ptr):
        self._check(self._L.locust_exchange_free(self._h, dev_ptr), "locust_exchange_free")

    def ipc_export(self, dev_ptr):
        h = ctypes.create_string_buffer(64)
        self._check(self._L.locust_ipc_export(self._h, dev_ptr, h), "locust_ipc_export")
        return h.raw

    def ipc_open(self, handle):
        p = ctypes.c_void_p()
        self._check(self._L.locust_ipc_open(self._h, bytes(handle), ctypes.byref(p)), "locust_ipc_open")
        return p.value

    def ipc_close(self, dev_ptr):
        self._check(self._L.locust_ipc_close(self._h, dev_ptr), "locust_ipc_close")

    def shuffle_pack_peers(self, dst_slots, slot_bytes, short_cap):
        arr = (ctypes.c_void_p * len(dst_slots))(*dst_slots)
        self._check(self._L.locust_shuffle_pack_peers(self._h, len(dst_slots), arr, slot_bytes, short_cap),
                    "locust_shuffle_pack_peers")

    def partition_histogram(self, dev_hist_ptr):
        """adds this device's keys to a uint64[PARTITION_BINS] histogram in device memory"""
        self._check(self._L.locust_partition_histogram(self._h, dev_hist_ptr), "locust_partition_histogram")

    def partition_by_range(self, dev_hist_ptr, nparts):
        """from the histogram summed over all ranks: later shuffle calls split by key range (0 parts = by hash)"""
        self._check(self._L.locust_partition_by_range(self._h, dev_hist_ptr, nparts), "locust_partition_by_range")

    def result_part_bytes(self):
        n = ctypes.c_uint64()
        self._check(self._L.locust_result_part_bytes(self._h, ctypes.byref(n)), "locust_result_part_bytes")
        return n.value

    def result_part_pack(self, dev_out_ptr):
        self._check(self._L.locust_result_part_pack(self._h, dev_out_ptr), "locust_result_part_pack")

    def result_concat(self, dev_part_ptrs, n_keys, blob_bytes, totals, keepalive=None):
        if keepalive is not None:
            self._keep.append(keepalive)
        k = len(dev_part_ptrs)
        u64s = ctypes.c_uint64 * k
        self._check(self._L.locust_result_concat(self._h, k, (ctypes.c_void_p * k)(*dev_part_ptrs), u64s(*n_keys),
                                                 u64s(*blob_bytes), u64s(*totals)), "locust_result_concat")

    def result_header(self, dev_header_ptr):
        self._check(self._L.locust_result_header(self._h, dev_header_ptr), "locust_result_header")

    def peer_put(self, dsts, dev_src_ptr, nbytes):
        arr = (ctypes.c_void_p * len(dsts))(*dsts)
        self._check(self._L.locust_peer_put(self._h, len(dsts), arr, dev_src_ptr, nbytes), "locust_peer_put")

    def partition_by_range_sum(self, dev_hists_ptr, nhists, nparts):
        self._check(self._L.locust_partition_by_range_sum(self._h, dev_hists_ptr, nhists, nparts),
                    "locust_partition_by_range_sum")

    def peer_barrier(self, flag_ptrs, rank, world, epoch):
        arr = (ctypes.c_void_p * len(flag_ptrs))(*flag_ptrs)
        self._check(self._L.locust_peer_barrier(self._h, arr, rank, world, epoch), "locust_peer_barrier")

    def result_adopt_slots(self, dev_slots_ptr, slot_bytes, nslots):
        self._check(self._L.locust_result_adopt_slots(self._h, dev_slots_ptr, slot_bytes, nslots),
                    "locust_result_adopt_slots")

    def shuffle_begin_merge(self):
        self._check(self._L.locust_shuffle_begin_merge(self._h), "locust_shuffle_begin_merge")

    def shuffle_merge(self, dev_ptr, nbytes, keepalive=None):
        if keepalive is not None:
            self._keep.append(keepalive)
        self._check(self._L.locust_shuffle_merge(self._h, dev_ptr, nbytes), "locust_shuffle_merge")


class MultiWordCount:
    """locust_multi_*: one job over several GPUs of one box, driven from this process.  The reference's
    multi mode (main.cu:367-385, :421-446; Distributor/slave.py) as one call; the result is read through
    the rank-0 context (result(), result_arrays(), format(), stats() as on WordCount)."""

    def __init__(self, devices, table_log2=0, flags=0, long_capacity=0):
        L = load()
        self._L = L
        devs = (ctypes.c_int * len(devices))(*devices)
        cfg = _Config(0, table_log2, flags, 0, long_capacity, None)
        h = ctypes.c_void_p()
        rc = L.locust_multi_create(devs, len(devices), ctypes.byref(cfg), ctypes.byref(h))
        self._h, self.devices = h, list(devices)
        if rc != 0:
            msg = L.locust_multi_last_error(h).decode() if h else "no handle"
            if h:
                L.locust_multi_destroy(h)
            self._h = None
            raise LocustError("locust_multi_create failed (%d): %s" % (rc, msg))
        self._keep = None

    def _check(self, rc, what):
        if rc != 0:
            e = LocustError("%s failed (%d): %s" % (what, rc, self._L.locust_multi_last_error(self._h).decode()))
            e.code = rc
            raise e

    def context(self, rank=0):
        """a WordCount view of rank's context (borrowed: closing it does nothing)"""
        wc = WordCount.__new__(WordCount)
        wc._L, wc._h, wc._keep = self._L, ctypes.c_void_p(self._L.locust_multi_context(self._h, rank)), []
        wc.close = lambda: None
        return wc

    def run(self, data):
        """host bytes (bytes, numpy uint8 array, pinned CPU tensor): cut on delimiters, one shard per GPU"""
        ptr, n, where, keep = _ptr_of(data)
        if where != HOST:
            raise LocustError("MultiWordCount.run takes host memory; use run_shards for device-resident shards")
        self._keep = keep
        self._check(self._L.locust_multi_wordcount(self._h, ptr, n), "locust_multi_wordcount")
        return self.context(0)

    def run_shards(self, tensors):
        """one contiguous CUDA uint8 tensor per rank, already on that rank's device"""
        n = len(self.devices)
        assert len(tensors) == n
        ptrs = (ctypes.c_void_p * n)(*[t.data_ptr() for t in tensors])
        sizes = (ctypes.c_uint64 * n)(*[t.numel() for t in tensors])
        self._keep = list(tensors)
        self._check(self._L.locust_multi_wordcount_device(self._h, ptrs, sizes), "locust_multi_wordcount_device")
        return self.context(0)

    def stats(self):
        ms, b = ctypes.c_float(), ctypes.c_uint64()
        self._L.locust_multi_stats(self._h, ctypes.byref(ms), ctypes.byref(b))
        return {"job_ms": ms.value, "record_bytes_exchanged": b.value}

    def close(self):
        if getattr(self, "_h", None):
            self._L.locust_multi_destroy(self._h)
            self._h = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass
