"""TEST INFRASTRUCTURE -- ctypes access to oracle/liboracle.so (the CPU restatement).

Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference
legs import this module.  The product package (locust_b200) never does.
"""
import ctypes
import os
import subprocess

HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(HERE, "liboracle.so")
REF_BIN = os.path.join(HERE, "_ref", "MapReduce_cpu")


class _WcResult(ctypes.Structure):
    _fields_ = [("n", ctypes.c_uint64), ("tokens", ctypes.c_uint64),
                ("key_blob", ctypes.POINTER(ctypes.c_uint8)),
                ("key_offsets", ctypes.POINTER(ctypes.c_uint64)),
                ("counts", ctypes.POINTER(ctypes.c_uint64))]


def build():
    """Compile the C restatement (and, when the reference tree is mounted, oracle/_ref)."""
    subprocess.run(["make", "-s", "-C", HERE, "liboracle.so", "wc_oracle"], check=True)
    subprocess.run(["bash", os.path.join(HERE, "build_ref.sh")], check=True,
                   stdout=subprocess.DEVNULL, stderr=subprocess.DEVNULL)


_lib = None


def lib():
    global _lib
    if _lib is None:
        if not os.path.exists(LIB_PATH):
            build()
        L = ctypes.CDLL(LIB_PATH)
        L.wc_oracle_sort.argtypes = [ctypes.c_void_p, ctypes.c_size_t, ctypes.c_int, ctypes.POINTER(_WcResult)]
        L.wc_oracle_sort.restype = ctypes.c_int
        L.wc_oracle_fast.argtypes = [ctypes.c_void_p, ctypes.c_size_t, ctypes.POINTER(_WcResult)]
        L.wc_oracle_fast.restype = ctypes.c_int
        L.wc_result_free.argtypes = [ctypes.POINTER(_WcResult)]
        L.wc_format_cpu.argtypes = [ctypes.POINTER(_WcResult), ctypes.c_void_p, ctypes.c_size_t]
        L.wc_format_cpu.restype = ctypes.c_size_t
        L.wc_format_tsv.argtypes = [ctypes.POINTER(_WcResult), ctypes.c_void_p, ctypes.c_size_t]
        L.wc_format_tsv.restype = ctypes.c_size_t
        L.pr_oracle.argtypes = [ctypes.c_void_p, ctypes.c_size_t, ctypes.c_uint32, ctypes.c_uint64, ctypes.POINTER(_WcResult)]
        L.pr_oracle.restype = ctypes.c_int
        _lib = L
    return _lib


class OracleResult:
    """keys (list of bytes, ascending byte order) and counts (list of int)."""

    def __init__(self, keys, counts, tokens, cpu_lines, tsv):
        self.keys, self.counts, self.tokens = keys, counts, tokens
        self.cpu_lines, self.tsv = cpu_lines, tsv

    def as_dict(self):
        return dict(zip(self.keys, self.counts))


def _buf_ptr(data):
    """(pointer, nbytes, keepalive) for bytes / bytearray / numpy uint8 array."""
    try:
        import numpy as np
        if isinstance(data, np.ndarray):
            a = np.ascontiguousarray(data, dtype=np.uint8)
            return a.ctypes.data, a.size, a
    except ImportError:
        pass
    b = bytes(data)
    return ctypes.cast(ctypes.c_char_p(b), ctypes.c_void_p).value, len(b), b


def wordcount(data, mode="sort", emits_per_line=0, formats=True):
    """Run the CPU restatement on `data`; mode 'sort' follows the reference's stages,
    'fast' is the hash-table checker for large inputs."""
    L = lib()
    ptr, n, keep = _buf_ptr(data)
    r = _WcResult()
    if mode == "sort":
        rc = L.wc_oracle_sort(ptr, n, emits_per_line, ctypes.byref(r))
    else:
        rc = L.wc_oracle_fast(ptr, n, ctypes.byref(r))
    if rc != 0:
        raise RuntimeError("oracle failed: %d" % rc)
    try:
        nu = r.n
        offs = [r.key_offsets[i] for i in range(nu + 1)] if nu < 4_000_000 else None
        if offs is None:
            import numpy as np
            offs = np.ctypeslib.as_array(r.key_offsets, shape=(nu + 1,)).copy()
            counts = np.ctypeslib.as_array(r.counts, shape=(nu,)).copy()
            blob = bytes(np.ctypeslib.as_array(r.key_blob, shape=(int(offs[nu]),)))
            keys = [blob[int(offs[i]):int(offs[i + 1])] for i in range(nu)]
            counts = counts.tolist()
        else:
            blob = ctypes.string_at(r.key_blob, offs[nu]) if nu else b""
            keys = [blob[offs[i]:offs[i + 1]] for i in range(nu)]
            counts = [r.counts[i] for i in range(nu)]
        cpu_lines = tsv = None
        if formats:
            need = L.wc_format_cpu(ctypes.byref(r), None, 0)
            b = ctypes.create_string_buffer(need if need else 1)
            L.wc_format_cpu(ctypes.byref(r), b, need)
            cpu_lines = b.raw[:need]
            need = L.wc_format_tsv(ctypes.byref(r), None, 0)
            b = ctypes.create_string_buffer(need if need else 1)
            L.wc_format_tsv(ctypes.byref(r), b, need)
            tsv = b.raw[:need]
        return OracleResult(keys, counts, r.tokens, cpu_lines, tsv)
    finally:
        L.wc_result_free(ctypes.byref(r))


def pagerank(data, iterations, damping_q32):
    """Second job (oracle/pagerank_oracle.c): (nodes in byte order, ranks, number of edges)."""
    L = lib()
    ptr, n, keep = _buf_ptr(data)
    r = _WcResult()
    rc = L.pr_oracle(ptr, n, iterations, damping_q32, ctypes.byref(r))
    if rc != 0:
        raise RuntimeError("pagerank oracle failed: %d" % rc)
    try:
        nu = r.n
        offs = [r.key_offsets[i] for i in range(nu + 1)]
        blob = ctypes.string_at(r.key_blob, offs[nu]) if nu else b""
        return [blob[offs[i]:offs[i + 1]] for i in range(nu)], [r.counts[i] for i in range(nu)], r.tokens
    finally:
        L.wc_result_free(ctypes.byref(r))


def run_reference_cpu(path):
    """`print key:` lines printed by the reference's own CPU branch, or None if
    oracle/_ref was never built (it needs /root/reference at build time)."""
    if not os.path.exists(REF_BIN):
        return None
    out = subprocess.run([REF_BIN, path], check=True, capture_output=True).stdout
    return b"".join(l + b"\n" for l in out.split(b"\n") if l.startswith(b"print key: "))
