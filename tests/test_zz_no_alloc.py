"""locust_count_static -- the entry point for callers that cannot allocate device memory
(a worker whose GPU heap already belongs to another process or to a framework's pooling
allocator, or a path that cannot take the synchronisation an allocation implies).  Its
combiner scratch is a fixed array inside liblocust.so instead of cudaMalloc'd buffers.

Checked here through the C ABI against the ordinary entry points (locust_create / map /
process / reduce) on the same bytes and against the CPU oracle: same words, same counts, bit
for bit.  Two calls in a row on different inputs show that the shared scratch is cleared
between calls.  The file
name makes pytest collect it last, so a run that stops at the first failure has already
run every other test; it also passes alone in a fresh process.
"""
import ctypes

import pytest

from conftest import golden_input

pytestmark = pytest.mark.gpu

torch = pytest.importorskip("torch")


@pytest.fixture(scope="module")
def lb():
    if not torch.cuda.is_available():
        pytest.skip("no CUDA device")
    import locust_b200
    locust_b200.load()
    return locust_b200


def static_count(lb, data, max_records=None, shift=0):
    """dict word -> count from locust_count_static on `data` (bytes); shift: the device copy starts
    that many bytes past an aligned address"""
    max_records = lb.STATIC_MAX_KEYS if max_records is None else max_records
    d = torch.zeros(len(data) + shift + 1, dtype=torch.uint8, device="cuda")
    d[shift:shift + len(data)] = torch.frombuffer(bytearray(data) if data else bytearray(1), dtype=torch.uint8)[:len(data)].cuda()
    d = d[shift:shift + len(data)]
    out = torch.zeros(max(max_records, 1) * 4, dtype=torch.int64, device="cuda")
    torch.cuda.synchronize()
    n = lb.count_static(d.data_ptr(), len(data), out.data_ptr(), max_records,
                        stream=torch.cuda.current_stream().cuda_stream)
    rec = out.cpu().numpy().view("uint64").reshape(-1, 4)[:n]
    res = {}
    for k0, k1, cnt, pad in rec.tolist():
        key = (int(k0).to_bytes(8, "little") + int(k1).to_bytes(8, "little")).rstrip(b"\x00")
        assert pad == 0 and key not in res
        res[key] = cnt
    return res


def regular_count(lb, data):
    wc = lb.WordCount(device=0, table_log2=16)
    try:
        wc.run(data)
        keys, counts, _ = wc.result()
        return dict(zip(keys, counts))
    finally:
        wc.close()


def oracle_count(data):
    from oracle import pyoracle
    return pyoracle.wordcount(data, formats=False).as_dict()


def test_two_calls_in_a_row_match_the_regular_path(lb):
    a = golden_input("hamlet")                    # 5 608 distinct words, longest 14 bytes
    b = golden_input("rand_heavy_s4")             # different words: nothing of call 1 may survive
    for data in (a, b):
        got = static_count(lb, data)
        assert got == regular_count(lb, data)
        assert got == oracle_count(data)
    # and once more the first input, after the second
    assert static_count(lb, a) == regular_count(lb, a)
    assert static_count(lb, b"") == {}


def test_unaligned_input_and_bytes_above_ascii(lb):
    """the device pointer need not be aligned, and bytes >= 0x80 and '\r' are word content"""
    text = golden_input("hamlet")[:20000] + " \xe9t\xe9 na\xefve r\xe9sum\xe9 \xe9t\xe9\r\nline\r two\r\n".encode("latin-1") * 7
    want = oracle_count(text)
    for shift in (0, 1, 7, 15):
        assert static_count(lb, text, shift=shift) == want


def test_limits_are_refused_with_a_reason(lb):
    # a word longer than 16 bytes does not fit the fixed table
    with pytest.raises(lb.LocustError, match="longer than 16 bytes") as ei:
        static_count(lb, b"short " + b"x" * 17 + b" short")
    assert ei.value.code == 3
    # more distinct words than LOCUST_STATIC_MAX_KEYS
    many = b" ".join(b"w%d" % i for i in range(lb.STATIC_MAX_KEYS + 50))
    with pytest.raises(lb.LocustError, match="distinct words") as ei:
        static_count(lb, many)
    assert ei.value.code == 3
    # an output buffer larger than the table can ever fill is fine
    assert static_count(lb, b"a b", max_records=lb.STATIC_MAX_KEYS + 1) == {b"a": 1, b"b": 1}
    # an input size beyond the maximum is refused before anything runs
    err = ctypes.create_string_buffer(200)
    n = ctypes.c_uint64()
    rc = lb.load().locust_count_static(1, lb.STATIC_MAX_BYTES + 1, 1, 1, ctypes.byref(n), None, err, 200)
    assert rc == 1 and b"LOCUST_STATIC_MAX_BYTES" in err.value
    # an output buffer that is too small reports how many records are needed
    with pytest.raises(lb.LocustError, match="too small") as ei:
        static_count(lb, b"a b c d e", max_records=3)
    assert ei.value.code == 3 and ei.value.needed == 5
    # the scratch is usable again after every refusal
    assert static_count(lb, b"b a b") == {b"a": 1, b"b": 2}
