"""Keys longer than 16 bytes: combined in MAP (hash of all bytes, matched by full byte compare) and put
in order by a parallel sort of the records that share a 16-byte prefix.  The reference treats long
keys like any other (KeyValue.h:6-33, main.cu:146, :415); the oracle is the checker."""
import random
import time

import numpy as np
import pytest

pytestmark = pytest.mark.gpu

torch = pytest.importorskip("torch")


@pytest.fixture(scope="module")
def lb():
    if not torch.cuda.is_available():
        pytest.skip("no CUDA device")
    import locust_b200
    locust_b200.load()
    return locust_b200


def arrays_equal_oracle(wc, data_np):
    from oracle import parity
    ref = parity.oracle_arrays(data_np)
    got = wc.result_arrays()
    why = parity.same_result(got, ref)
    assert why is None, why
    assert wc.stats()["tokens"] == ref[3]
    vals = got[3]
    if len(got[2]):
        assert np.array_equal(vals, np.concatenate([[0], np.cumsum(got[2])[:-1]]).astype(np.uint64))
    return ref


def test_two_long_keys_a_million_times_each(lb):
    """two 24-byte keys that share their first 16 bytes, interleaved, 10^6 occurrences each: the long-key
    list ends up with (about) two entries, and PROCESS + REDUCE have nothing to chew on"""
    a, b = b"https_//example_org/aaaa", b"https_//example_org/aaab"      # ':' and '.' would split them
    assert len(a) == 24 and a[:16] == b[:16]
    unit = a + b" " + b + b"\n" + b"the " + b + b" " + a + b" of\n"
    data = np.frombuffer(unit * 500_000, dtype=np.uint8)
    dev = torch.from_numpy(data.copy()).cuda()
    wc = lb.WordCount(device=0)
    try:
        for _ in range(2):
            wc.run(dev)
        ms = wc.stage_ms()
        keys, counts, _ = wc.result()
        assert dict(zip(keys, counts)) == {a: 1_000_000, b: 1_000_000, b"the": 500_000, b"of": 500_000}
        assert keys == sorted(keys)
        assert ms[1] + ms[2] < 5.0, "PROCESS + REDUCE took %.2f ms" % (ms[1] + ms[2])
        arrays_equal_oracle(wc, data)
    finally:
        wc.close()


def test_many_distinct_long_keys_with_one_prefix(lb):
    """10^5 distinct 40-byte keys with a common 16-byte prefix (one tied group of 10^5 records), some of
    them prefixes of others, plus the 16-byte prefix itself as a key"""
    rng = random.Random(7)
    prefix = b"user_id=00000000"                                   # 16 bytes; '=' and '_' are not delimiters
    assert len(prefix) == 16
    words = set()
    while len(words) < 100_000:
        words.add(prefix + bytes(rng.choice(b"abcdefghijklmnopqrstuvwxyz0123456789") for _ in range(24)))
    words = list(words)
    extra = [prefix, words[0][:17], words[1][:30], words[2] + b"x", words[3] + b"yz" * 40]
    toks = words + extra + [rng.choice(words) for _ in range(50_000)] + extra
    rng.shuffle(toks)
    data = np.frombuffer(b" ".join(toks) + b"\n", dtype=np.uint8)
    wc = lb.WordCount(device=0)
    try:
        t0 = time.perf_counter()
        wc.run(torch.from_numpy(data.copy()).cuda())
        dt = time.perf_counter() - t0
        ref = arrays_equal_oracle(wc, data)
        assert len(ref[2]) == len(set(toks))
        assert dt < 20.0, "job took %.1f s" % dt
    finally:
        wc.close()


@pytest.mark.parametrize("dist,bound", [("uniform", 3.0), ("zipf", 6.0)])
def test_every_token_long(lb, dist, bound):
    """64 MiB in which every token has 17..64 bytes (20 000 distinct words): parity, and MAP (position list +
    long_tokens_kernel) within `bound` x of its rate on BASELINE config 2 text of the same size.  With words
    drawn uniformly the bound is 3; with Zipf(1.3) draws a quarter of all tokens are one word and its counter
    in L2 is the limit (equal keys of a warp are added up first, which is what keeps it within 6)."""
    from locust_b200.textgen import TextGen
    rng = np.random.default_rng(11)
    nwords = 20_000
    lens = rng.integers(17, 65, size=nwords)
    vocab = [bytes(rng.integers(97, 123, size=int(n), dtype=np.uint8)) for n in lens]
    picks = (rng.zipf(1.3, size=1_800_000) % nwords) if dist == "zipf" else rng.integers(0, nwords, size=1_800_000)
    n = 64 << 20
    text = b" ".join(vocab[int(i)] for i in picks)[:n]
    data = np.frombuffer(text[: text.rfind(b" ")] + b"\n", dtype=np.uint8)
    dev = torch.from_numpy(data.copy()).cuda()
    c2 = torch.empty(n, dtype=torch.uint8, device="cuda")
    TextGen.config("C2", total_bytes=n, device=0).fill_device(c2)
    torch.cuda.synchronize()
    wc = lb.WordCount(device=0)
    try:
        for _ in range(3):
            wc.run(c2)
        c2_rate = n / wc.stage_ms()[0]
        for _ in range(3):
            wc.run(dev)
        rate = len(data) / wc.stage_ms()[0]
        arrays_equal_oracle(wc, data)
        assert rate * bound >= c2_rate, "MAP on long tokens %.1f MB/ms, on config 2 text %.1f MB/ms" % (rate / 1e6, c2_rate / 1e6)
    finally:
        wc.close()


def test_long_keys_grow_the_list(lb):
    """more distinct long keys than long_capacity: the list grows and the job is redone"""
    rng = random.Random(3)
    words = [b"k%06d_" % i + bytes(rng.choice(b"abcdef") for _ in range(20)) for i in range(5000)]
    data = b" ".join(words + words[:100]) + b"\n"
    wc = lb.WordCount(device=0, table_log2=12, long_capacity=64)
    try:
        wc.run(data)
        keys, counts, _ = wc.result()
        from oracle import pyoracle
        ref = pyoracle.wordcount(data)
        assert keys == ref.keys and counts == ref.counts
    finally:
        wc.close()
