"""Parity of the CUDA path (through the C ABI of liblocust.so) with the CPU oracle and with
the golden output of the reference's own CPU branch.  Bit-exact: same key set, same counts,
same order.  Needs a B200; run with `pytest -m gpu`."""
import os
import random
import re
import subprocess

import pytest

from conftest import ROOT, golden_cases, golden_input, golden_ref_cpu

pytestmark = pytest.mark.gpu

torch = pytest.importorskip("torch")

DELIMS = b" ,.-;:'()\"\t\n"
CASES = golden_cases()


@pytest.fixture(scope="module")
def lb():
    if not torch.cuda.is_available():
        pytest.skip("no CUDA device")
    import locust_b200
    locust_b200.load()          # raises if the extension is missing: no silent fallback
    return locust_b200


@pytest.fixture(scope="module")
def oracle():
    from oracle import pyoracle
    return pyoracle


def gpu_count(lb, data, where="host", **kw):
    wc = lb.WordCount(device=0, **kw)
    try:
        if where == "device":
            t = torch.frombuffer(bytearray(data) if len(data) else bytearray(1), dtype=torch.uint8)[:len(data)].cuda()
            wc.run(t)
        else:
            wc.run(data)
        keys, counts, vals = wc.result()
        return keys, counts, vals, wc.format(cpu_format=True), wc.stats()
    finally:
        wc.close()


def check_against_oracle(lb, oracle, data, where="host", **kw):
    keys, counts, vals, text, st = gpu_count(lb, data, where, **kw)
    ref = oracle.wordcount(data)
    assert keys == ref.keys
    assert counts == ref.counts
    assert text == ref.cpu_lines
    acc = 0
    for c, v in zip(counts, vals):           # `val` = exclusive prefix sum of counts (main.cu:132)
        assert v == acc
        acc += c
    assert st["tokens"] == ref.tokens == acc and st["unique_keys"] == len(ref.keys)
    return ref


@pytest.mark.parametrize("case", CASES, ids=[c[0] for c in CASES])
@pytest.mark.parametrize("where", ["host", "device"])
def test_golden_vectors(lb, oracle, case, where):
    name, _, _, _, _, warn = case
    data = golden_input(name)
    keys, counts, vals, text, st = gpu_count(lb, data, where, table_log2=14)
    if not warn:
        # byte-identical to what the reference's CPU branch printed
        assert text == golden_ref_cpu(name)
    else:
        # the reference drops tokens 21+ of a line (main.cu:300-303); that cap is lifted here
        assert text == oracle.wordcount(data).cpu_lines
        assert text != golden_ref_cpu(name)


def test_hamlet_known_answers(lb):
    keys, counts, vals, text, st = gpu_count(lb, golden_input("hamlet"))
    d = dict(zip(keys, counts))
    assert st["tokens"] == 32940 and len(keys) == 5608
    assert d[b"the"] == 930 and d[b"and"] == 692 and d[b"Hamlet"] == 96 and d[b"zone"] == 1
    assert keys[:4] == [b"&c", b"1", b"1604", b"?"] and keys[-1] == b"zone"


def test_gpu_branch_line_format(lb):
    wc = lb.WordCount(device=0, table_log2=14)
    wc.run(golden_input("hamlet"))
    keys, counts, vals = wc.result()
    lines = wc.format(cpu_format=False).split(b"\n")
    assert lines[-1] == b"" and len(lines) == len(keys) + 1
    for k, c, v, line in zip(keys, counts, vals, lines):
        assert line == b"print key: %s \t val: %d \t count: %d" % (k, v, c)     # main.cu:132
    wc.close()


FLAG_SETS = [0, 1, 2, 3, 4, 6, 7, 8, 9]


@pytest.mark.parametrize("flags", FLAG_SETS)
@pytest.mark.parametrize("name", ["hamlet", "edge_basic", "edge_long_tokens", "rand_heavy_s4", "rand_print_s3"])
def test_every_code_path_agrees(lb, oracle, name, flags):
    # 1: byte-serial tokeniser, 2: no shared-memory combiner, 4: 8-bit hash (forced collisions),
    # 8: alternative path (every token a record, radix sort + run-length reduce, no combiner)
    check_against_oracle(lb, oracle, golden_input(name), flags=flags, table_log2=14)


def random_bytes(rng, n, alphabet, p_delim):
    out = bytearray(n)
    for i in range(n):
        out[i] = rng.choice(DELIMS) if rng.random() < p_delim else rng.choice(alphabet)
    return bytes(out)


@pytest.mark.parametrize("n", [0, 1, 2, 15, 16, 17, 31, 32, 33, 511, 512, 513, 1023, 1025, 4096, 65537, 300001])
def test_sizes_and_ragged_ends(lb, oracle, n):
    rng = random.Random(n)
    data = random_bytes(rng, n, b"abcdeXYZ019_!", 0.25)
    for where in ("host", "device"):
        check_against_oracle(lb, oracle, data, where, table_log2=12)
    if n:
        # no trailing delimiter and no leading delimiter
        data2 = b"q" + data[1:-1] + b"q" if n > 1 else b"q"
        check_against_oracle(lb, oracle, data2, "device", table_log2=12)


@pytest.mark.parametrize("p_delim", [0.0, 0.02, 0.5, 0.95, 1.0])
def test_delimiter_densities(lb, oracle, p_delim):
    rng = random.Random(int(p_delim * 100))
    data = random_bytes(rng, 50000, b"ab", p_delim)
    check_against_oracle(lb, oracle, data, "device", table_log2=12)


def test_every_byte_value_except_nul(lb, oracle):
    rng = random.Random(7)
    alphabet = bytes(range(1, 256))
    data = random_bytes(rng, 200000, alphabet, 0.1)
    check_against_oracle(lb, oracle, data, "device", table_log2=16)
    check_against_oracle(lb, oracle, data, "host", table_log2=16, flags=1)


def test_unaligned_device_pointers(lb, oracle):
    rng = random.Random(11)
    data = random_bytes(rng, 9000, b"abcdefgh", 0.2)
    base = torch.frombuffer(bytearray(b" " * 64 + data + b" " * 64), dtype=torch.uint8).cuda()
    wc = lb.WordCount(device=0, table_log2=12)
    for off in range(0, 34):
        for n in (len(data) - off, 4097, 777):
            view = base[64 + off:64 + off + n]
            wc.run(view)
            keys, counts, _ = wc.result()
            ref = oracle.wordcount(data[off:off + n])
            assert keys == ref.keys and counts == ref.counts, (off, n)
    wc.close()


def test_long_tokens(lb, oracle):
    rng = random.Random(5)
    words = []
    stem = b"x" * 16
    for i in range(300):
        L = rng.choice([15, 16, 17, 18, 31, 32, 33, 100, 511, 512, 513, 600, 5000])
        w = bytes(rng.choice(b"ab") for _ in range(L))
        words += [w] * rng.randint(1, 4)
        # many keys that agree on their first 16 bytes and differ later / in length
        words.append(stem + bytes(rng.choice(b"abc") for _ in range(rng.randint(0, 5))))
    rng.shuffle(words)
    data = b" ".join(words) + b"\n" + b"z" * 70000      # one token longer than many rows, no final delimiter
    ref = check_against_oracle(lb, oracle, data, "device", table_log2=12)
    assert max(len(k) for k in ref.keys) == 70000
    check_against_oracle(lb, oracle, data, "host", table_log2=12, long_capacity=4)      # forces the long buffer to grow


def test_sort_path_matches_and_grows(lb, oracle):
    # the alternative path needs one record per token: 16 slots + 4 long entries must grow a lot
    data = golden_input("rand_heavy_s4") + golden_input("edge_long_tokens")
    check_against_oracle(lb, oracle, data, "device", table_log2=4, long_capacity=4, flags=8)


def test_table_growth(lb, oracle):
    rng = random.Random(3)
    words = [bytes(rng.choice(b"abcdefghijklmnopqrstuvwxyz") for _ in range(rng.randint(1, 12))) for _ in range(20000)]
    data = b" ".join(rng.choice(words) for _ in range(100000))
    check_against_oracle(lb, oracle, data, "device", table_log2=4)       # 16 slots: must grow several times
    check_against_oracle(lb, oracle, data, "host", table_log2=4, flags=2)


def test_map_in_several_calls_is_linear(lb, oracle):
    a, b, c = golden_input("hamlet"), golden_input("rand_heavy_s4"), golden_input("edge_basic") + b"\n"
    wc = lb.WordCount(device=0, table_log2=14)
    wc.reset()
    ta = torch.frombuffer(bytearray(a), dtype=torch.uint8).cuda()
    wc.map(ta)
    wc.map(b)                       # host bytes
    wc.map(torch.frombuffer(bytearray(c), dtype=torch.uint8).cuda())
    wc.process()
    wc.reduce()
    keys, counts, _ = wc.result()
    ref = oracle.wordcount(a + b + c)
    assert keys == ref.keys and counts == ref.counts
    # counts are additive over inputs
    da, db, dc = (oracle.wordcount(x).as_dict() for x in (a, b, c))
    for k, v in zip(keys, counts):
        assert v == da.get(k, 0) + db.get(k, 0) + dc.get(k, 0)
    # a context can be reused
    wc.run(a)
    assert wc.result()[0] == oracle.wordcount(a).keys
    wc.close()


def test_stage_order_is_enforced(lb):
    wc = lb.WordCount(device=0, table_log2=12)
    wc.reset()
    with pytest.raises(lb.LocustError):
        wc.reduce()
    wc.map(b"a b a")
    wc.process()
    with pytest.raises(lb.LocustError):
        wc.map(b"c")
    wc.reduce()
    assert wc.result()[:2] == ([b"a", b"b"], [2, 1])
    wc.close()


def test_empty_job(lb):
    wc = lb.WordCount(device=0, table_log2=12)
    wc.run(b"")
    assert wc.result() == ([], [], [])
    assert wc.format() == b""
    wc.run(b" \n\t ...")
    assert wc.result() == ([], [], [])
    wc.close()


def test_textgen_device_matches_host(lb):
    from locust_b200.textgen import TextGen, BLOCK
    for name in ("C2", "C4", "C5"):
        n = 256 * BLOCK
        g = TextGen.config(name, total_bytes=4 * n, device=0, vocab=20000 if name != "C4" else None)
        d = torch.empty(n, dtype=torch.uint8, device="cuda")
        g.fill_device(d, byte_offset=n)
        torch.cuda.synchronize()
        assert bytes(d.cpu().numpy()) == bytes(g.host(n, byte_offset=n)), name


@pytest.mark.parametrize("name,mib", [("C2", 64), ("C4", 32), ("C5", 64)])
def test_synthetic_configs_against_oracle(lb, oracle, name, mib):
    from locust_b200.textgen import TextGen
    n = mib << 20
    g = TextGen.config(name, total_bytes=n, device=0, vocab=200000 if name == "C5" else None)
    d = torch.empty(n, dtype=torch.uint8, device="cuda")
    g.fill_device(d)
    torch.cuda.synchronize()
    host = d.cpu().numpy()
    ref = oracle.wordcount(host, mode="fast", formats=False)
    wc = lb.WordCount(device=0, table_log2=18)
    wc.run(d)
    keys, counts, _ = wc.result()
    assert keys == ref.keys and counts == ref.counts
    # the same bytes handed over as a host buffer (chunked copy overlapped with map)
    pinned = torch.from_numpy(host).pin_memory()
    wc.run(pinned)
    k2, c2, _ = wc.result()
    assert k2 == keys and c2 == counts
    wc.close()


def test_full_size_c2_properties(lb, oracle):
    """BASELINE.json configs[1] at full size (1 GiB): checked against the oracle's fast
    path on the same bytes and through size-independent properties."""
    import numpy as np
    from locust_b200.textgen import TextGen
    n = 1 << 30
    g = TextGen.config("C2", device=0)
    d = torch.empty(n, dtype=torch.uint8, device="cuda")
    g.fill_device(d)
    torch.cuda.synchronize()
    wc = lb.WordCount(device=0, table_log2=17)
    wc.run(d)
    blob, offs, counts, vals = wc.result_arrays()
    st = wc.stats()
    # sortedness and uniqueness of the keys
    keys = [bytes(blob[int(offs[i]):int(offs[i + 1])]) for i in range(len(counts))]
    assert all(keys[i] < keys[i + 1] for i in range(len(keys) - 1))
    # checksum of checksums: counts add up to the token count, vals are their prefix sums
    assert int(counts.sum()) == st["tokens"]
    assert (vals == np.concatenate([[0], np.cumsum(counts)[:-1]]).astype(np.uint64)).all()
    # linearity: the two halves counted separately add up to the whole
    half = n // 2
    wc2 = lb.WordCount(device=0, table_log2=17)
    wc2.run(d[:half]); ka, ca, _ = wc2.result()
    wc2.run(d[half:]); kb, cb, _ = wc2.result()
    merged = dict(zip(ka, ca))
    for k, c in zip(kb, cb):
        merged[k] = merged.get(k, 0) + c
    assert merged == dict(zip(keys, counts.tolist()))
    # idempotence: a second run on the same context gives the same records
    wc.run(d)
    b2, o2, c2, _ = wc.result_arrays()
    assert (b2 == blob).all() and (o2 == offs).all() and (c2 == counts).all()
    # and the oracle on the same bytes
    ref = oracle.wordcount(d.cpu().numpy(), mode="fast", formats=False)
    assert ref.keys == keys and ref.counts == counts.tolist() and ref.tokens == st["tokens"]
    wc.close(); wc2.close()


def test_cli_matches_reference_stdout(lb, tmp_path):
    exe = os.path.join(ROOT, "locust_b200", "MapReduce")
    path = os.path.join(ROOT, "tests", "golden", "hamlet.txt")
    env = dict(os.environ, LOCUST_FORMAT="cpu", LOCUST_TABLE_LOG2="14")
    out = subprocess.run([exe, path], capture_output=True, env=env, check=True).stdout
    lines = out.split(b"\n")
    assert lines[0] == b"Running" and lines[1] == b"Length: 4463"
    assert re.fullmatch(rb"GPU mapping \d+ nanoseconds ", lines[2])
    assert re.fullmatch(rb"GPU stream compaction and sorting \d+ nanoseconds ", lines[3])
    assert re.fullmatch(rb"GPU reduce \d+ nanoseconds ", lines[4])
    assert out.endswith(b"\n\nDone\n")
    keylines = b"".join(l + b"\n" for l in lines if l.startswith(b"print key: "))
    assert keylines == golden_ref_cpu("hamlet")
    # usage text and exit code of main.cu:362-366
    r = subprocess.run([exe], capture_output=True)
    assert r.returncode == 255 and r.stdout == (b"Running\nMissing or invalid arguments.\n"
                                                b"mapreduce <filename> [line_start] [line_end] [node_num] [stage]\n")
    # [line_start, line_end) window
    r = subprocess.run([exe, path, "10", "20"], capture_output=True, env=env, check=True).stdout
    with open(path, "rb") as f:
        window = b"".join(l + b"\n" for l in f.read().split(b"\n")[10:20])
    from oracle import pyoracle
    want = pyoracle.wordcount(window).cpu_lines
    got = b"".join(l + b"\n" for l in r.split(b"\n") if l.startswith(b"print key: "))
    assert got == want and b"Using custom start and end locations: (10, 20)\nLength: 10\n" in r


def test_intermediate_format_round_trip(lb, oracle):
    """writeKeyIntValues / loadIntermediateFile (main.cu:116-124, :66-103): partials written as
    "%s \\t%d\\n" lines and read back reduce to the same result; a file in the reference's own
    stage-1 form (one line per token, value 1) does too."""
    a, b = golden_input("hamlet"), golden_input("edge_long_tokens") + golden_input("rand_heavy_s4")
    wc = lb.WordCount(device=0, table_log2=14)
    wc.run(a)
    inter_a = wc.format(intermediate=True)
    keys, counts, _ = wc.result()
    assert inter_a == b"".join(b"%s \t%d\n" % (k, c) for k, c in zip(keys, counts))
    wc.run(b)
    inter_b = wc.format(intermediate=True)
    # two nodes' files merged by a reduce stage
    wc.reset()
    wc.map_intermediate(inter_a)
    wc.map_intermediate(inter_b)
    wc.process()
    wc.reduce()
    ref = oracle.wordcount(a + b"\n" + b)
    assert wc.result()[:2] == (ref.keys, ref.counts)
    # the reference's stage-1 file: every token on its own line with value 1, sorted
    toks = sorted(t for t in re.split(b"[" + re.escape(DELIMS) + b"]+", b) if t)
    ref_style = b"".join(t + b" \t1\n" for t in toks)
    wc.reset()
    wc.map_intermediate(ref_style)
    wc.map(a)                                   # and it mixes with ordinary map calls
    wc.process()
    wc.reduce()
    ref2 = oracle.wordcount(b + b"\n" + a)
    assert wc.result()[:2] == (ref2.keys, ref2.counts)
    wc.close()


def test_cli_multi_mode_stages(lb, tmp_path):
    exe = os.path.join(ROOT, "locust_b200", "MapReduce")
    path = os.path.join(ROOT, "tests", "golden", "hamlet.txt")
    inter = str(tmp_path / "out.txt")
    env = dict(os.environ, LOCUST_FORMAT="cpu", LOCUST_TABLE_LOG2="14", LOCUST_INTERMEDIATE=inter)
    s1 = subprocess.run([exe, path, "0", "100000", "0", "1"], capture_output=True, env=env, check=True).stdout
    assert s1.endswith(b"MODE_MULTI: Finished map\n") and b"Done" not in s1 and b"print key" not in s1
    with open(inter, "rb") as f:
        first = f.readline()
    assert first == b"&c \t1\n"
    s2 = subprocess.run([exe, path, "0", "100000", "0", "2"], capture_output=True, env=env, check=True).stdout
    assert s2.startswith(b"Running\nUsing custom start and end locations: (0, 100000)\nGPU reduce ") and s2.endswith(b"\n\nDone\n")
    keylines = b"".join(l + b"\n" for l in s2.split(b"\n") if l.startswith(b"print key: "))
    assert keylines == golden_ref_cpu("hamlet")


def _packed_keys(blob, offs):
    """(n, 2) big-endian uint64 view of keys of <= 16 bytes, zero padded: integer order = key order"""
    import numpy as np
    offs = offs.astype(np.int64)
    lens = np.diff(offs)
    assert lens.min() >= 1 and lens.max() <= 16
    keys = np.zeros((len(lens), 16), dtype=np.uint8)
    idx = np.repeat(np.arange(len(lens)), lens)
    within = np.arange(len(blob)) - np.repeat(offs[:-1], lens)
    keys[idx, within] = blob
    return keys.view(">u8").reshape(-1, 2).astype(np.uint64)


@pytest.mark.parametrize("name,mib,log2", [("C3", 256, 20), ("C4", 128, 20), ("C5", 256, 20)])
def test_larger_configs_by_properties(lb, name, mib, log2):
    """Size-independent checks on the other BASELINE configurations at sizes the oracle would
    need minutes for: strictly ascending keys, counts that add up to the token count, vals as
    prefix sums, the two halves of the input adding up to the whole, and the combiner table
    growing by itself (C4 starts 16x too small)."""
    import numpy as np
    from locust_b200.textgen import TextGen
    n = mib << 20
    g = TextGen.config(name, total_bytes=n, device=0)
    d = torch.empty(n, dtype=torch.uint8, device="cuda")
    g.fill_device(d)
    torch.cuda.synchronize()
    wc = lb.WordCount(device=0, table_log2=log2)

    def run(view):
        wc.run(view)
        blob, offs, counts, vals = wc.result_arrays()
        st = wc.stats()
        assert int(counts.sum()) == st["tokens"] and len(counts) == st["unique_keys"]
        assert (vals == np.concatenate([[0], np.cumsum(counts)[:-1]]).astype(np.uint64)).all()
        k = _packed_keys(blob, offs)
        a, b = k[:-1], k[1:]
        assert ((a[:, 0] < b[:, 0]) | ((a[:, 0] == b[:, 0]) & (a[:, 1] < b[:, 1]))).all()      # sorted, unique
        return k, counts

    kw, cw = run(d)
    half = n // 2                                  # a generator block boundary = a token boundary
    k1, c1 = run(d[:half])
    k2, c2 = run(d[half:])
    k = np.concatenate([k1, k2])
    c = np.concatenate([c1, c2])
    order = np.lexsort((k[:, 1], k[:, 0]))
    k, c = k[order], c[order]
    head = np.ones(len(k), dtype=bool)
    head[1:] = (k[1:] != k[:-1]).any(axis=1)
    sums = np.add.reduceat(c, np.flatnonzero(head))
    assert (k[head] == kw).all() and (sums == cw).all()
    wc.close()


def test_streaming_map_kernel_variants(lb):
    """the bulk-async / inserter-warp MAP kernel (LOCUST_MAP_VARIANT >= 0, map_stream_kernel.cu) is not the
    default, but it stays correct: a ladder of inputs through it in child processes, against the oracle"""
    import sys
    out = subprocess.run([sys.executable, os.path.join(ROOT, "tools", "dbg_map.py"), "0,3", "60", "small"],
                         capture_output=True, text=True, timeout=900).stdout
    lines = [l for l in out.splitlines() if l.startswith("variant")]
    assert len(lines) >= 16, out
    assert all(": ok " in l for l in lines), out


def test_cli_reads_the_file_in_chunks(lb, tmp_path):
    """single mode reads the file through a ring of pinned chunks (disk read of chunk k overlapping copy + map of
    chunk k - 1); the stdout must not depend on the chunk size, nor differ from the whole-file path"""
    from locust_b200.textgen import TextGen
    exe = os.path.join(ROOT, "locust_b200", "MapReduce")
    data = bytes(TextGen.config("C2", total_bytes=5 << 20).host(5 << 20))[:-3] + b"tail"       # no newline at the end
    path = tmp_path / "five_mib.txt"
    path.write_bytes(data)
    env = dict(os.environ, LOCUST_FORMAT="cpu")
    keep = lambda out: [l for l in out.split(b"\n") if l.startswith(b"print key: ") or l.startswith(b"Length")]
    whole = keep(subprocess.run([exe, str(path)], env=dict(env, LOCUST_WHOLE_FILE="1"), capture_output=True, check=True).stdout)
    for mib in ("1", "2", "64"):
        got = keep(subprocess.run([exe, str(path)], env=dict(env, LOCUST_CHUNK_MIB=mib), capture_output=True, check=True).stdout)
        assert got == whole, "chunk size %s MiB" % mib
    assert len(whole) > 1000 and whole[0].startswith(b"Length: ")
