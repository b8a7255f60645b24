"""locust_multi_* through the C ABI (ctypes): one process, several ranks.  On the one-GPU test box the
ranks are contexts on cuda:0 (devices = [0, 0, 0]); with more GPUs visible every GPU is a rank.  The
oracle is the checker: same keys, counts, order and vals as the CPU restatement on the whole input."""
import os
import subprocess

import numpy as np
import pytest

from conftest import ROOT, golden_input, golden_ref_cpu

pytestmark = pytest.mark.gpu

torch = pytest.importorskip("torch")


@pytest.fixture(scope="module")
def lb():
    if not torch.cuda.is_available():
        pytest.skip("no CUDA device")
    import locust_b200
    locust_b200.load()
    return locust_b200


def check(wc, data):
    from oracle import parity
    ref = parity.oracle_arrays(np.frombuffer(bytes(data), dtype=np.uint8))
    got = wc.result_arrays()
    why = parity.same_result(got, ref)
    assert why is None, why
    if len(got[2]):
        assert np.array_equal(got[3], np.concatenate([[0], np.cumsum(got[2])[:-1]]).astype(np.uint64))


def device_lists():
    n = torch.cuda.device_count() if torch.cuda.is_available() else 1
    lists = [[0], [0, 0], [0, 0, 0], [0] * 8]
    if n > 1:
        lists.append(list(range(n)))
    return lists


@pytest.mark.parametrize("devices", device_lists(), ids=lambda d: "x".join(map(str, d)))
def test_multi_matches_the_oracle(lb, devices):
    data = golden_input("hamlet") * 3 + golden_input("edge_long_tokens") + b" " + golden_input("rand_longtok_s5") + golden_input("rand_heavy_s4")
    m = lb.MultiWordCount(devices, table_log2=14)
    try:
        for _ in range(2):                                   # the handle is reusable
            wc = m.run(data)
            check(wc, data)
        st = m.stats()
        assert st["job_ms"] > 0
        assert (st["record_bytes_exchanged"] > 0) == (len(devices) > 1)
        # tiny and empty inputs: most ranks get nothing
        for tiny in (b"", b"word", b"a b a\n"):
            check(m.run(tiny), tiny)
        # formatted output of the drop-in surface comes from rank 0
        wc = m.run(golden_input("hamlet"))
        assert wc.format(cpu_format=True) == golden_ref_cpu("hamlet")
    finally:
        m.close()


def test_multi_device_shards_and_growth(lb):
    """shards already on the device; a table far too small for the shard grows during the count pass"""
    from locust_b200.shuffle import plan_shards
    from locust_b200.textgen import TextGen
    data = bytes(TextGen.config("C3", total_bytes=8 << 20).host(8 << 20))
    devices = [0, 0, 0, 0]
    shards = [torch.frombuffer(bytearray(data[lo:hi]), dtype=torch.uint8).cuda() for lo, hi in plan_shards(data, len(devices))]
    m = lb.MultiWordCount(devices, table_log2=10)
    try:
        check(m.run_shards(shards), data)
        check(m.run_shards(shards), data)
    finally:
        m.close()


def test_cli_on_several_ranks(lb, tmp_path):
    """./MapReduce with LOCUST_DEVICES prints the same stdout as on one GPU"""
    exe = os.path.join(ROOT, "locust_b200", "MapReduce")
    hamlet = os.path.join(ROOT, "tests", "golden", "hamlet.txt")
    env = dict(os.environ, LOCUST_FORMAT="cpu")
    one = subprocess.run([exe, hamlet], env=env, capture_output=True, check=True).stdout
    many = subprocess.run([exe, hamlet], env=dict(env, LOCUST_DEVICES="0,0,0"), capture_output=True, check=True).stdout
    keep = lambda out: [l for l in out.split(b"\n") if l.startswith(b"print key: ")]
    assert keep(many) == keep(one) and len(keep(one)) == 5608
    assert b"".join(l + b"\n" for l in keep(many)) == golden_ref_cpu("hamlet")


def test_peer_exchange_under_torchrun(lb):
    """the real thing: one process per GPU, IPC-mapped peer buffers, flag barriers across processes, every
    exchange path against the oracle (tools/check_peer_exchange.py).  Needs two GPUs; the one-GPU box skips."""
    import sys
    n = torch.cuda.device_count()
    if n < 2:
        pytest.skip("one GPU: the ranks of this path are separate processes on separate GPUs")
    n = 2 if n < 4 else 4
    r = subprocess.run([sys.executable, "-m", "torch.distributed.run", "--nnodes=1", "--nproc-per-node", str(n), "--master-addr", "127.0.0.1",
                        "--master-port", "29733", os.path.join(ROOT, "tools", "check_peer_exchange.py")], capture_output=True, text=True, timeout=600)
    assert r.returncode == 0, r.stdout[-2000:] + r.stderr[-2000:]
    assert r.stdout.count("keys match the oracle") >= 10 and r.stdout.count("ok") >= n
