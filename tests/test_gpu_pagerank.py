"""Second job through the C ABI (locust_pagerank) against oracle/pagerank_oracle.c, bit for bit.  The reference
names PageRank as a milestone and holds no code for it (docs/PROPOSAL.md:21): parity is against our own
definition ("unpinned"), which tests/test_pagerank_oracle.py checks against exact Python integers."""
import numpy as np
import pytest

from pagerank_cases import D85, SMALL, random_graph

pytestmark = pytest.mark.gpu

torch = pytest.importorskip("torch")


@pytest.fixture(scope="module")
def lb():
    if not torch.cuda.is_available():
        pytest.skip("no CUDA device")
    import locust_b200
    locust_b200.load()
    return locust_b200


def check(wc, data, iterations, damping=D85):
    from oracle import pyoracle
    keys, ranks, edges = pyoracle.pagerank(data, iterations, damping)
    wc.pagerank(data, iterations, damping)
    got_keys, got_ranks, got_vals = wc.result()
    assert got_keys == keys
    assert got_ranks == ranks
    assert got_vals == [int(v) for v in np.concatenate([[0], np.cumsum(np.array(ranks, dtype=np.uint64))[:-1]])] if ranks else got_vals == []
    st = wc.pagerank_stats()
    assert st == {"nodes": len(keys), "edges": edges}
    return keys, ranks


@pytest.mark.parametrize("name", sorted(SMALL))
def test_small_cases(lb, name):
    with lb.WordCount() as wc:
        for iterations in (0, 1, 2, 20):
            check(wc, SMALL[name], iterations)


@pytest.mark.parametrize("damping", [0, 1 << 31, D85, 1 << 32])
def test_damping_values(lb, damping):
    with lb.WordCount() as wc:
        check(wc, random_graph(200, 1500, seed=3), 7, damping)


@pytest.mark.parametrize("zipf", [False, True])
def test_random_graphs(lb, zipf):
    data = random_graph(100_000, 1_000_000, seed=5, zipf=zipf)
    with lb.WordCount() as wc:
        check(wc, data, 10)
        check(wc, data, 3)            # the same context again: nothing of the last job is left behind


def test_many_heavily_linked_nodes(lb):
    """edges into nodes with >= 8192 in-edges are added up per block in a 1024-entry table in shared memory; with 60
    such nodes some of them meet in one entry and take the direct path -- every path must give the same integers"""
    import random
    rng = random.Random(17)
    lines = []
    for h in range(60):
        for i in range(9000 + 100 * h):
            lines.append("s%d hub%d" % (rng.randrange(50_000), h))
    for i in range(100_000):
        lines.append("s%d t%d" % (rng.randrange(50_000), rng.randrange(50_000)))
    rng.shuffle(lines)
    data = ("\n".join(lines) + "\n").encode()
    with lb.WordCount() as wc:
        check(wc, data, 6)


def test_table_grows_with_the_graph(lb):
    data = random_graph(60_000, 200_000, seed=9)
    with lb.WordCount(table_log2=12) as wc:         # 4096 slots for ~60 000 nodes
        check(wc, data, 5)


def test_device_input_at_any_alignment(lb):
    data = random_graph(5000, 40_000, seed=11)
    with lb.WordCount() as wc:
        for shift in (0, 1, 7, 15):
            t = torch.zeros(len(data) + 64, dtype=torch.uint8, device="cuda")
            t[shift:shift + len(data)] = torch.frombuffer(bytearray(data), dtype=torch.uint8).cuda()
            torch.cuda.synchronize()
            from oracle import pyoracle
            keys, ranks, edges = pyoracle.pagerank(data, 4, D85)
            wc.pagerank(t[shift:shift + len(data)], 4, D85)
            got = wc.result()
            assert got[0] == keys and got[1] == ranks


def test_long_node_name_is_refused(lb):
    with lb.WordCount() as wc:
        with pytest.raises(RuntimeError, match="16 bytes"):
            wc.pagerank(b"a b\nthisnameiswaytoolong b\n", 1)
        # the context is still good for both jobs
        check(wc, SMALL["dangling"], 3)
        from oracle import pyoracle
        text = b"the cat and the hat and the bat\n"
        wc.run(text)
        ref = pyoracle.wordcount(text, mode="fast", formats=False)
        assert wc.result()[:2] == (ref.keys, ref.counts)
        with pytest.raises(RuntimeError):
            wc.pagerank_stats()


def test_formatted_output_lists_nodes_and_ranks(lb):
    with lb.WordCount() as wc:
        keys, ranks = check(wc, SMALL["triangle"], 2)
        text = wc.format(cpu_format=True)
        for k, r in zip(keys, ranks):
            assert k in text and str(r).encode() in text
