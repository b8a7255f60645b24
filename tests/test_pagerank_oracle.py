"""CPU: oracle/pagerank_oracle.c (the definition of the second job; parity unpinned -- the reference holds no
PageRank code, docs/PROPOSAL.md:21) against an independent pure-Python statement with exact integers."""
import pytest

from oracle import pyoracle
from pagerank_cases import D85, MASS, SMALL, py_pagerank, random_graph


@pytest.mark.parametrize("name", sorted(SMALL))
@pytest.mark.parametrize("iterations", [0, 1, 2, 20])
def test_small_cases(name, iterations):
    data = SMALL[name]
    assert pyoracle.pagerank(data, iterations, D85) == tuple(py_pagerank(data, iterations, D85))


@pytest.mark.parametrize("damping", [0, 1 << 31, D85, 1 << 32])
def test_damping_values(damping):
    data = random_graph(200, 1500, seed=3)
    assert pyoracle.pagerank(data, 7, damping) == tuple(py_pagerank(data, 7, damping))


def test_random_graphs_and_mass():
    for seed, zipf in ((1, False), (2, True)):
        data = random_graph(3000, 20000, seed=seed, zipf=zipf)
        keys, ranks, edges = pyoracle.pagerank(data, 10, D85)
        assert (keys, ranks, edges) == tuple(py_pagerank(data, 10, D85))
        assert keys == sorted(keys) and len(set(keys)) == len(keys)
        # integer divisions only ever round down: the mass shrinks by less than one unit per node, edge and iteration
        assert MASS - 10 * (len(keys) * 3 + edges) <= sum(ranks) <= MASS


def test_long_name_is_an_error():
    with pytest.raises(RuntimeError):
        pyoracle.pagerank(b"a b\nthisnameiswaytoolong b\n", 1, D85)
    with pytest.raises(ValueError):
        py_pagerank(b"a b\nthisnameiswaytoolong b\n", 1, D85)
