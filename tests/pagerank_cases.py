"""Shared by the CPU and GPU PageRank tests: a pure-Python statement of the job (independent of the C oracle;
exact Python integers), and generators of edge-list text."""
import random

DELIMS = b" ,.-;:'()\"\t"
MASS = 1 << 48
D85 = int(0.85 * (1 << 32))


def py_pagerank(data, iterations, damping_q32):
    """(node names in byte order, ranks, edges) -- the job as include/locust.h:locust_pagerank states it"""
    edges = []
    for line in bytes(data).split(b"\n"):
        toks, cur = [], bytearray()
        for ch in line:
            if ch in DELIMS:
                if cur:
                    toks.append(bytes(cur)); cur = bytearray()
                    if len(toks) == 2:
                        break
            else:
                cur.append(ch)
        if cur and len(toks) < 2:
            toks.append(bytes(cur))
        if any(len(t) > 16 for t in toks):
            raise ValueError("node name longer than 16 bytes")
        if len(toks) == 2:
            edges.append((toks[0], toks[1]))
    nodes = sorted({n for e in edges for n in e})
    n = len(nodes)
    if n == 0:
        return [], [], 0
    outdeg = dict.fromkeys(nodes, 0)
    for s, _ in edges:
        outdeg[s] += 1
    rank = dict.fromkeys(nodes, MASS // n)
    for _ in range(iterations):
        contrib = {v: (rank[v] // outdeg[v] if outdeg[v] else 0) for v in nodes}
        dangling = sum(rank[v] for v in nodes if not outdeg[v])
        acc = dict.fromkeys(nodes, 0)
        for s, d in edges:
            acc[d] += contrib[s]
        teleport = ((MASS - ((MASS * damping_q32) >> 32)) + ((dangling * damping_q32) >> 32)) // n
        rank = {v: teleport + ((acc[v] * damping_q32) >> 32) for v in nodes}
    return nodes, [rank[v] for v in nodes], len(edges)


SMALL = {
    "empty": b"",
    "no_edges": b"lonely\n\n  \nalso,\n",
    "triangle": b"a b\nb c\nc a\n",
    "dangling": b"a b\nb c\nc a\nd a\ne sink\n",
    "self_loop_and_repeats": b"a a\na b\na b\nb a\n",
    "delimiters": b"x, y z\n  (p)\t\"q\" ignored words\n;;r:s\nlast line-without newline",
    "one_token_lines_add_no_node": b"a b\nghost\nb a\n",
    "names_of_16_bytes": b"0123456789abcdef fedcba9876543210\nfedcba9876543210 0123456789abcdef\n0123456789abcde 0123456789abcdef\n",
    "high_bytes": "été naïve\nnaïve été\nété z\n".encode("utf-8"),
    "blank_lines": b"\n\n\na b\n\n\nb c\n\n",
    "crlf_is_part_of_the_name": b"a b\r\nb\r a\n",
}


def random_graph(n_nodes, n_edges, seed, zipf=False, name_len=(1, 16)):
    rng = random.Random(seed)
    alphabet = "abcdefghijklmnopqrstuvwxyz0123456789_/"
    names = set()
    while len(names) < n_nodes:
        names.add("".join(rng.choice(alphabet) for _ in range(rng.randint(*name_len))))
    names = sorted(names)
    rng.shuffle(names)
    seps = [" ", "\t", ", ", " - ", ":"]
    out = []
    for _ in range(n_edges):
        if zipf:
            s = names[min(int(rng.paretovariate(1.1)) - 1, n_nodes - 1)]
            d = names[min(int(rng.paretovariate(0.9)) - 1, n_nodes - 1)]
        else:
            s, d = rng.choice(names), rng.choice(names)
        out.append(s + rng.choice(seps) + d)
    return ("\n".join(out) + "\n").encode()
