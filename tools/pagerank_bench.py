#!/usr/bin/env python
"""Second job, measured: a synthetic edge list (names "n<id>", uniform or power-law endpoints) generated on the
GPU with torch, locust_pagerank timed with device events, a bounded sample checked against the oracle.
TEST/BENCH TOOL: the product never imports the oracle."""
import argparse, json, os, sys, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
import torch
import locust_b200


def edge_text(n_nodes, n_edges, seed, zipf):
    """fixed-width lines "nXXXXXXX nXXXXXXX\\n" (hex ids) built on the device"""
    g = torch.Generator(device="cuda"); g.manual_seed(seed)
    if zipf:
        u = torch.rand(2, n_edges, device="cuda", generator=g)
        ids = (n_nodes ** u).long().clamp_(1, n_nodes) - 1          # log-uniform: a few hubs, a long tail
    else:
        ids = torch.randint(0, n_nodes, (2, n_edges), device="cuda", generator=g)
    width = 7
    hexd = torch.tensor(list(b"0123456789abcdef"), dtype=torch.uint8, device="cuda")
    line = torch.empty(n_edges, 2 * (width + 1) + 2, dtype=torch.uint8, device="cuda")
    for side in range(2):
        o = side * (width + 2)
        line[:, o] = ord("n")
        for k in range(width):
            line[:, o + 1 + k] = hexd[(ids[side] >> (4 * (width - 1 - k))) & 15]
    line[:, width + 1] = ord(" ")
    line[:, -1] = ord("\n")
    return line.reshape(-1)


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--nodes", type=int, default=1_000_000)
    ap.add_argument("--edges", type=int, default=32_000_000)
    ap.add_argument("--iterations", type=int, default=20)
    ap.add_argument("--steps", type=int, default=5)
    ap.add_argument("--zipf", action="store_true")
    ap.add_argument("--check-edges", type=int, default=2_000_000, help="size of the sample checked against the oracle")
    a = ap.parse_args()
    text = edge_text(a.nodes, a.edges, 1, a.zipf)
    torch.cuda.synchronize()
    D = int(0.85 * (1 << 32))
    out = {"job": "pagerank", "nodes_asked": a.nodes, "edges": a.edges, "bytes": int(text.numel()), "iterations": a.iterations,
           "endpoints": "log-uniform" if a.zipf else "uniform"}
    with locust_b200.WordCount(table_log2=22) as wc:
        wc.pagerank(text, a.iterations, D)                       # warm-up, sizes the table
        ms, parse = [], []
        for _ in range(a.steps):
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record(); wc.pagerank(text, a.iterations, D); e1.record(); torch.cuda.synchronize()
            ms.append(e0.elapsed_time(e1))
        t0 = []
        for _ in range(a.steps):
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record(); wc.pagerank(text, 0, D); e1.record(); torch.cuda.synchronize()
            t0.append(e0.elapsed_time(e1))
        st = wc.pagerank_stats()
        ms_med, ms0 = float(np.median(ms)), float(np.median(t0))
        out.update(nodes=st["nodes"], ms_per_job=ms_med, ms_all=ms, ms_parse_sort_only=ms0,
                   ms_per_iteration=(ms_med - ms0) / max(a.iterations, 1),
                   edges_per_s_per_iteration=a.edges / ((ms_med - ms0) / max(a.iterations, 1)) * 1e3 if a.iterations else None,
                   input_gbs_parse_only=text.numel() / ms0 / 1e6, input_gbs_whole_job=text.numel() / ms_med / 1e6)
        # parity on a bounded sample (the oracle is a single host thread)
        from oracle import pyoracle
        n = min(a.check_edges, a.edges)
        sample = text[: n * 18].cpu().numpy()
        t = time.perf_counter()
        keys, ranks, edges = pyoracle.pagerank(sample, a.iterations, D)
        cpu_s = time.perf_counter() - t
        wc.pagerank(text[: n * 18], a.iterations, D)
        got = wc.result()
        out["parity"] = {"edges_checked": edges, "nodes": len(keys), "ok": got[0] == keys and got[1] == ranks,
                         "oracle_seconds": cpu_s, "oracle_input_gbs": n * 18 / cpu_s / 1e9}
    print(json.dumps(out))
    if not out["parity"]["ok"]:
        sys.exit(1)


if __name__ == "__main__":
    main()
