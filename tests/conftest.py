import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)

GOLDEN = os.path.join(ROOT, "tests", "golden")


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA device (run on the B200 box)")


def golden_cases():
    """[(name, input_bytes, keys, ref_cpu_sha256, warn_lines)] from tests/golden/MANIFEST.tsv"""
    rows = []
    with open(os.path.join(GOLDEN, "MANIFEST.tsv")) as f:
        next(f)
        for line in f:
            name, nbytes, insha, keys, outsha, warn = line.rstrip("\n").split("\t")
            rows.append((name, int(nbytes), insha, int(keys), outsha, int(warn)))
    return rows


def golden_input(name):
    with open(os.path.join(GOLDEN, name + ".txt"), "rb") as f:
        return f.read()


def golden_ref_cpu(name):
    with open(os.path.join(GOLDEN, name + ".ref_cpu.txt"), "rb") as f:
        return f.read()
