#!/usr/bin/env python
"""Regenerates the golden vectors in this directory.

Runs the reference's own CPU branch (oracle/_ref/MapReduce_cpu, built by
oracle/build_ref.sh from the sources under /root/reference; main.cu with
GPU_IMPLEMENTATION 0) on each input below and stores the `print key:` lines it
prints (main.cu:286) next to the input as `<name>.ref_cpu.txt`.

Inputs stay inside the reference's fixed capacities (main.cu:18-20, KeyValue.h:8-9:
at most 5 800 lines, lines shorter than 100 bytes) so the verbatim binary is a valid
oracle for them; `cap22` is the one case where the 20-tokens-per-line cap
(main.cu:300-303) fires.

Only needed when /root/reference is mounted; the outputs are committed.
"""
import hashlib
import os
import random
import shutil
import subprocess
import sys

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))
REF_BIN = os.path.join(ROOT, "oracle", "_ref", "MapReduce_cpu")
REF_HAMLET = "/root/reference/hamlet.txt"

DELIMS = " ,.-;:'()\"\t"


def random_text(seed, nlines, alphabet, max_tok_per_line=18, max_tok_len=12, heavy=None):
    rng = random.Random(seed)
    lines = []
    for _ in range(nlines):
        ntok = rng.randint(0, max_tok_per_line)
        parts = []
        if rng.random() < 0.3:
            parts.append("".join(rng.choice(DELIMS) for _ in range(rng.randint(1, 3))))
        for _ in range(ntok):
            if heavy and rng.random() < 0.5:
                tok = rng.choice(heavy)
            else:
                tok = "".join(rng.choice(alphabet) for _ in range(rng.randint(1, max_tok_len)))
            parts.append(tok)
            parts.append("".join(rng.choice(DELIMS) for _ in range(rng.choice([1, 1, 1, 2, 3]))))
        line = "".join(parts)[:98]
        lines.append(line)
    return ("\n".join(lines) + "\n").encode("ascii")


def inputs():
    with open(REF_HAMLET, "rb") as f:
        yield "hamlet", f.read()
    # edge probes of SURVEY.md section 8c
    yield "edge_basic", (
        b"x--y;;z\n"
        b"(foo)\"bar\"'baz'\n"
        b"\n"
        b" ,.-;:'()\"\t\n"
        b"Z z z\n"
        b"tab\tseparated\twords tab\n"
        b"a ab abc abcd abc ab a\n"
        b"!bang ?query [bracket] under_score slash/ 123 4.5 6-7\n"
        b"tilde~ caret^ back\\slash pipe| at@ hash# dollar$ percent% amp& star* plus+ eq=\n"
        b"last line without newline")
    yield "edge_cap22", (
        b"a b c d e f g h i j k l m n o p q r s t u v\n"
        b"u v w\n")
    yield "edge_cr", b"one two\r\nthree four\r\nfive\r\n"
    yield "edge_single", b"word"
    yield "edge_delims_only", b" \t \n...\n---\n"
    yield "edge_long_tokens", (
        b"supercalifragilisticexpialidocious antidisestablishmentarianism\n"
        b"supercalifragilisticexpialidocious supercalifragilisticexpialidocioux\n"
        b"abcdefghijklmnop abcdefghijklmnopq abcdefghijklmnopqr abcdefghijklmno\n"
        b"abcdefgh abcdefghi abcdefg abcdefgh\n"
        + b"a" * 97 + b"\n" + b"a" * 96 + b"b\n" + b"a" * 97 + b"\n" + b"a" * 96 + b" a\n")
    lower = "abcdefghijklmnopqrstuvwxyz"
    yield "rand_lower_s1", random_text(1, 400, lower, max_tok_len=4)
    yield "rand_lower_s2", random_text(2, 800, "abc", max_tok_len=3)
    printable = "".join(chr(c) for c in range(0x21, 0x7f) if chr(c) not in DELIMS)
    yield "rand_print_s3", random_text(3, 600, printable)
    yield "rand_heavy_s4", random_text(4, 1000, lower, heavy=["the", "a", "of", "zebra"])
    yield "rand_longtok_s5", random_text(5, 500, "ab", max_tok_per_line=4, max_tok_len=40)


def main():
    if not os.path.exists(REF_BIN):
        sys.exit("run oracle/build_ref.sh first (needs /root/reference)")
    manifest = []
    for name, data in inputs():
        path = os.path.join(HERE, name + ".txt")
        with open(path, "wb") as f:
            f.write(data)
        out = subprocess.run([REF_BIN, path], check=True, capture_output=True).stdout
        lines = [l for l in out.split(b"\n") if l.startswith(b"print key: ")]
        golden = b"".join(l + b"\n" for l in lines)
        with open(os.path.join(HERE, name + ".ref_cpu.txt"), "wb") as f:
            f.write(golden)
        warn = out.count(b"WARN: Exceeded emit limit")
        manifest.append("%s\t%d\t%s\t%d\t%s\t%d" % (
            name, len(data), hashlib.sha256(data).hexdigest(), len(lines),
            hashlib.sha256(golden).hexdigest(), warn))
    with open(os.path.join(HERE, "MANIFEST.tsv"), "w") as f:
        f.write("name\tinput_bytes\tinput_sha256\tkeys\tref_cpu_sha256\twarn_lines\n")
        f.write("\n".join(manifest) + "\n")
    print("\n".join(manifest))


if __name__ == "__main__":
    main()
