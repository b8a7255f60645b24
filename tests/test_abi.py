"""CPU-side checks of the boundary: liblocust.so loads, exports every entry point that
include/locust.h declares, refuses to run without a GPU (no CPU fallback), and the bit
tricks shared by the kernels hold on the host."""
import ctypes
import os
import re
import subprocess

import pytest

from conftest import ROOT


@pytest.fixture(scope="module")
def built():
    import __graft_entry__ as g
    if not (os.path.exists(os.path.join(ROOT, "locust_b200", "liblocust.so"))
            and os.path.exists(os.path.join(ROOT, "locust_b200", "liblocust_textgen.so"))):
        g.build()
    return True


def _declared(header):
    with open(os.path.join(ROOT, "include", header)) as f:
        src = re.sub(r"/\*.*?\*/", "", f.read(), flags=re.S)
    return sorted(set(re.findall(r"\b(locust_[a-z_0-9]+)\s*\(", src)))


def test_library_exports_every_declared_symbol(built):
    import locust_b200
    L = ctypes.CDLL(locust_b200.lib_path())
    names = _declared("locust.h")
    assert len(names) >= 18
    for n in names:
        assert hasattr(L, n), "liblocust.so does not export %s" % n
    assert set(names) == set(locust_b200.EXPORTS), "python binding and header disagree"
    L.locust_version.restype = ctypes.c_char_p
    assert b"sm_100a" in L.locust_version()


def test_header_is_plain_c(built, tmp_path):
    src = tmp_path / "t.c"
    src.write_text('#include "locust.h"\nint main(void){locust_config c; (void)c; return LOCUST_OK;}\n')
    subprocess.run(["gcc", "-std=c99", "-Wall", "-Werror", "-I", os.path.join(ROOT, "include"), "-c", str(src),
                    "-o", str(tmp_path / "t.o")], check=True)


def test_no_cpu_fallback_without_gpu(built):
    import torch
    if torch.cuda.is_available():
        pytest.skip("a GPU is present")
    import locust_b200
    with pytest.raises(locust_b200.LocustError):
        locust_b200.WordCount(device=0)


def test_product_does_not_touch_the_oracle():
    # the oracle is test infrastructure: nothing under locust_b200/ may name it
    for dirpath, _, files in os.walk(os.path.join(ROOT, "locust_b200")):
        for fn in files:
            if fn.endswith((".py", ".cu", ".cuh", ".cpp", ".h")):
                with open(os.path.join(dirpath, fn), errors="ignore") as f:
                    text = f.read()
                assert "pyoracle" not in text and "liboracle" not in text and "wordcount_oracle" not in text, fn


def test_swar_helpers_on_host(tmp_path):
    exe = tmp_path / "swar"
    subprocess.run(["g++", "-O2", "-std=c++17", "-I/usr/local/cuda/include",
                    os.path.join(ROOT, "tests", "host_swar_check.cpp"), "-o", str(exe)], check=True)
    out = subprocess.run([str(exe)], capture_output=True, text=True)
    assert out.returncode == 0 and out.stdout.strip() == "ok", out.stdout + out.stderr


def test_textgen_host_is_deterministic_and_block_independent(built):
    from locust_b200.textgen import TextGen, BLOCK
    g = TextGen.config("C2", total_bytes=64 * BLOCK)
    a = g.host(16 * BLOCK)
    b = g.host(4 * BLOCK, byte_offset=8 * BLOCK)
    assert bytes(a[8 * BLOCK:12 * BLOCK]) == bytes(b)
    assert bytes(g.host(16 * BLOCK)) == bytes(a)
    raw = bytes(a)
    assert b"\x00" not in raw and max(raw) < 0x80
    for k in range(1, 16):          # every block ends in a delimiter: any block boundary is a shard cut
        assert raw[k * BLOCK - 1] in b" \n"
    for name in ("C4", "C5"):
        t = TextGen.config(name, total_bytes=64 * BLOCK).host(8 * BLOCK)
        assert b"\x00" not in bytes(t)
