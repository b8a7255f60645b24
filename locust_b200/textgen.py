"""ctypes binding of liblocust_textgen.so: deterministic synthetic text for the
BASELINE.json configurations.  Bench/test infrastructure."""
import ctypes
import os

_HERE = os.path.dirname(os.path.abspath(__file__))
BLOCK = 4096
ZIPF, UNIFORM, SKEWED = 0, 1, 2

# name -> (mode, vocab, zipf_s, seed, vocab_seed, full size in bytes)
CONFIGS = {
    "C2": (ZIPF, 50_000, 1.1, 0x10C0572, 0x10C0571, 1 << 30),
    "C3": (ZIPF, 1_000_000, 1.1, 0x10C0573, 0x10C0571, 1 << 34),
    "C4": (UNIFORM, 50_000_000, 0.0, 0x10C0574, 0x10C0571, 1 << 33),
    "C5": (SKEWED, 1_000_000, 1.1, 0x10C0575, 0x10C0571, 1 << 33),
}

_lib = None


def _load():
    global _lib
    if _lib is None:
        p = os.path.join(_HERE, "liblocust_textgen.so")
        if not os.path.exists(p):
            raise RuntimeError("liblocust_textgen.so is not built: run __graft_entry__.build()")
        L = ctypes.CDLL(p)
        L.textgen_create.argtypes = [ctypes.c_int, ctypes.c_uint32, ctypes.c_double, ctypes.c_uint64, ctypes.c_uint64,
                                     ctypes.c_uint64, ctypes.c_int, ctypes.POINTER(ctypes.c_void_p)]
        L.textgen_destroy.argtypes = [ctypes.c_void_p]
        L.textgen_fill_host.argtypes = [ctypes.c_void_p, ctypes.c_void_p, ctypes.c_uint64, ctypes.c_uint64]
        L.textgen_fill_device.argtypes = [ctypes.c_void_p, ctypes.c_void_p, ctypes.c_uint64, ctypes.c_uint64,
                                          ctypes.c_void_p]
        L.textgen_last_error.restype = ctypes.c_char_p
        _lib = L
    return _lib


class TextGen:
    def __init__(self, mode, vocab, zipf_s, seed, vocab_seed, total_bytes, device=-1):
        L = _load()
        assert total_bytes % BLOCK == 0
        h = ctypes.c_void_p()
        rc = L.textgen_create(mode, vocab, zipf_s, seed, vocab_seed, total_bytes, device, ctypes.byref(h))
        if rc:
            raise RuntimeError("textgen_create: " + L.textgen_last_error().decode())
        self._L, self._h, self.total_bytes = L, h, total_bytes

    @classmethod
    def config(cls, name, total_bytes=None, device=-1, vocab=None):
        mode, v, s, seed, vseed, full = CONFIGS[name]
        return cls(mode, vocab or v, s, seed, vseed, total_bytes or full, device)

    def host(self, nbytes, byte_offset=0):
        """numpy uint8 array with bytes [byte_offset, byte_offset + nbytes) of the text"""
        import numpy as np
        out = np.empty(nbytes, dtype=np.uint8)
        rc = self._L.textgen_fill_host(self._h, out.ctypes.data, nbytes, byte_offset)
        if rc:
            raise RuntimeError("textgen_fill_host: " + self._L.textgen_last_error().decode())
        return out

    def fill_device(self, tensor, byte_offset=0, stream=None):
        """fills a contiguous torch uint8 CUDA tensor in place"""
        n = tensor.numel()
        rc = self._L.textgen_fill_device(self._h, tensor.data_ptr(), n, byte_offset, stream)
        if rc:
            raise RuntimeError("textgen_fill_device: " + self._L.textgen_last_error().decode())
        return tensor

    def close(self):
        if self._h:
            self._L.textgen_destroy(self._h)
            self._h = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass
