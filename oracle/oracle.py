"""TEST INFRASTRUCTURE ONLY -- ctypes loaders for the CPU oracle.

Two checkers live here:

* ``Oracle``    -- oracle/_build/liboracle.so, our C restatement of the reference host codec
                   (oracle/snappy_oracle.c; each function cites the reference file:line).
* ``Reference`` -- oracle/_ref/libsnappy_ref.so, the reference's OWN host codec compiled
                   unmodified from /root/reference/snappy/*.c (oracle/Makefile, target ``ref``).

Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / ``--impl reference`` legs may
import this module.  The product package never does (tests/test_no_cpu_fallback.py enforces it).
"""
from __future__ import annotations

import ctypes as C
import os
import subprocess

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
ORACLE_SO = os.path.join(_HERE, "_build", "liboracle.so")
REF_SO = os.path.join(_HERE, "_ref", "libsnappy_ref.so")
REF_BIN = os.path.join(_HERE, "_ref", "dpu_snappy_ref")
REFMAIN_BIN = os.path.join(_HERE, "_ref", "dpu_snappy_refmain")  # the reference's main() linked against libpim_snappy.so
REFERENCE_ROOT = "/root/reference"

OK, INVALID_INPUT, BUFFER_TOO_SMALL = 0, 1, 2


def build(with_ref: bool | None = None) -> None:
    """Compile the restatement and, when /root/reference is present, the reference itself."""
    subprocess.run(["make", "-s", "-C", _HERE, "all"], check=True)
    if with_ref is None:
        with_ref = os.path.isdir(os.path.join(REFERENCE_ROOT, "snappy"))
    if with_ref:
        subprocess.run(["make", "-s", "-C", _HERE, "ref"], check=True)
        if os.path.exists(os.path.join(_HERE, "..", "pim-compression_b200", "libpim_snappy.so")):
            subprocess.run(["make", "-s", "-C", _HERE, "refmain"], check=True)  # the reference's main() bound to the product library


def _u8(buf) -> np.ndarray:
    a = np.frombuffer(buf, dtype=np.uint8) if not isinstance(buf, np.ndarray) else buf
    return np.ascontiguousarray(a, dtype=np.uint8)


def _ptr(a: np.ndarray):
    return a.ctypes.data_as(C.POINTER(C.c_uint8))


class Oracle:
    """Our restatement (kind = "port")."""

    def __init__(self) -> None:
        if not os.path.exists(ORACLE_SO):
            build(with_ref=False)
        L = C.CDLL(ORACLE_SO)
        L.oracle_max_compressed_length.restype = C.c_size_t
        L.oracle_max_compressed_length.argtypes = [C.c_size_t]
        L.oracle_file_bound.restype = C.c_size_t
        L.oracle_file_bound.argtypes = [C.c_size_t, C.c_uint32]
        L.oracle_compress.restype = C.c_int
        L.oracle_compress.argtypes = [C.POINTER(C.c_uint8), C.c_size_t, C.c_uint32, C.POINTER(C.c_uint8), C.c_size_t, C.POINTER(C.c_size_t)]
        for name in ("oracle_decompress", "oracle_decompress_strict"):
            f = getattr(L, name)
            f.restype = C.c_int
            f.argtypes = [C.POINTER(C.c_uint8), C.c_size_t, C.POINTER(C.c_uint8), C.c_size_t, C.POINTER(C.c_size_t)]
        L.oracle_read_header.restype = C.c_size_t
        L.oracle_read_header.argtypes = [C.POINTER(C.c_uint8), C.c_size_t, C.POINTER(C.c_uint32), C.POINTER(C.c_uint32)]
        self.lib = L

    def file_bound(self, n: int, block_size: int) -> int:
        return int(self.lib.oracle_file_bound(n, block_size))

    def compress(self, data, block_size: int = 32768) -> np.ndarray:
        src = _u8(data)
        cap = self.file_bound(src.size, block_size)
        out = np.empty(cap, dtype=np.uint8)
        n = C.c_size_t(0)
        rc = self.lib.oracle_compress(_ptr(src), src.size, block_size, _ptr(out), cap, C.byref(n))
        if rc != OK:
            raise RuntimeError(f"oracle_compress rc={rc}")
        return out[: n.value].copy()

    def read_header(self, stream):
        src = _u8(stream)
        dlen, bs = C.c_uint32(0), C.c_uint32(0)
        k = self.lib.oracle_read_header(_ptr(src), src.size, C.byref(dlen), C.byref(bs))
        if k == 0:
            raise ValueError("malformed header")
        return int(k), int(dlen.value), int(bs.value)

    def _dec(self, fn, stream, cap=None):
        src = _u8(stream)
        if cap is None:
            _, dlen, _ = self.read_header(src)
            cap = dlen
        out = np.zeros(max(cap, 1), dtype=np.uint8)
        n = C.c_size_t(0)
        rc = fn(_ptr(src), src.size, _ptr(out), cap, C.byref(n))
        return rc, out[: n.value] if rc == OK else out[:0]

    def decompress(self, stream, cap=None):
        """Reference semantics (with its quirks). Returns (status, bytes)."""
        return self._dec(self.lib.oracle_decompress, stream, cap)

    def decompress_strict(self, stream, cap=None):
        return self._dec(self.lib.oracle_decompress_strict, stream, cap)


class _HostBuf(C.Structure):  # dpu_snappy.h:37-44
    _fields_ = [("file_name", C.c_char_p), ("buffer", C.POINTER(C.c_uint8)), ("curr", C.POINTER(C.c_uint8)),
                ("length", C.c_ulong), ("max", C.c_ulong)]


class _Runtime(C.Structure):  # dpu_snappy.h:47-55
    _fields_ = [(k, C.c_double) for k in ("pre", "d_alloc", "load", "copy_in", "run", "copy_out", "d_free")]


class Reference:
    """The reference's own host codec (kind = "reference"); exists only if oracle/_ref was built."""

    @staticmethod
    def available() -> bool:
        return os.path.exists(REF_SO)

    def __init__(self) -> None:
        if not self.available():
            build()
        L = C.CDLL(REF_SO)
        L.setup_compression.restype = None
        L.setup_compression.argtypes = [C.POINTER(_HostBuf), C.POINTER(_HostBuf), C.POINTER(_Runtime)]
        L.snappy_compress_host.restype = C.c_int
        L.snappy_compress_host.argtypes = [C.POINTER(_HostBuf), C.POINTER(_HostBuf), C.c_uint32]
        L.setup_decompression.restype = C.c_int
        L.setup_decompression.argtypes = [C.POINTER(_HostBuf), C.POINTER(_HostBuf), C.POINTER(_Runtime)]
        L.snappy_decompress_host.restype = C.c_int
        L.snappy_decompress_host.argtypes = [C.POINTER(_HostBuf), C.POINTER(_HostBuf)]
        self.lib = L
        self.libc = C.CDLL(None)
        self.libc.free.argtypes = [C.c_void_p]

    def compress(self, data, block_size: int = 32768) -> np.ndarray:
        src = _u8(data)
        # the reference's own bound (32 + n + n/6) ignores the 4 B block headers:
        # only call this with inputs where that is safe (tests keep block_size >= 64 or tiny n).
        inp, out, rt = _HostBuf(), _HostBuf(), _Runtime()
        inp.buffer = _ptr(src)
        inp.curr = inp.buffer
        inp.length = src.size
        inp.max = 2**63
        out.max = 2**63
        self.lib.setup_compression(C.byref(inp), C.byref(out), C.byref(rt))
        if src.size == 0:  # setup mallocs 0 bytes; give it room for the 4..6 header bytes
            self.libc.free(C.cast(out.buffer, C.c_void_p))
            keep = np.zeros(16, dtype=np.uint8)
            out.buffer = _ptr(keep)
            out.curr = out.buffer
            rc = self.lib.snappy_compress_host(C.byref(inp), C.byref(out), block_size)
            assert rc == OK
            return keep[: out.length].copy()
        rc = self.lib.snappy_compress_host(C.byref(inp), C.byref(out), block_size)
        assert rc == OK
        res = np.ctypeslib.as_array(out.buffer, shape=(out.length,)).copy()
        self.libc.free(C.cast(out.buffer, C.c_void_p))
        return res

    def decompress(self, stream):
        src = _u8(stream)
        inp, out, rt = _HostBuf(), _HostBuf(), _Runtime()
        inp.buffer = _ptr(src)
        inp.curr = inp.buffer
        inp.length = src.size
        inp.max = 2**63
        out.max = 2**63
        rc = self.lib.setup_decompression(C.byref(inp), C.byref(out), C.byref(rt))
        if rc != OK:
            return rc, np.zeros(0, dtype=np.uint8)
        rc = self.lib.snappy_decompress_host(C.byref(inp), C.byref(out))
        res = np.ctypeslib.as_array(out.buffer, shape=(max(out.length, 1),))[: out.length].copy() if rc == OK else np.zeros(0, dtype=np.uint8)
        self.libc.free(C.cast(out.buffer, C.c_void_p))
        return rc, res
