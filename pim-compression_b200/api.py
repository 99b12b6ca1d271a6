"""ctypes binding of include/pim_snappy.h -- mirrors the reference's C interface one to one.

Reference interface being mirrored (relative to /root/reference/snappy/):
  host_buffer_context / program_runtime / snappy_status   dpu_snappy.h:21-55
  setup_compression / snappy_compress_dpu                 snappy_compress.h:16,37
  setup_decompression / snappy_decompress_dpu             snappy_decompress.h:15,34
"""
from __future__ import annotations

import ctypes as C
import os
import subprocess

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(_HERE, "libpim_snappy.so")
SYNTH_LIB_PATH = os.path.join(_HERE, "libpim_synth.so")  # workload generators (include/pim_synth.h): no codec inside

SNAPPY_OK, SNAPPY_INVALID_INPUT, SNAPPY_BUFFER_TOO_SMALL = 0, 1, 2
SYNTH_TEXT, SYNTH_RANDOM, SYNTH_MIX, SYNTH_BIN = 0, 1, 2, 3


class SnappyError(RuntimeError):
    def __init__(self, status: int, what: str = ""):
        super().__init__(f"snappy_status={status} {what}".strip())
        self.status = status


class HostBufferContext(C.Structure):  # dpu_snappy.h:37-44
    _fields_ = [("file_name", C.c_char_p), ("buffer", C.c_void_p), ("curr", C.c_void_p),
                ("length", C.c_ulong), ("max", C.c_ulong)]


class ProgramRuntime(C.Structure):  # dpu_snappy.h:47-55
    _fields_ = [(k, C.c_double) for k in ("pre", "d_alloc", "load", "copy_in", "run", "copy_out", "d_free")]

    def as_dict(self):
        return {k: getattr(self, k) for k, _ in self._fields_}


def build(force: bool = False) -> str:
    """Compile libpim_snappy.so and the dpu_snappy CLI for sm_100a (nvcc cross-compiles without a GPU)."""
    if force:
        subprocess.run(["make", "-s", "-C", _HERE, "clean"], check=True)
    subprocess.run(["make", "-s", "-C", _HERE, "all"], check=True)
    return LIB_PATH


_lib = None


def lib() -> C.CDLL:
    """Load the native library.  No fallback: a missing library is an error."""
    global _lib
    if _lib is not None:
        return _lib
    if not os.path.exists(LIB_PATH):
        raise ImportError(f"{LIB_PATH} is not built; run `python -c 'import __graft_entry__ as g; g.build()'` "
                          f"or `make -C pim-compression_b200`.  There is no Python/CPU fallback.")
    L = C.CDLL(LIB_PATH)
    P = C.POINTER
    hb, rt = P(HostBufferContext), P(ProgramRuntime)
    sig = {
        "setup_compression": (None, [hb, hb, rt]),
        "setup_decompression": (C.c_int, [hb, hb, rt]),
        "snappy_compress_dpu": (C.c_int, [hb, hb, C.c_uint32, rt]),
        "snappy_decompress_dpu": (C.c_int, [hb, hb, rt]),
        "snappy_compress_gpu": (C.c_int, [hb, hb, C.c_uint32, rt]),
        "snappy_decompress_gpu": (C.c_int, [hb, hb, rt]),
        "pim_snappy_host_alloc": (C.c_void_p, [C.c_size_t]),
        "pim_snappy_host_free": (None, [C.c_void_p]),
        "pim_snappy_set_devices": (C.c_int, [P(C.c_int), C.c_int]),
        "pim_snappy_device_count": (C.c_int, []),
        "pim_snappy_set_chunk_bytes": (None, [C.c_uint64]),
        "pim_snappy_set_write_file": (None, [C.c_int]),
        "pim_snappy_set_extended_blocks": (None, [C.c_int]),
        "pim_snappy_shard_range": (None, [C.c_uint64, C.c_int, C.c_int, C.POINTER(C.c_uint64), C.POINTER(C.c_uint64)]),
        "pim_snappy_num_blocks": (C.c_uint64, [C.c_uint64, C.c_uint32]),
        "pim_snappy_compress_bound": (C.c_uint64, [C.c_uint64, C.c_uint32]),
        "pim_snappy_compress_scratch_bytes": (C.c_uint64, [C.c_uint64, C.c_uint32]),
        "pim_snappy_compress_device": (C.c_int, [C.c_void_p, C.c_uint64, C.c_uint32, C.c_void_p, C.c_uint64, C.c_void_p, C.c_void_p, C.c_void_p]),
        "pim_snappy_compress_device_noalloc": (C.c_int, [C.c_void_p, C.c_uint64, C.c_uint32, C.c_void_p, C.c_void_p, C.c_void_p]),
        "pim_snappy_decompress_device": (C.c_int, [C.c_void_p, C.c_void_p, C.c_uint64, C.c_uint32, C.c_uint64, C.c_void_p, C.c_void_p, C.c_void_p]),
        "pim_snappy_compress_batch_device": (C.c_int, [C.c_int, P(C.c_void_p), P(C.c_uint64), C.c_uint32, P(C.c_void_p), P(C.c_uint64), P(C.c_void_p),
                                                       P(C.c_void_p), C.c_void_p]),
        "pim_snappy_decompress_batch_device": (C.c_int, [C.c_int, P(C.c_void_p), P(C.c_void_p), P(C.c_uint64), C.c_uint32, P(C.c_uint64), P(C.c_void_p),
                                                         C.c_void_p, C.c_void_p]),
        "pim_snappy_gather_shards": (C.c_int, [C.c_int, P(C.c_int), P(C.c_void_p), P(C.c_uint64), P(C.c_void_p), P(C.c_uint64), C.c_int,
                                               C.c_void_p, C.c_void_p, C.c_void_p]),
        "pim_snappy_launch_count": (C.c_uint64, []),
        "pim_snappy_profile": (None, [C.c_int]),
        "pim_snappy_profile_read": (C.c_int, [P(C.c_double), P(C.c_uint64)]),
        "pim_snappy_version": (C.c_char_p, []),
        "get_runtime": (C.c_double, [C.c_void_p, C.c_void_p]),
    }
    for name, (res, args) in sig.items():
        f = getattr(L, name)  # AttributeError if the library does not export what the header declares
        f.restype, f.argtypes = res, args
    _lib = L
    return L


EXPORTED_SYMBOLS = [
    "setup_compression", "setup_decompression", "snappy_compress_dpu", "snappy_decompress_dpu",
    "snappy_compress_gpu", "snappy_decompress_gpu", "get_runtime", "pim_snappy_host_alloc", "pim_snappy_host_free",
    "pim_snappy_set_devices", "pim_snappy_device_count", "pim_snappy_set_chunk_bytes", "pim_snappy_set_write_file", "pim_snappy_set_extended_blocks", "pim_snappy_shard_range", "pim_snappy_num_blocks", "pim_snappy_compress_bound",
    "pim_snappy_compress_scratch_bytes", "pim_snappy_compress_device", "pim_snappy_compress_device_noalloc", "pim_snappy_decompress_device", "pim_snappy_compress_batch_device",
    "pim_snappy_decompress_batch_device", "pim_snappy_gather_shards",
    "pim_snappy_launch_count", "pim_snappy_profile", "pim_snappy_profile_read", "pim_snappy_version",
]


def set_devices(ids) -> int:
    arr = (C.c_int * len(ids))(*ids)
    return int(lib().pim_snappy_set_devices(arr, len(ids)))


def shard_range(nblocks: int, n_shards: int, shard: int) -> tuple[int, int]:
    """Contiguous block range of one shard (one GPU / one rank), as the reference splits blocks over DPUs."""
    lo, hi = C.c_uint64(0), C.c_uint64(0)
    lib().pim_snappy_shard_range(nblocks, n_shards, shard, C.byref(lo), C.byref(hi))
    return int(lo.value), int(hi.value)


def set_chunk_bytes(n: int) -> None:
    """Chunk size of the streamed host path (0: whole shard in one piece)."""
    lib().pim_snappy_set_chunk_bytes(n)


def set_extended_blocks(enable: bool) -> None:
    """Extended-domain mode of the compressor (block sizes up to 1 MiB, the reference's bytes incl. its 16-bit truncations)."""
    lib().pim_snappy_set_extended_blocks(1 if enable else 0)


def device_count() -> int:
    return int(lib().pim_snappy_device_count())


def launch_count() -> int:
    return int(lib().pim_snappy_launch_count())


def profile(enable: bool) -> None:
    lib().pim_snappy_profile(1 if enable else 0)


def profile_read() -> dict:
    ms, n = (C.c_double * 4)(), (C.c_uint64 * 4)()
    lib().pim_snappy_profile_read(ms, n)
    names = ("compress", "scan", "compact", "decompress")
    return {k: {"ms": ms[i], "launches": int(n[i])} for i, k in enumerate(names)}


class PinnedBuffer:
    """Page-locked host bytes from pim_snappy_host_alloc, viewed as a numpy uint8 array."""

    def __init__(self, nbytes: int):
        self.nbytes = int(nbytes)
        self.ptr = lib().pim_snappy_host_alloc(max(self.nbytes, 1))
        if not self.ptr:
            raise MemoryError("pim_snappy_host_alloc failed (no CUDA device?)")
        self.array = np.ctypeslib.as_array(C.cast(self.ptr, C.POINTER(C.c_uint8)), shape=(max(self.nbytes, 1),))[: self.nbytes]

    def free(self):
        if self.ptr:
            self.array = None
            lib().pim_snappy_host_free(self.ptr)
            self.ptr = None

    def __del__(self):
        try:
            self.free()
        except Exception:
            pass


def _as_u8(data) -> np.ndarray:
    a = data if isinstance(data, np.ndarray) else np.frombuffer(data, dtype=np.uint8)
    return np.ascontiguousarray(a, dtype=np.uint8)


def compress(data, block_size: int = 32768, runtime: ProgramRuntime | None = None, in_ptr: int | None = None,
             in_len: int | None = None) -> np.ndarray:
    """setup_compression + snappy_compress_dpu on HOST buffers; returns the file image (a copy)."""
    L = lib()
    if in_ptr is None:
        src = _as_u8(data)
        in_ptr, in_len = src.ctypes.data, src.size
    inp, out = HostBufferContext(), HostBufferContext()
    rt = runtime if runtime is not None else ProgramRuntime()
    inp.buffer = inp.curr = in_ptr
    inp.length = in_len
    inp.max = out.max = 2**63
    L.setup_compression(C.byref(inp), C.byref(out), C.byref(rt))
    if not out.buffer:
        raise SnappyError(SNAPPY_INVALID_INPUT, "setup_compression could not allocate page-locked memory")
    try:
        st = L.snappy_compress_dpu(C.byref(inp), C.byref(out), block_size, C.byref(rt))
        if st != SNAPPY_OK:
            raise SnappyError(st, "snappy_compress_dpu")
        return np.ctypeslib.as_array(C.cast(out.buffer, C.POINTER(C.c_uint8)), shape=(max(out.length, 1),))[: out.length].copy()
    finally:
        L.pim_snappy_host_free(out.buffer)


def decompress(stream, runtime: ProgramRuntime | None = None, max_len: int = 2**63) -> np.ndarray:
    """setup_decompression + snappy_decompress_dpu on HOST buffers; returns the decoded bytes (a copy)."""
    L = lib()
    src = _as_u8(stream)
    inp, out = HostBufferContext(), HostBufferContext()
    rt = runtime if runtime is not None else ProgramRuntime()
    inp.buffer = inp.curr = src.ctypes.data
    inp.length = src.size
    inp.max = 2**63
    out.max = max_len
    st = L.setup_decompression(C.byref(inp), C.byref(out), C.byref(rt))
    if st != SNAPPY_OK:
        raise SnappyError(st, "setup_decompression")
    try:
        st = L.snappy_decompress_dpu(C.byref(inp), C.byref(out), C.byref(rt))
        if st != SNAPPY_OK:
            raise SnappyError(st, "snappy_decompress_dpu")
        return np.ctypeslib.as_array(C.cast(out.buffer, C.POINTER(C.c_uint8)), shape=(max(out.length, 1),))[: out.length].copy()
    finally:
        L.pim_snappy_host_free(out.buffer)


def block_offsets(stream) -> tuple[int, int, int, np.ndarray]:
    """Host header walk (snappy_decompress.c:317-340): returns (header_bytes, dlen, block_size, off[nblocks+1])
    with offsets relative to the first block header.  Pure bookkeeping -- no element is decoded."""
    src = _as_u8(stream)
    pos, vals = 0, []
    for _ in range(2):
        v, shift = 0, 0
        for k in range(5):
            if pos >= src.size:
                raise SnappyError(SNAPPY_INVALID_INPUT, "truncated header")
            c = int(src[pos]); pos += 1
            v |= (c & 0x7F) << shift
            shift += 7
            if not c & 0x80:
                break
        else:
            raise SnappyError(SNAPPY_INVALID_INPUT, "bad varint")
        vals.append(v)
    dlen, bs = vals
    nb = (dlen + bs - 1) // bs if bs else 0
    off = np.zeros(nb + 1, dtype=np.uint64)
    p = pos
    for i in range(nb):
        off[i] = p - pos
        if p + 4 > src.size:
            raise SnappyError(SNAPPY_INVALID_INPUT, "truncated block header")
        cs = int(src[p]) | int(src[p + 1]) << 8 | int(src[p + 2]) << 16 | int(src[p + 3]) << 24
        p += 4 + cs
    off[nb] = p - pos
    return pos, dlen, bs, off


# ---- device-resident API (raw device pointers; torch tensors are only the allocator) ----------

def compress_device(d_in_ptr: int, n: int, block_size: int, d_scratch_ptr: int, scratch_bytes: int, d_out_ptr: int,
                    d_off_ptr: int, stream: int = 0) -> None:
    st = lib().pim_snappy_compress_device(d_in_ptr, n, block_size, d_scratch_ptr, scratch_bytes, d_out_ptr, d_off_ptr, stream)
    if st != SNAPPY_OK:
        raise SnappyError(st, "pim_snappy_compress_device")


NOALLOC_SCRATCH_BYTES = 64 << 20  # PIM_SNAPPY_NOALLOC_SCRATCH_BYTES


def compress_device_noalloc(d_in_ptr: int, n: int, block_size: int, d_out_ptr: int, d_off_ptr: int, stream: int = 0) -> None:
    """pim_snappy_compress_device_noalloc: scratch in the library's built-in device array, no allocation in the call."""
    st = lib().pim_snappy_compress_device_noalloc(d_in_ptr, n, block_size, d_out_ptr, d_off_ptr, stream)
    if st != SNAPPY_OK:
        raise SnappyError(st, "pim_snappy_compress_device_noalloc")


def compress_batch_device(d_in_ptrs, lens, block_size: int, d_scratch_ptrs, scratch_bytes, d_out_ptrs, d_off_ptrs, stream: int = 0) -> None:
    """pim_snappy_compress_batch_device: several device-resident streams in one compress launch."""
    n = len(d_in_ptrs)
    st = lib().pim_snappy_compress_batch_device(n, (C.c_void_p * n)(*d_in_ptrs), (C.c_uint64 * n)(*lens), block_size, (C.c_void_p * n)(*d_scratch_ptrs),
                                                (C.c_uint64 * n)(*scratch_bytes), (C.c_void_p * n)(*d_out_ptrs), (C.c_void_p * n)(*d_off_ptrs), stream)
    if st != SNAPPY_OK:
        raise SnappyError(st, "pim_snappy_compress_batch_device")


def decompress_batch_device(d_comp_ptrs, d_off_ptrs, nblocks, block_size: int, out_lens, d_out_ptrs, d_status_ptr: int, stream: int = 0) -> None:
    """pim_snappy_decompress_batch_device: several block streams in one decompress launch; d_status: one u32 per stream."""
    n = len(d_comp_ptrs)
    st = lib().pim_snappy_decompress_batch_device(n, (C.c_void_p * n)(*d_comp_ptrs), (C.c_void_p * n)(*d_off_ptrs), (C.c_uint64 * n)(*nblocks), block_size,
                                                  (C.c_uint64 * n)(*out_lens), (C.c_void_p * n)(*d_out_ptrs), d_status_ptr, stream)
    if st != SNAPPY_OK:
        raise SnappyError(st, "pim_snappy_decompress_batch_device")


def gather_shards(devices, shard_ptrs, shard_bytes, off_ptrs, nblocks, dst_device: int, d_dst_ptr: int, d_dst_off_ptr: int, stream: int = 0) -> None:
    """pim_snappy_gather_shards: compressed shards on several GPUs -> one block stream + offsets on dst_device."""
    n = len(devices)
    st = lib().pim_snappy_gather_shards(n, (C.c_int * n)(*devices), (C.c_void_p * n)(*shard_ptrs), (C.c_uint64 * n)(*shard_bytes),
                                        (C.c_void_p * n)(*off_ptrs), (C.c_uint64 * n)(*nblocks), dst_device, d_dst_ptr, d_dst_off_ptr, stream)
    if st != SNAPPY_OK:
        raise SnappyError(st, "pim_snappy_gather_shards")


def decompress_device(d_comp_ptr: int, d_off_ptr: int, nblocks: int, block_size: int, out_len: int, d_out_ptr: int,
                      d_status_ptr: int, stream: int = 0) -> None:
    st = lib().pim_snappy_decompress_device(d_comp_ptr, d_off_ptr, nblocks, block_size, out_len, d_out_ptr, d_status_ptr, stream)
    if st != SNAPPY_OK:
        raise SnappyError(st, "pim_snappy_decompress_device")


_synth = None


def synth_lib() -> C.CDLL:
    """libpim_synth.so: the seeded workload generators (no codec; the reference arm of bench.py loads only this)."""
    global _synth
    if _synth is None:
        if not os.path.exists(SYNTH_LIB_PATH):
            raise ImportError(f"{SYNTH_LIB_PATH} is not built; run `make -C pim-compression_b200`")
        S = C.CDLL(SYNTH_LIB_PATH)
        S.pim_snappy_synth_host.restype = C.c_int
        S.pim_snappy_synth_host.argtypes = [C.c_int, C.c_uint64, C.c_void_p, C.c_uint64]
        S.pim_snappy_synth_device.restype = C.c_int
        S.pim_snappy_synth_device.argtypes = [C.c_int, C.c_uint64, C.c_void_p, C.c_uint64, C.c_void_p]
        _synth = S
    return _synth


def synth_host(kind: int, seed: int, n: int, out: np.ndarray | None = None) -> np.ndarray:
    buf = out if out is not None else np.empty(n, dtype=np.uint8)
    assert buf.size >= n and buf.dtype == np.uint8
    rc = synth_lib().pim_snappy_synth_host(kind, seed, buf.ctypes.data, n)
    if rc:
        raise ValueError("pim_snappy_synth_host")
    return buf[:n]


def synth_device(kind: int, seed: int, d_ptr: int, n: int, stream: int = 0) -> None:
    rc = synth_lib().pim_snappy_synth_device(kind, seed, d_ptr, n, stream)
    if rc:
        raise ValueError("pim_snappy_synth_device")
