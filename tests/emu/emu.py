"""TEST INFRASTRUCTURE: builds and loads the lock-step warp emulator (tests/emu/warp_emu.cpp), which runs
pim-compression_b200/csrc/compress_warp.cuh -- the source the CUDA compress kernel includes -- on the CPU."""
import ctypes as C
import os
import subprocess

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))
SO = os.path.join(HERE, "_build", "libwarp_emu.so")
SRC = [os.path.join(HERE, "warp_emu.cpp"),
       os.path.join(ROOT, "pim-compression_b200", "csrc", "compress_warp.cuh"),
       os.path.join(ROOT, "pim-compression_b200", "csrc", "decompress_warp.cuh"),
       os.path.join(ROOT, "pim-compression_b200", "csrc", "warp_prims.cuh")]


def build(defines=()):
    """defines: e.g. ("PIM_DW_RB=2048", "PIM_DW_BPL=4") -- decoder geometry variants get their own library."""
    so = SO if not defines else SO.replace(".so", "_" + "_".join(d.replace("=", "") for d in defines) + ".so")
    if os.path.exists(so) and all(os.path.getmtime(so) >= os.path.getmtime(s) for s in SRC):
        return so
    os.makedirs(os.path.dirname(so), exist_ok=True)
    subprocess.run(["g++", "-O1", "-g", "-shared", "-fPIC", "-std=c++17", "-DPIM_WARP_EMU", *["-D" + d for d in defines], SRC[0], "-o", so],
                   check=True)
    return so


class Emu:
    def __init__(self, defines=()):
        L = C.CDLL(build(tuple(defines)))
        L.emu_decompress.restype = C.c_long
        L.emu_decompress.argtypes = [C.c_void_p, C.c_uint64, C.c_void_p, C.c_uint64, C.c_int]
        L.emu_dw_config.restype = C.c_uint32
        L.emu_compress.restype = C.c_uint64
        L.emu_compress.argtypes = [C.c_void_p, C.c_uint64, C.c_uint32, C.c_void_p]
        L.emu_set_order.argtypes = [C.c_int]
        L.emu_set_global_table.argtypes = [C.c_int]
        self.L = L

    def set_order(self, order: int):
        """Which lane's store lands last when lanes of one instruction hit the same slot (unspecified on the GPU):
        0 = highest lane, 1 = lowest lane, 2 = shuffled."""
        self.L.emu_set_order(order)

    def set_global_table(self, on: bool):
        """Run the parser variant of the pairs whose hash table lives in global memory (parse_block_warp<true>)."""
        self.L.emu_set_global_table(1 if on else 0)

    def compress(self, data, block_size: int) -> np.ndarray:
        d = np.ascontiguousarray(np.frombuffer(data, dtype=np.uint8) if not isinstance(data, np.ndarray) else data)
        nb = (d.size + block_size - 1) // block_size
        out = np.zeros(64 + d.size + d.size // 6 + 40 * nb, dtype=np.uint8)
        n = self.L.emu_compress(d.ctypes.data, d.size, block_size, out.ctypes.data)
        return out[:n].copy()

    def stats(self, reset=True):
        a = (C.c_ulong * 8)()
        self.L.emu_stats(a, 1 if reset else 0)
        return list(a)

    def decompress(self, stream, out_len: int, shift: int = 0):
        """Decode a file image with the warp decoder (decompress_warp.cuh).  Returns the bytes, or the negative
        status (-1 malformed structure, -2 a block failed, -3 wrote outside the output)."""
        d = np.ascontiguousarray(np.frombuffer(stream, dtype=np.uint8) if not isinstance(stream, np.ndarray) else stream)
        out = np.zeros(max(out_len, 1), dtype=np.uint8)
        n = self.L.emu_decompress(d.ctypes.data, d.size, out.ctypes.data, out_len, shift)
        return out[:n].copy() if n >= 0 else int(n)

    def dw_stats(self, reset=True):
        a = (C.c_ulong * 16)()
        self.L.emu_dw_stats(a, 1 if reset else 0)
        return list(a)

    def dw_config(self):
        return tuple(int(self.L.emu_dw_config(i)) for i in range(3))
