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
       os.path.join(ROOT, "pim-compression_b200", "csrc", "warp_prims.cuh")]


def build():
    """$PIM_EMU_DEFINES (e.g. "PIM_CW_WINDOW=1 PIM_CW_CHAIN_MAX=24"): build variants of the kernel source get their own library."""
    global SO
    defines = os.environ.get("PIM_EMU_DEFINES", "").split()
    if defines:
        SO = os.path.join(HERE, "_build", "libwarp_emu_" + "_".join(d.replace("=", "") for d in defines) + ".so")
    if os.path.exists(SO) and all(os.path.getmtime(SO) >= os.path.getmtime(s) for s in SRC):
        return
    os.makedirs(os.path.dirname(SO), exist_ok=True)
    subprocess.run(["g++", "-O1", "-g", "-shared", "-fPIC", "-std=c++17", "-DPIM_WARP_EMU", *["-D" + d for d in defines], SRC[0], "-o", SO], check=True)


class Emu:
    def __init__(self):
        build()
        L = C.CDLL(SO)
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
