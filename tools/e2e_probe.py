#!/usr/bin/env python3
"""Development aid: wall time of snappy_compress_dpu / snappy_decompress_dpu on pinned host buffers for one
stream, per chunk size, with the runtime spans the shim reports (copy_in / run / copy_out)."""
import ctypes as C
import os
import sys
import time

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import numpy as np  # noqa: E402
import torch  # noqa: E402
import pim_compression_b200 as P  # noqa: E402

n, bs = 1 << 30, 32768
L = P.lib()
assert P.set_devices([0]) == 1
dev = torch.device("cuda:0")
t = torch.empty(n, dtype=torch.uint8, device=dev)
P.synth_device(P.SYNTH_MIX, 0x64000000, t.data_ptr(), n, torch.cuda.current_stream().cuda_stream)
h_in = P.PinnedBuffer(n)
torch.from_numpy(h_in.array).copy_(t)
h_comp = P.PinnedBuffer(int(L.pim_snappy_compress_bound(n, bs)) + 64)
h_back = P.PinnedBuffer(n + 4096)
torch.cuda.synchronize()
for mib in [int(x) for x in (sys.argv[1] if len(sys.argv) > 1 else "128,64,32,16").split(",")]:
    L.pim_snappy_set_chunk_bytes(mib << 20)
    res = []
    for rep in range(4):
        inp, out, rt = P.HostBufferContext(), P.HostBufferContext(), P.ProgramRuntime()
        inp.buffer = inp.curr = h_in.ptr
        inp.length, inp.max, out.max = n, 2**63, 2**63
        out.buffer = out.curr = h_comp.ptr
        t0 = time.perf_counter()
        assert L.snappy_compress_dpu(C.byref(inp), C.byref(out), bs, C.byref(rt)) == 0
        t1 = time.perf_counter()
        clen = int(out.length)
        hdr = bytes(h_comp.array[:5])
        k = next(i for i in range(5) if not hdr[i] & 0x80) + 1
        dinp, dout, rt2 = P.HostBufferContext(), P.HostBufferContext(), P.ProgramRuntime()
        dinp.buffer, dinp.curr, dinp.length, dinp.max = h_comp.ptr, h_comp.ptr + k, clen, 2**63
        dout.buffer = dout.curr = h_back.ptr
        dout.length, dout.max = n, 2**63
        t2 = time.perf_counter()
        assert L.snappy_decompress_dpu(C.byref(dinp), C.byref(dout), C.byref(rt2)) == 0
        t3 = time.perf_counter()
        res.append((t1 - t0, t3 - t2, rt.as_dict(), rt2.as_dict()))
    c, d, r1, r2 = min(res[1:], key=lambda x: x[0] + x[1])
    f = lambda r: {k: round(v * 1e3, 2) for k, v in r.items() if v}
    print(f"chunk {mib:4d} MiB: compress {c*1e3:6.2f} ms ({n/c/1e9:5.1f} GB/s) {f(r1)}  decompress {d*1e3:6.2f} ms ({n/d/1e9:5.1f} GB/s) {f(r2)}", flush=True)
assert np.array_equal(h_back.array[:n], h_in.array)
