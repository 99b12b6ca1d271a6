#!/usr/bin/env python3
"""Debug build only (make EXTRA=-DPIM_CLOCKS): per-phase cycle counters of the decompress kernel."""
import ctypes as C, os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import torch
import pim_compression_b200 as P
L = P.lib()
kind = {"text": P.SYNTH_TEXT, "mix": P.SYNTH_MIX, "bin": P.SYNTH_BIN, "random": P.SYNTH_RANDOM}[sys.argv[1] if len(sys.argv) > 1 else "text"]
n, bs = 64 << 20, 32768
dev = torch.device("cuda:0"); st = torch.cuda.current_stream().cuda_stream
nb = int(L.pim_snappy_num_blocks(n, bs))
d_in = torch.empty(n, dtype=torch.uint8, device=dev); P.synth_device(kind, 0x640, d_in.data_ptr(), n, st)
scratch = torch.empty(int(L.pim_snappy_compress_scratch_bytes(n, bs)), dtype=torch.uint8, device=dev)
d_comp = torch.empty(int(L.pim_snappy_compress_bound(n, bs)), dtype=torch.uint8, device=dev)
d_off = torch.empty(nb + 1, dtype=torch.int64, device=dev); d_back = torch.zeros(n, dtype=torch.uint8, device=dev)
d_status = torch.zeros(1, dtype=torch.int32, device=dev)
P.compress_device(d_in.data_ptr(), n, bs, scratch.data_ptr(), scratch.numel(), d_comp.data_ptr(), d_off.data_ptr(), st)
buf = (C.c_ulonglong * 32)()
L.pim_snappy_debug_clocks(buf, 1)
e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
e0.record(); P.decompress_device(d_comp.data_ptr(), d_off.data_ptr(), nb, bs, n, d_back.data_ptr(), d_status.data_ptr(), st); e1.record()
torch.cuda.synchronize()
L.pim_snappy_debug_clocks(buf, 0)
v = list(buf)
print(f"kernel {e0.elapsed_time(e1)*1e3:.0f} us, ok={bool(torch.equal(d_back, d_in))}, blocks {nb}")
pw = 2 * nb
print(f"parser (per warp avg cycles): total {v[0]/pw:.0f} slotwait {v[1]/pw:.0f} stage+DP {v[2]/pw:.0f} entrywait {v[3]/pw:.0f} chain {v[4]/pw:.0f} emit {v[5]/pw:.0f} chunks {v[6]/pw:.1f}")
if v[6]: print(f"  per chunk: slot {v[1]/v[6]:.0f} stage+DP {v[2]/v[6]:.0f} entry {v[3]/v[6]:.0f} chain {v[4]/v[6]:.0f} emit {v[5]/v[6]:.0f}")
print(f"copier (per block avg cycles): total {v[8]/nb:.0f} wait {v[9]/nb:.0f} refill {v[10]/nb:.0f} run {v[11]/nb:.0f} iters {v[12]/nb:.1f} trips {v[13]/nb:.1f} elems {v[14]/nb:.0f}")
if v[12]: print(f"  per iteration: refill {v[10]/v[12]:.0f} run {v[11]/v[12]:.0f} cycles; {v[14]/v[12]:.1f} elems, {v[13]/v[12]:.2f} trips")
