#!/usr/bin/env python3
"""Development aid: where a parser's pass spends its cycles (build with EXTRA=-DPIM_CW_PROFILE; shared-memory pairs only).
Usage: PIM_SNAPPY_NSM=.. PIM_SNAPPY_NGT=.. python tools/cw_prof.py [kind]"""
import ctypes as C
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import torch  # noqa: E402
import pim_compression_b200 as P  # noqa: E402

L = P.lib()
kind = {"text": P.SYNTH_TEXT, "mix": P.SYNTH_MIX, "bin": P.SYNTH_BIN}[sys.argv[1] if len(sys.argv) > 1 else "text"]
n, bs = 256 << 20, 32768
dev = torch.device("cuda:0")
st = torch.cuda.current_stream().cuda_stream
nb = int(L.pim_snappy_num_blocks(n, bs))
d_in = torch.empty(n, dtype=torch.uint8, device=dev)
P.synth_device(kind, 0x640, d_in.data_ptr(), n, st)
scratch = torch.empty(int(L.pim_snappy_compress_scratch_bytes(n, bs)), dtype=torch.uint8, device=dev)
d_comp = torch.empty(int(L.pim_snappy_compress_bound(n, bs)), dtype=torch.uint8, device=dev)
d_off = torch.empty(nb + 1, dtype=torch.int64, device=dev)
out = (C.c_ulonglong * 16)()
for rep in range(2):
    torch.cuda.synchronize()
    L.pim_snappy_debug_cw_prof(out, 1)
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    P.compress_device(d_in.data_ptr(), n, bs, scratch.data_ptr(), scratch.numel(), d_comp.data_ptr(), d_off.data_ptr(), st)
    e1.record()
    torch.cuda.synchronize()
    L.pim_snappy_debug_cw_prof(out, 0)
names = ["table entry arrived (LDS)", "insert + read-back, candidate bytes arrived, match lengths", "shared slots resolved", "pair walk + forwarded-candidate check",
         "table updates + pass end position", "literal bounds, long match, record to the emitter", "own bytes arrived, hash", "loop control, prefetch, lane layout"]
passes = out[8]
tot = sum(out[i] for i in range(8))
print(f"env NSM={os.environ.get('PIM_SNAPPY_NSM')} NGT={os.environ.get('PIM_SNAPPY_NGT')}: {n / (e0.elapsed_time(e1) * 1e-3) / 1e9:.1f} GB/s, "
      f"{passes} passes of shared-memory pairs, {tot / max(passes, 1):.0f} cycles per pass")
for i in range(8):
    if out[i]:
        print(f"  {out[i] / max(passes, 1):7.0f} cycles {100 * out[i] / tot:5.1f}%  {names[i]}")
