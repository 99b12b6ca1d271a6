#!/usr/bin/env python3
"""Small round trips through the C ABI for `compute-sanitizer --tool memcheck` (development aid): ragged sizes,
several block sizes, streamed and one-piece host paths.  Exits non-zero on a mismatch."""
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import numpy as np  # noqa: E402
import pim_compression_b200 as P  # noqa: E402

assert P.set_devices([0]) == 1
bad = 0
for kind, n in ((P.SYNTH_MIX, (2 << 20) + 4321), (P.SYNTH_BIN, 300000), (P.SYNTH_TEXT, 70001), (P.SYNTH_RANDOM, 40000)):
    data = P.synth_host(kind, 0x5A17 + kind, n)
    for bs, chunk in ((32768, 0), (4096, 1 << 19), (65536, 0), (1000, 1 << 18), (15, 0) if n < 100000 else (20000, 1 << 19)):
        P.set_chunk_bytes(chunk)
        comp = P.compress(data, bs)
        back = P.decompress(comp)
        ok = back.size == data.size and np.array_equal(back, data)
        print(f"kind {kind} n {n} bs {bs} chunk {chunk}: C/U {comp.size / n:.3f} {'ok' if ok else 'MISMATCH'}", flush=True)
        bad += not ok
sys.exit(1 if bad else 0)
