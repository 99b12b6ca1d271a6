#!/usr/bin/env python3
"""Small round trips through the C ABI for `compute-sanitizer --tool memcheck|racecheck` (tools/gpu_sanitize.sh):
ragged sizes, block sizes 15 .. 65536, streamed and one-piece host paths, the stream's last block (tail copy),
periodic runs / zero pages, and a malformed stream.  With PIM_SNAPPY_NSM=1 in the environment the compressor runs one
shared-memory-table pair and two global-memory-table pairs per CTA, so both table placements are covered.
Exits non-zero on a mismatch."""
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import numpy as np  # noqa: E402
import pim_compression_b200 as P  # noqa: E402

SMALL = os.environ.get("SANITIZE_SMALL", "0") == "1"  # racecheck is slow: fewer / smaller cases
assert P.set_devices([0]) == 1
bad = 0
rng = np.random.default_rng(7)
cases = [("mix", P.synth_host(P.SYNTH_MIX, 0x5A19, (1 << 20) + 4321 if SMALL else (2 << 20) + 4321)),
         ("bin", P.synth_host(P.SYNTH_BIN, 0x5A1A, 300000)),
         ("text", P.synth_host(P.SYNTH_TEXT, 0x5A17, 70001)),
         ("random", P.synth_host(P.SYNTH_RANDOM, 0x5A18, 40000)),
         ("zeros", np.zeros(100000, dtype=np.uint8)),
         ("period7", np.tile(np.frombuffer(b"abcdefg", dtype=np.uint8), 9000)),
         ("period300", np.tile(rng.integers(0, 256, 300, dtype=np.uint8), 200))]
for name, data in cases:
    n = data.size
    plans = [(32768, 0), (4096, 1 << 19), (65536, 0), (1000, 1 << 18), (15, 0) if n < 100000 else (20000, 1 << 19)]
    if SMALL:
        plans = plans[:3]
    for bs, chunk in plans:
        P.set_chunk_bytes(chunk)
        comp = P.compress(data, bs)
        back = P.decompress(comp)
        ok = back.size == data.size and np.array_equal(back, data)
        print(f"{name} n {n} bs {bs} chunk {chunk}: C/U {comp.size / n:.3f} {'ok' if ok else 'MISMATCH'}", flush=True)
        bad += not ok
# malformed: a copy4 tag with a wild offset at the start of a block; a body that ends early
P.set_chunk_bytes(0)
good = P.compress(cases[2][1], 4096)
for what, s in (("wild copy", np.concatenate([good[:9], np.frombuffer(b"\xff\xff\xff\xff\xff", dtype=np.uint8), good[14:]])), ("truncated", good[: good.size // 2])):
    try:
        P.decompress(s)
        print(f"malformed ({what}): accepted", flush=True)
        bad += 1
    except P.SnappyError as e:
        print(f"malformed ({what}): rejected with status {e.status}", flush=True)
back = P.decompress(good)
bad += not np.array_equal(back, cases[2][1])
print("memcheck_small:", "OK" if bad == 0 else f"{bad} problems", flush=True)
sys.exit(1 if bad else 0)
