#!/usr/bin/env python3
"""Quick on-GPU diagnostics: compress and decompress checked separately against the oracle,
with the first differing block / byte printed.  Development aid, not a test."""
import os
import sys
import time

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))
import pim_compression_b200 as P  # noqa: E402
from oracle.oracle import Oracle  # noqa: E402
from conftest import edge_inputs  # noqa: E402


def first_diff(a, b):
    m = min(a.size, b.size)
    d = np.flatnonzero(a[:m] != b[:m])
    return int(d[0]) if d.size else (m if a.size != b.size else -1)


def main():
    o = Oracle()
    print("devices:", P.set_devices([0]), flush=True)
    cases = dict(edge_inputs())
    cases["mix3m"] = P.synth_host(P.SYNTH_MIX, 1, 3 << 20)
    cases["text1m"] = P.synth_host(P.SYNTH_TEXT, 2, 1 << 20)
    cases["bin1m"] = P.synth_host(P.SYNTH_BIN, 3, 1 << 20)
    bad = 0
    for name, data in cases.items():
        for bs in (32768, 1000, 65536):
            want = o.compress(data, bs)
            hlen, dlen, _, off = P.block_offsets(want)
            try:
                t0 = time.time()
                back = P.decompress(want)
                fd = first_diff(back, data)
                if fd >= 0:
                    bad += 1
                    print(f"DECOMP MISMATCH {name} bs={bs}: first diff at {fd} (block {fd // bs}, in-block {fd % bs}) sizes {back.size}/{data.size}", flush=True)
            except Exception as e:
                bad += 1
                print(f"DECOMP ERROR {name} bs={bs}: {e}", flush=True)
            try:
                got = P.compress(data, bs)
                fd = first_diff(got, want)
                if fd >= 0:
                    bad += 1
                    blk = int(np.searchsorted(off, max(fd - hlen, 0), side="right") - 1)
                    print(f"COMP MISMATCH {name} bs={bs}: first diff at {fd} (block {blk}, at +{fd - hlen - int(off[blk])}) sizes {got.size}/{want.size}", flush=True)
            except Exception as e:
                bad += 1
                print(f"COMP ERROR {name} bs={bs}: {e}", flush=True)
    print("sanity:", "OK" if bad == 0 else f"{bad} problems", flush=True)
    return 1 if bad else 0


if __name__ == "__main__":
    sys.exit(main())
