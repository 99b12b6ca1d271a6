#!/usr/bin/env python3
"""Device-resident kernel timings per data kind (development aid; the judged numbers come from
bench.py).  For each workload: compress on the GPU, decompress, verify the round trip, and time
both with CUDA events (L2 flushed between repetitions).  Prints one line per workload."""
import argparse
import gzip
import json
import os
import statistics
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import numpy as np  # noqa: E402
import torch  # noqa: E402
import pim_compression_b200 as P  # noqa: E402


def corpus(n):
    parts = []
    for name in ("world192", "plrabn12", "terror2"):
        with gzip.open(os.path.join(ROOT, "tests", "golden", name + ".txt.gz"), "rb") as f:
            parts.append(np.frombuffer(f.read(), dtype=np.uint8))
    p = np.concatenate(parts)
    return np.tile(p, n // p.size + 1)[:n]


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--mib", type=int, default=256)
    ap.add_argument("--reps", type=int, default=5)
    ap.add_argument("--block-size", type=int, default=32768)
    ap.add_argument("--kinds", default="text,random,mix,bin,corpus")
    ap.add_argument("--no-compress-timing", action="store_true")
    a = ap.parse_args()
    L = P.lib()
    dev = torch.device("cuda:0")
    n, bs = a.mib << 20, a.block_size
    st = torch.cuda.current_stream().cuda_stream
    nb = int(L.pim_snappy_num_blocks(n, bs))
    peak = 6454.6
    try:
        peak = json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))["hbm_gbs"]
    except Exception:
        pass
    d_in = torch.empty(n, dtype=torch.uint8, device=dev)
    scratch = torch.empty(int(L.pim_snappy_compress_scratch_bytes(n, bs)), dtype=torch.uint8, device=dev)
    d_comp = torch.empty(int(L.pim_snappy_compress_bound(n, bs)), dtype=torch.uint8, device=dev)
    d_off = torch.empty(nb + 1, dtype=torch.int64, device=dev)
    d_back = torch.zeros(n, dtype=torch.uint8, device=dev)
    d_status = torch.zeros(1, dtype=torch.int32, device=dev)
    flush = torch.empty(256 << 20, dtype=torch.uint8, device=dev)
    kinds = {"text": P.SYNTH_TEXT, "random": P.SYNTH_RANDOM, "mix": P.SYNTH_MIX, "bin": P.SYNTH_BIN}
    for name in a.kinds.split(","):
        if name == "corpus":
            d_in.copy_(torch.from_numpy(corpus(n)))
        else:
            P.synth_device(kinds[name], 0x640, d_in.data_ptr(), n, st)
        tc, td = [], []
        for r in range(a.reps + 1):
            flush.fill_(r & 255)
            e = [torch.cuda.Event(enable_timing=True) for _ in range(4)]
            e[0].record()
            if r == 0 or not a.no_compress_timing:
                P.compress_device(d_in.data_ptr(), n, bs, scratch.data_ptr(), scratch.numel(), d_comp.data_ptr(), d_off.data_ptr(), st)
            e[1].record()
            flush.fill_((r + 1) & 255)
            d_back.zero_()
            e[2].record()
            P.decompress_device(d_comp.data_ptr(), d_off.data_ptr(), nb, bs, n, d_back.data_ptr(), d_status.data_ptr(), st)
            e[3].record()
            torch.cuda.synchronize()
            if r >= 1:
                tc.append(e[0].elapsed_time(e[1]))
                td.append(e[2].elapsed_time(e[3]))
        ok = int(d_status.item()) == 0 and bool(torch.equal(d_back, d_in))
        C = int(d_off[-1].item())
        mc, md = statistics.median(tc), statistics.median(td)
        print(json.dumps({"kind": name, "mib": a.mib, "bs": bs, "ok": ok, "C_over_U": round(C / n, 4),
                          "comp_ms": round(mc, 3), "comp_GBps": round(n / mc / 1e6, 1),
                          "decomp_ms": round(md, 3), "decomp_GBps": round(n / md / 1e6, 1),
                          "decomp_hbm_frac": round((n + C) / md / 1e6 / peak, 4),
                          "comp_hbm_frac": round((n + C) / mc / 1e6 / peak, 4)}), flush=True)
        if not ok:
            bad = torch.nonzero(d_back != d_in)[:1]
            print("  ROUND TRIP MISMATCH status", int(d_status.item()), "first diff", bad.flatten().tolist(), flush=True)


if __name__ == "__main__":
    main()
