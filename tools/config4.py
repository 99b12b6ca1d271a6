#!/usr/bin/env python3
"""BASELINE.json config 4: synthetic 64 MiB mixed input (SYN-MIX-64M, 2048 blocks @32 KB),
device-resident decompress (and compress), >= 20 repetitions with an L2 flush in between.
Also the short command used for ncu captures (--reps 3).  Prints one JSON line."""
import argparse
import json
import os
import statistics
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import torch  # noqa: E402
import pim_compression_b200 as P  # noqa: E402


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--mib", type=int, default=64)
    ap.add_argument("--reps", type=int, default=20)
    ap.add_argument("--kind", type=int, default=P.SYNTH_MIX)
    ap.add_argument("--block-size", type=int, default=32768)
    ap.add_argument("--seed", type=lambda x: int(x, 0), default=0x640)
    a = ap.parse_args()
    L = P.lib()
    dev = torch.device("cuda:0")
    n, bs = a.mib << 20, a.block_size
    st = torch.cuda.current_stream().cuda_stream
    nb = int(L.pim_snappy_num_blocks(n, bs))
    d_in = torch.empty(n, dtype=torch.uint8, device=dev)
    P.synth_device(a.kind, a.seed, d_in.data_ptr(), n, st)
    scratch = torch.empty(int(L.pim_snappy_compress_scratch_bytes(n, bs)), dtype=torch.uint8, device=dev)
    d_comp = torch.empty(int(L.pim_snappy_compress_bound(n, bs)), dtype=torch.uint8, device=dev)
    d_off = torch.empty(nb + 1, dtype=torch.int64, device=dev)
    d_back = torch.zeros(n, dtype=torch.uint8, device=dev)
    d_status = torch.zeros(1, dtype=torch.int32, device=dev)
    flush = torch.empty(256 << 20, dtype=torch.uint8, device=dev)  # > 126 MB L2
    tc, td = [], []
    for r in range(a.reps + 3):
        flush.fill_(r & 255)
        e = [torch.cuda.Event(enable_timing=True) for _ in range(4)]
        e[0].record()
        P.compress_device(d_in.data_ptr(), n, bs, scratch.data_ptr(), scratch.numel(), d_comp.data_ptr(), d_off.data_ptr(), st)
        e[1].record()
        flush.fill_((r + 1) & 255)
        e[2].record()
        P.decompress_device(d_comp.data_ptr(), d_off.data_ptr(), nb, bs, n, d_back.data_ptr(), d_status.data_ptr(), st)
        e[3].record()
        torch.cuda.synchronize()
        if r >= 3:
            tc.append(e[0].elapsed_time(e[1]))
            td.append(e[2].elapsed_time(e[3]))
    assert int(d_status.item()) == 0 and torch.equal(d_back, d_in)
    C = int(d_off[-1].item())
    peak = 6544.7
    try:
        peak = json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))["hbm_gbs"]
    except Exception:
        pass

    def row(ts):
        med, best = statistics.median(ts), min(ts)
        return {"median_ms": med, "best_ms": best, "uncompressed_gbps_median": n / med / 1e6, "uncompressed_gbps_best": n / best / 1e6,
                "hbm_frac_median": (n + C) / med / 1e6 / peak, "hbm_frac_best": (n + C) / best / 1e6 / peak}

    print(json.dumps({"workload": f"kind={a.kind} {a.mib} MiB, {nb} blocks @ {bs}", "U": n, "C": C, "C_over_U": C / n, "reps": a.reps,
                      "peak_gbps": peak, "compress": row(tc), "decompress": row(td)}))


if __name__ == "__main__":
    main()
