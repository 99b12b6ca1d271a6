#!/usr/bin/env python3
"""BASELINE.json configs 1 and 2 as measurements (run on the GPU box; prints one JSON object per config).

config 1  host CPU snappy decompress of test/terror2.snappy (105 KB) by the reference itself
          (oracle/_ref/dpu_snappy_ref, "Host time", dpu_snappy.c:195-202), verified against terror2.txt --
          and the same file through the GPU path (CLI and C ABI) with its latency beside it.
config 2  dickens.txt is absent from the reference checkout: stand-ins SYN-TEXT-10M (10,192,446 B, seed 0xD1C0)
          and CORPUS-10M (world192 + plrabn12 + terror2 + xml tiled), compress + decompress round trip at
          32 KB blocks on one B200, bit-exact against the oracle, with host-API and kernel times.
"""
import gzip
import json
import os
import re
import subprocess
import sys
import tempfile
import time

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import numpy as np  # noqa: E402
import torch  # noqa: E402
import pim_compression_b200 as P  # noqa: E402
from oracle.oracle import Oracle, REF_BIN  # noqa: E402  (the checker, and the reference binary for config 1)

GOLD = os.path.join(ROOT, "tests", "golden")
CLI = os.path.join(ROOT, "pim-compression_b200", "dpu_snappy")


def gz(name):
    return np.frombuffer(gzip.open(os.path.join(GOLD, name + ".gz"), "rb").read(), dtype=np.uint8)


def cli_times(out):
    t = {}
    for key in ("Pre-processing time", "Host time", "Alloc time", "Load time", "Copy in time", "Copy out time", "Free time"):
        m = re.search(rf"{key}:\s*([0-9.eE+-]+)", out)
        if m:
            t[key] = float(m.group(1))
    return t


def config1(tmp):
    sn, txt = gz("terror2.snappy"), gz("terror2.txt")
    fsn = os.path.join(tmp, "terror2.snappy")
    sn.tofile(fsn)
    res = {"config": "host CPU snappy decompress of test/terror2.snappy (52,525 B -> 105,438 B, 4 blocks), verified against terror2.txt"}
    if os.path.exists(REF_BIN):
        host = []
        for _ in range(7):
            out = subprocess.run([REF_BIN, "-i", fsn, "-o", os.path.join(tmp, "ref.out")], capture_output=True, text=True).stdout
            host.append(cli_times(out).get("Host time"))
        ok = np.array_equal(np.fromfile(os.path.join(tmp, "ref.out"), dtype=np.uint8), txt)
        host = [h for h in host if h is not None]
        res["reference_host"] = {"binary": "oracle/_ref/dpu_snappy_ref (the reference's dpu_snappy.c + snappy_decompress.c, unmodified)",
                                 "host_time_s_best": min(host), "host_time_s_median": sorted(host)[len(host) // 2], "runs": len(host),
                                 "matches_terror2_txt": bool(ok), "uncompressed_MBps_best": txt.size / min(host) / 1e6}
    # GPU path: the CLI (process start + context creation included in wall time; its own breakdown beside it) ...
    t0 = time.perf_counter()
    out = subprocess.run([CLI, "-d", "-i", fsn, "-o", os.path.join(tmp, "gpu.out")], capture_output=True, text=True).stdout
    wall = time.perf_counter() - t0
    ok = np.array_equal(np.fromfile(os.path.join(tmp, "gpu.out"), dtype=np.uint8), txt)
    res["gpu_cli"] = {"command": "dpu_snappy -d -i terror2.snappy", "wall_s": wall, "breakdown_s": cli_times(out), "matches_terror2_txt": bool(ok)}
    # ... and the C ABI in a warm process (setup_decompression + snappy_decompress_dpu, host buffers)
    lat = []
    for _ in range(20):
        t0 = time.perf_counter()
        back = P.decompress(sn)
        lat.append(time.perf_counter() - t0)
    res["gpu_c_abi_warm"] = {"latency_s_best": min(lat), "latency_s_median": sorted(lat)[10], "matches_terror2_txt": bool(np.array_equal(back, txt)),
                             "note": "4 blocks occupy 4 warps of one B200: this is launch + copy latency, not throughput"}
    return res


def config2():
    o = Oracle()
    n = 10192446
    parts = [gz(f + ".txt") for f in ("world192", "plrabn12", "terror2")] + [P.decompress(gz("xml.snappy"))]
    corpus = np.concatenate(parts)
    inputs = {"SYN-TEXT-10M (seed 0xD1C0)": P.synth_host(P.SYNTH_TEXT, 0xD1C0, n), "CORPUS-10M": np.tile(corpus, n // corpus.size + 1)[:n].copy()}
    res = {"config": "dickens.txt stand-ins, 10,192,446 B, 312 blocks @ 32 KB, 1 B200, compress + decompress round trip", "inputs": {}}
    L = P.lib()
    dev = torch.device("cuda:0")
    st = torch.cuda.current_stream().cuda_stream
    bs = 32768
    nb = int(L.pim_snappy_num_blocks(n, bs))
    for name, data in inputs.items():
        want = o.compress(data, bs)
        tc, td = [], []
        for _ in range(7):
            t0 = time.perf_counter()
            got = P.compress(data, bs)
            t1 = time.perf_counter()
            back = P.decompress(got)
            t2 = time.perf_counter()
            tc.append(t1 - t0)
            td.append(t2 - t1)
        row = {"compressed_bytes": int(got.size), "compress_equals_oracle": bool(got.size == want.size and np.array_equal(got, want)),
               "round_trip_equals_input": bool(np.array_equal(back, data)),
               "host_api_s": {"compress_best": min(tc), "decompress_best": min(td), "round_trip_best": min(a + b for a, b in zip(tc, td))}}
        # kernels alone, device resident
        d_in = torch.from_numpy(data).to(dev)
        scratch = torch.empty(int(L.pim_snappy_compress_scratch_bytes(n, bs)), dtype=torch.uint8, device=dev)
        d_comp = torch.empty(int(L.pim_snappy_compress_bound(n, bs)), dtype=torch.uint8, device=dev)
        d_off = torch.empty(nb + 1, dtype=torch.int64, device=dev)
        d_back = torch.zeros(n, dtype=torch.uint8, device=dev)
        d_status = torch.zeros(1, dtype=torch.int32, device=dev)
        kc, kd = [], []
        for _ in range(8):
            e = [torch.cuda.Event(enable_timing=True) for _ in range(3)]
            e[0].record()
            P.compress_device(d_in.data_ptr(), n, bs, scratch.data_ptr(), scratch.numel(), d_comp.data_ptr(), d_off.data_ptr(), st)
            e[1].record()
            P.decompress_device(d_comp.data_ptr(), d_off.data_ptr(), nb, bs, n, d_back.data_ptr(), d_status.data_ptr(), st)
            e[2].record()
            torch.cuda.synchronize()
            kc.append(e[0].elapsed_time(e[1]))
            kd.append(e[1].elapsed_time(e[2]))
        assert int(d_status.item()) == 0 and torch.equal(d_back, d_in)
        row["device_resident_ms"] = {"compress_best": min(kc[1:]), "decompress_best": min(kd[1:]),
                                     "note": "312 blocks on 2,220 compress pair slots / 3,552 decoder warps: one block's latency, not throughput"}
        res["inputs"][name] = row
    return res


def main():
    assert P.set_devices([0]) == 1
    with tempfile.TemporaryDirectory() as tmp:
        print(json.dumps({"config1": config1(tmp)}), flush=True)
    print(json.dumps({"config2": config2()}), flush=True)


if __name__ == "__main__":
    main()
