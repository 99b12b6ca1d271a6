#!/usr/bin/env python3
"""bench.py -- headline benchmark: block-parallel Snappy (PIM-compression multi-block format).

Workload (BASELINE.json configs[4], the configuration the metric is quoted on): a synthetic
64 GiB corpus of 64 independent 1 GiB SYN-MIX streams (1 MiB text-like / 1 MiB random
alternating, 32 KB blocks) sharded over 8 B200 -- i.e. 8 streams (8 GiB) per GPU.  Weak scaling:
every rank always works on 8 GiB (fewer only if --streams-per-gpu says so or memory forces it;
the JSON line states what ran).

One step = compress + decompress of every stream of the rank.
  value  device-resident: inputs, offsets and outputs in HBM, kernels only (CUDA events)
  e2e    the same through the reference-facing C ABI (snappy_compress_dpu/snappy_decompress_dpu)
         on pinned HOST buffers, H2D and D2H inside the timed region
  --impl reference   the reference's own host codec (oracle/_ref, else the oracle port) on the
         box's host cores, all threads, on a bounded sample of the same streams.
"""
from __future__ import annotations

import argparse
import json
import os
import statistics
import subprocess
import sys
import threading
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)

GIB = 1 << 30
SEED_BASE = 0x64000000
METRIC = "uncompressed GB/s, compress + decompress round trip (block-parallel Snappy)"


def parse_args():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=3)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--streams-per-gpu", type=int, default=8)
    ap.add_argument("--stream-bytes", type=int, default=GIB)
    ap.add_argument("--block-size", type=int, default=32768)
    ap.add_argument("--cpu-sample-mib", type=int, default=1024, help="cpu_baseline sample size (1 thread)")
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--no-e2e", action="store_true")
    ap.add_argument("--no-by-input", dest="by_input", action="store_false", help="skip the text / corpus / binary one-launch measurements")
    return ap.parse_args()


def peaks():
    try:
        with open(os.path.join(ROOT, "MEASURED_PEAKS.json")) as f:
            return float(json.load(f)["hbm_gbs"]), "measured (MEASURED_PEAKS.json hbm_gbs)"
    except Exception:
        return 6650.0, "fallback (B200_PROFILING.md 6.65 TB/s)"


def workload_name(args, n_gpus):
    return (f"SYN-MIX corpus, {n_gpus * args.streams_per_gpu} streams x {args.stream_bytes / GIB:g} GiB "
            f"({n_gpus * args.streams_per_gpu * args.stream_bytes / GIB:g} GiB total, {args.streams_per_gpu} streams per GPU), "
            f"{args.block_size} B blocks, contiguous block ranges per GPU")


# ----------------------------------------------------------------------------- clocks
class ClockSampler:
    """nvidia-smi clocks / throttle reasons sampled DURING the timed region."""
    Q = ("clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.active,clocks_event_reasons.hw_slowdown,"
         "clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap")

    def __init__(self, index):
        self.index, self.proc, self.lines = index, None, []

    def start(self):
        try:
            self.proc = subprocess.Popen(["nvidia-smi", f"--query-gpu={self.Q}", "--format=csv,noheader,nounits", "-lms", "100",
                                          "-i", str(self.index)], stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
            self.t = threading.Thread(target=lambda: self.lines.extend(self.proc.stdout), daemon=True)
            self.t.start()
        except Exception:
            self.proc = None

    def stop(self):
        if not self.proc:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        time.sleep(0.15)
        self.proc.terminate()
        self.t.join(timeout=2)
        sm, mx, reasons = [], [], set()
        for ln in self.lines:
            f = [x.strip() for x in ln.split(",")]
            if len(f) < 8:
                continue
            try:
                sm.append(float(f[0])); mx.append(float(f[1]))
            except ValueError:
                continue
            for name, v in zip(("hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"), f[4:8]):
                if v.lower().startswith("active"):
                    reasons.add(name)
        # samples under load = the upper half of the clock readings
        return {"sm_mhz": statistics.median(sm) if sm else None, "sm_max_mhz": max(mx) if mx else None,
                "reasons": sorted(reasons), "samples": len(sm)}


# ----------------------------------------------------------------------------- CPU legs
def cpu_codec():
    from oracle.oracle import Oracle, Reference
    if Reference.available():
        return Reference(), "reference"
    return Oracle(), "port"


def cpu_baseline_1thread(P, args):
    """Reference host codec, as-is (single thread), on a bounded sample of stream 0."""
    codec, kind = cpu_codec()
    n = min(args.cpu_sample_mib << 20, args.stream_bytes)
    data = P.synth_host(P.SYNTH_MIX, SEED_BASE, n)
    t0 = time.perf_counter()
    comp = codec.compress(data, args.block_size)
    t1 = time.perf_counter()
    rc, back = codec.decompress(comp)
    t2 = time.perf_counter()
    assert rc == 0 and back.size == n
    return {"value": n / (t2 - t0) / 1e9, "unit": "GB/s", "cores": 1, "kind": kind,
            "sample": f"first {n >> 20} MiB of stream 0 (SYN-MIX), compress {n / (t1 - t0) / 1e9:.3f} GB/s + decompress {n / (t2 - t1) / 1e9:.3f} GB/s, 1 thread",
            "compress_gbps": n / (t1 - t0) / 1e9, "decompress_gbps": n / (t2 - t1) / 1e9}


def host_cores():
    """(threads this process may run on, distinct physical cores among them) -- the second from /proc/cpuinfo."""
    aff = sorted(os.sched_getaffinity(0))
    phys = set()
    try:
        cur = {}
        for ln in open("/proc/cpuinfo"):
            if ":" in ln:
                k, v = [x.strip() for x in ln.split(":", 1)]
                cur[k] = v
            elif cur:
                if int(cur.get("processor", -1)) in aff:
                    phys.add((cur.get("physical id", "0"), cur.get("core id", cur.get("processor"))))
                cur = {}
    except Exception:
        pass
    return len(aff), (len(phys) if phys else None)


def run_reference_arm(args):
    """--impl reference: the reference's host codec on all host threads (rank 0 only).  Loads oracle/_ref and the
    workload generator library only -- no codec of this repository is in the process."""
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return 0
    import pim_compression_b200 as P
    from concurrent.futures import ThreadPoolExecutor
    codec, kind = cpu_codec()
    cores, phys = host_cores()
    slice_bytes = 32 << 20  # thread t codes the first 32 MiB of stream t of the corpus, every step
    slices = [P.synth_host(P.SYNTH_MIX, SEED_BASE + t, slice_bytes) for t in range(cores)]

    def work(d):
        c = codec.compress(d, args.block_size)
        rc, b = codec.decompress(c)
        assert rc == 0 and b.size == d.size
        return d.size

    times = []
    with ThreadPoolExecutor(cores) as ex:
        for it in range(args.warmup + args.steps):
            t0 = time.perf_counter()
            total = sum(ex.map(work, slices))
            dt = time.perf_counter() - t0
            if it >= args.warmup:
                times.append(dt)
    t = sum(times) / len(times)
    v = total / t / 1e9
    sample = (f"{cores} threads ({phys if phys else '?'} physical cores; the reference codec is single-threaded, one independent slice per thread) x "
              f"the first {slice_bytes >> 20} MiB of streams 0..{cores - 1} of the corpus per step, compress + decompress each (ctypes releases the GIL)")
    line = {"impl": "reference", "metric": METRIC, "value": v, "unit": "GB/s", "n_gpus": args.gpus, "steps": args.steps,
            "warmup": args.warmup, "ms_per_step": t * 1e3, "higher_is_better": True, "scaling": "weak", "vs_baseline": None,
            "dtype": "u8", "data": "synthetic",
            "config": {"workload": workload_name(args, args.gpus), "cpu_sample_bytes_per_step": total, "threads": cores, "physical_cores": phys},
            "cpu_baseline": {"value": v, "unit": "GB/s", "cores": cores, "kind": kind, "sample": sample},
            "e2e": {"value": v, "unit": "GB/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0}, "gpu_launches": 0}
    print(json.dumps(line))
    return 0


# ----------------------------------------------------------------------------- GPU arm
def mem_available_bytes():
    try:
        for ln in open("/proc/meminfo"):
            if ln.startswith("MemAvailable:"):
                return int(ln.split()[1]) * 1024
    except Exception:
        pass
    return 0


def run_ours(args):
    import ctypes as C
    import numpy as np
    import torch
    import pim_compression_b200 as P

    rank = int(os.environ.get("RANK", "0"))
    local_rank = int(os.environ.get("LOCAL_RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    if not torch.cuda.is_available():
        raise SystemExit("bench.py: no CUDA device; the product has no CPU path (use --impl reference for the CPU arm)")
    torch.cuda.set_device(local_rank)
    dev = torch.device("cuda", local_rank)
    # Library banners (NCCL prints its version on stdout at communicator creation) go to stderr: file descriptor 1 points at
    # stderr until the JSON line is printed, so that stdout carries exactly that one line.
    sys.stdout.flush()
    saved_stdout = os.dup(1)
    os.dup2(2, 1)
    dist = None
    if world > 1:
        import torch.distributed as dist
        dist.init_process_group("nccl", device_id=dev)
    L = P.lib()
    assert P.set_devices([local_rank]) == 1, "no usable sm_100 GPU"

    S, n, bs = args.streams_per_gpu, args.stream_bytes, args.block_size
    nb = int(L.pim_snappy_num_blocks(n, bs))
    st = torch.cuda.current_stream().cuda_stream

    # ---- resident inputs: stream s of this rank is global stream rank*S+s of the corpus
    d_in = []
    for s in range(S):
        t = torch.empty(n, dtype=torch.uint8, device=dev)
        P.synth_device(P.SYNTH_MIX, SEED_BASE + rank * S + s, t.data_ptr(), n, st)
        d_in.append(t)
    # one set of work buffers per stream: the S streams of a step go through ONE compress launch and ONE decompress
    # launch (pim_snappy_compress_batch_device / pim_snappy_decompress_batch_device)
    scratch = [torch.empty(int(L.pim_snappy_compress_scratch_bytes(n, bs)), dtype=torch.uint8, device=dev) for _ in range(S)]
    d_comp = [torch.empty(int(L.pim_snappy_compress_bound(n, bs)), dtype=torch.uint8, device=dev) for _ in range(S)]
    d_off = [torch.empty(nb + 1, dtype=torch.int64, device=dev) for _ in range(S)]
    d_back = [torch.empty(n, dtype=torch.uint8, device=dev) for _ in range(S)]
    d_status = torch.zeros(S, dtype=torch.int32, device=dev)
    torch.cuda.synchronize()
    ptrs = lambda ts: [t.data_ptr() for t in ts]

    def device_step():
        P.compress_batch_device(ptrs(d_in), [n] * S, bs, ptrs(scratch), [t.numel() for t in scratch], ptrs(d_comp), ptrs(d_off), st)
        P.decompress_batch_device(ptrs(d_comp), ptrs(d_off), [nb] * S, bs, [n] * S, ptrs(d_back), d_status.data_ptr(), st)

    def sync_all():
        torch.cuda.synchronize()
        if dist is not None:
            dist.barrier()
            torch.cuda.synchronize()

    def max_over_ranks(x):
        if dist is None:
            return x
        t = torch.tensor([x], dtype=torch.float64, device=dev)
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        return float(t.item())

    def sum_over_ranks(x):
        if dist is None:
            return x
        t = torch.tensor([x], dtype=torch.float64, device=dev)
        dist.all_reduce(t, op=dist.ReduceOp.SUM)
        return float(t.item())

    # ---- warm-up + one correctness check (round trip of the last stream; compressed sizes)
    comp_bytes = 0
    for w in range(max(args.warmup, 1)):
        if w == 0:
            device_step()
            torch.cuda.synchronize()
            assert int(d_status.abs().sum().item()) == 0, "malformed stream reported"
            for s in range(S):
                comp_bytes += int(d_off[s][-1].item())
                assert torch.equal(d_back[s], d_in[s]), f"round trip failed on stream {s}"
                d_back[s].zero_()
        else:
            device_step()
    U = S * n

    # ---- timed region, device resident
    sampler = ClockSampler(torch.cuda.current_device() if "CUDA_VISIBLE_DEVICES" not in os.environ else local_rank)
    sync_all()
    P.profile(True)
    P.profile_read()
    launches0 = P.launch_count()
    sampler.start()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(args.steps):
        device_step()
    e1.record()
    sync_all()
    dev_ms = max_over_ranks(e0.elapsed_time(e1)) / args.steps
    clocks = sampler.stop()
    # what the LAST timed step left behind is checked too (the warm-up check zeroed d_back): every stream's status word,
    # compressed size and round trip
    assert int(d_status.abs().sum().item()) == 0, "malformed stream reported in the timed steps"
    assert sum(int(d_off[s][-1].item()) for s in range(S)) == comp_bytes, "compressed size changed in the timed steps"
    for s in range(S):
        assert torch.equal(d_back[s], d_in[s]), f"round trip of the timed steps failed on stream {s}"
    launches = P.launch_count() - launches0
    prof = P.profile_read()
    P.profile(False)

    total_U = sum_over_ranks(U)
    value = total_U / (dev_ms * 1e-3) / 1e9
    peak, peak_src = peaks()
    kern = {}
    for k in ("compress", "decompress"):
        ms = prof[k]["ms"] / max(prof[k]["launches"], 1)  # per launch = the S streams of one step
        alg = S * n + comp_bytes                            # U + C per launch
        kern[k] = {"ms_per_launch": ms, "streams_per_launch": S, "uncompressed_gbps": S * n / (ms * 1e-3) / 1e9, "achieved_gbps": alg / (ms * 1e-3) / 1e9,
                   "frac_of_hbm_peak": alg / (ms * 1e-3) / 1e9 / peak, "launches": prof[k]["launches"]}
    dom = max(("compress", "decompress"), key=lambda k: prof[k]["ms"])
    traffic = None  # DRAM bytes per launch, scaled from the committed ncu --set full capture (profiles/traffic.json)
    try:
        with open(os.path.join(ROOT, "profiles", "traffic.json")) as f:
            tj = json.load(f)
        if args.block_size == 32768:
            tk = tj[f"snappy_{dom}_kernel"]
            traffic = (tk["dram_read"] + tk["dram_write"]) * S * n / tj["capture_input_bytes"]
    except Exception:
        traffic = None
    roofline = {"kernel": f"snappy_{dom}_kernel", "bound": "hbm", "achieved": kern[dom]["achieved_gbps"], "peak": peak,
                "unit": "GB/s", "frac": kern[dom]["frac_of_hbm_peak"], "traffic": traffic, "peak_source": peak_src,
                "algorithmic_bytes_per_launch": S * n + comp_bytes,
                "share_of_step": prof[dom]["ms"] / (e0.elapsed_time(e1))}

    # ---- end to end through the C ABI on pinned host buffers
    e2e = None
    if not args.no_e2e:
        avail = mem_available_bytes()
        per_stream = n
        fixed = int(L.pim_snappy_compress_bound(n, bs)) + 64 + n + 4096
        distinct = S
        if avail:
            budget = avail // max(world, 1) // 2
            distinct = int(max(1, min(S, (budget - fixed) // per_stream)))
        h_in = [P.PinnedBuffer(n) for _ in range(distinct)]
        for s in range(distinct):
            torch.from_numpy(h_in[s].array).copy_(d_in[s])
        h_comp = P.PinnedBuffer(int(L.pim_snappy_compress_bound(n, bs)) + 64)
        h_back = P.PinnedBuffer(n + 4096)
        torch.cuda.synchronize()

        phase = {"compress": 0.0, "decompress": 0.0}

        def e2e_stream(s):
            tp0 = time.perf_counter()
            inp, out, rt = P.HostBufferContext(), P.HostBufferContext(), P.ProgramRuntime()
            inp.buffer = inp.curr = h_in[s % distinct].ptr
            inp.length, inp.max, out.max = n, 2**63, 2**63
            out.buffer = out.curr = h_comp.ptr
            rc = L.snappy_compress_dpu(C.byref(inp), C.byref(out), bs, C.byref(rt))
            assert rc == 0, f"snappy_compress_dpu rc={rc}"
            clen = int(out.length)
            tp1 = time.perf_counter()
            # what setup_decompression does, minus its allocation: consume the length varint
            hdr = bytes(h_comp.array[:5])
            k = next(i for i in range(5) if not hdr[i] & 0x80) + 1
            dinp, dout = P.HostBufferContext(), P.HostBufferContext()
            dinp.buffer, dinp.curr, dinp.length, dinp.max = h_comp.ptr, h_comp.ptr + k, clen, 2**63
            dout.buffer = dout.curr = h_back.ptr
            dout.length, dout.max = n, 2**63
            rc = L.snappy_decompress_dpu(C.byref(dinp), C.byref(dout), C.byref(rt))
            assert rc == 0, f"snappy_decompress_dpu rc={rc}"
            phase["compress"] += tp1 - tp0
            phase["decompress"] += time.perf_counter() - tp1
            return clen

        for _ in range(max(1, min(args.warmup, 2))):
            clens = [e2e_stream(s) for s in range(S)]
        assert np.array_equal(h_back.array[:n], h_in[(S - 1) % distinct].array), "e2e round trip mismatch"
        # the link the end-to-end number is bounded by: plain pinned copies of one stream, each direction
        d_tmp = d_back[0]
        pcie = {}
        rtlib = C.CDLL("libcudart.so.12")  # the runtime torch already loaded
        rtlib.cudaMemcpyAsync.argtypes = [C.c_void_p, C.c_void_p, C.c_size_t, C.c_int, C.c_void_p]
        for name, fn in (("h2d", lambda: rtlib.cudaMemcpyAsync(d_tmp.data_ptr(), h_in[0].ptr, n, 4, st)),
                         ("d2h", lambda: rtlib.cudaMemcpyAsync(h_back.ptr, d_tmp.data_ptr(), n, 4, st))):
            assert fn() == 0
            torch.cuda.synchronize()
            c0, c1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            c0.record()
            assert fn() == 0
            c1.record()
            torch.cuda.synchronize()
            pcie[name + "_gbps"] = n / (c0.elapsed_time(c1) * 1e-3) / 1e9
        # ... and the same with every rank copying both ways at once (what the host side gives N GPUs together)
        s_up, s_dn = torch.cuda.Stream(), torch.cuda.Stream()
        sync_all()
        tb0 = time.perf_counter()
        assert rtlib.cudaMemcpyAsync(d_tmp.data_ptr(), h_in[0].ptr, n, 4, s_up.cuda_stream) == 0
        assert rtlib.cudaMemcpyAsync(h_back.ptr, d_in[0].data_ptr(), n, 4, s_dn.cuda_stream) == 0
        torch.cuda.synchronize()
        both = n / (time.perf_counter() - tb0) / 1e9
        pcie["both_directions_all_ranks_gbps_per_direction"] = {"slowest_rank": -max_over_ranks(-both), "sum_over_ranks": sum_over_ranks(both)}
        assert np.array_equal(h_back.array[:n], h_in[0].array)  # (h_back now holds stream 0 again)
        sync_all()
        phase["compress"] = phase["decompress"] = 0.0
        t0 = time.perf_counter()
        for _ in range(args.steps):
            for s in range(S):
                e2e_stream(s)
        torch.cuda.synchronize()
        dt = max_over_ranks(time.perf_counter() - t0) / args.steps
        sync_all()
        csum = sum(clens)
        e2e = {"value": total_U / dt / 1e9, "unit": "GB/s", "h2d_bytes_per_step": int(sum_over_ranks(U + csum)), "d2h_bytes_per_step": int(sum_over_ranks(csum + U)),
               "ms_per_step": dt * 1e3, "distinct_host_streams": distinct,
               "compress_gbps": U * args.steps / phase["compress"] / 1e9, "decompress_gbps": U * args.steps / phase["decompress"] / 1e9,
               "pinned_copy": pcie,
               "api": "snappy_compress_dpu + snappy_decompress_dpu (C ABI) on page-locked host buffers, one GPU per rank"}

    # ---- several GPUs serving ONE stream through the C ABI (contiguous block ranges per GPU, host stitch): rank 0 alone,
    # the other ranks parked at the barrier
    e2e_sharded = None
    if world > 1 and not args.no_e2e:
        sync_all()
        if rank == 0:
            def one_stream():
                t0 = time.perf_counter()
                clen = e2e_stream(0)
                return clen, time.perf_counter() - t0
            try:  # (never let this extra leg take the headline run down; the other ranks wait at the barrier below)
                seen = P.set_devices(list(range(world)))
                if seen != world:
                    e2e_sharded = {"skipped": f"rank 0 sees {seen} of {world} GPUs"}
                else:
                    one_stream()
                    times = []
                    for _ in range(3):
                        clen_n, dt1 = one_stream()
                        times.append(dt1)
                    comp_n = h_comp.array[:clen_n].copy()
                    back_ok = bool(np.array_equal(h_back.array[:n], h_in[0].array))
                    assert P.set_devices([local_rank]) == 1
                    clen_1, _ = one_stream()
                    e2e_sharded = {"value": n / min(times) / 1e9, "unit": "GB/s", "n_gpus": world, "stream_bytes": n,
                                   "bytes_equal_to_1gpu": bool(clen_1 == clen_n and np.array_equal(h_comp.array[:clen_1], comp_n)) and back_ok,
                                   "api": "snappy_compress_dpu + snappy_decompress_dpu on ONE stream with pim_snappy_set_devices([0..N-1]), best of 3"}
            except Exception as ex:  # noqa: BLE001
                e2e_sharded = {"error": repr(ex)[:300]}
            finally:
                P.set_devices([local_rank])
        sync_all()

    # ---- the same two kernels on other inputs (one 1-stream launch each, after one warm launch): text, the reference's
    # fixtures tiled to the stream size, the binary-like mix
    by_input = None
    if rank == 0 and args.by_input:
        import gzip
        gold = os.path.join(ROOT, "tests", "golden")
        def corpus_fill():
            parts = [np.frombuffer(gzip.open(os.path.join(gold, f + ".txt.gz"), "rb").read(), dtype=np.uint8) for f in ("world192", "plrabn12", "terror2")]
            xml = P.decompress(np.frombuffer(gzip.open(os.path.join(gold, "xml.snappy.gz"), "rb").read(), dtype=np.uint8))  # (xml.txt is not shipped)
            p = np.concatenate(parts + [xml])
            d_in[0].copy_(torch.from_numpy(np.tile(p, n // p.size + 1)[:n]))
        fills = {"SYN-TEXT": lambda: P.synth_device(P.SYNTH_TEXT, 0x5BA0, d_in[0].data_ptr(), n, st),
                 "CORPUS (world192+plrabn12+terror2+xml tiled)": corpus_fill,
                 "SYN-BIN": lambda: P.synth_device(P.SYNTH_BIN, 0x0B1A, d_in[0].data_ptr(), n, st)}
        by_input = {}
        for name, fill in fills.items():
          try:  # (an extra leg: a failure is recorded, it does not take the headline run down)
            fill()
            torch.cuda.synchronize()
            P.profile(True)
            for r in range(2):
                P.profile_read()
                P.compress_device(d_in[0].data_ptr(), n, bs, scratch[0].data_ptr(), scratch[0].numel(), d_comp[0].data_ptr(), d_off[0].data_ptr(), st)
                P.decompress_device(d_comp[0].data_ptr(), d_off[0].data_ptr(), nb, bs, n, d_back[0].data_ptr(), d_status.data_ptr(), st)
                torch.cuda.synchronize()
                pr = P.profile_read()
            P.profile(False)
            assert int(d_status[0].item()) == 0 and torch.equal(d_back[0], d_in[0]), f"round trip failed on {name}"
            cb = int(d_off[0][-1].item())
            row = {"stream_bytes": n, "compressed_fraction": cb / n}
            for k in ("compress", "decompress"):
                ms = pr[k]["ms"]
                row[k] = {"ms": ms, "uncompressed_gbps": n / (ms * 1e-3) / 1e9, "frac_of_hbm_peak": (n + cb) / (ms * 1e-3) / 1e9 / peak}
            row["compress"]["scan_compact_ms"] = pr["scan"]["ms"] + pr["compact"]["ms"]
            by_input[name] = row
          except Exception as ex:  # noqa: BLE001
            P.profile(False)
            by_input[name] = {"error": repr(ex)[:300]}

    cpu = None
    if rank == 0 and world == 1 and not args.no_cpu_baseline:
        cpu = cpu_baseline_1thread(P, args)

    if rank == 0:
        line = {"metric": METRIC, "value": value, "unit": "GB/s", "n_gpus": world, "steps": args.steps, "warmup": max(args.warmup, 1),
                "ms_per_step": dev_ms, "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "u8",
                "data": "synthetic",
                "config": {"workload": workload_name(args, world), "l2": f"inputs ({U / GIB:g} GiB per GPU, every stream read once per step) far exceed the 126 MB L2; no flush needed",
                           "compressed_fraction": comp_bytes / U, "bytes_per_gpu": U},
                "roofline": roofline, "kernels": kern, "by_input": by_input, "cpu_baseline": cpu, "e2e": e2e, "e2e_sharded": e2e_sharded,
                "gpu_launches": int(launches), "clocks": clocks}
        sys.stdout.flush()
        os.dup2(saved_stdout, 1)
        print(json.dumps(line), flush=True)
        os.dup2(2, 1)
    if dist is not None:
        dist.destroy_process_group()
    return 0


def main():
    args = parse_args()
    if args.impl == "reference":
        return run_reference_arm(args)
    return run_ours(args)


if __name__ == "__main__":
    sys.exit(main())
