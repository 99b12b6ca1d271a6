#!/usr/bin/env python3
"""Development aid: host-link bandwidth with N GPUs copying at once, with and without binding each process to
its GPU's NUMA node before the pinned buffers are allocated.  Usage: numa_probe.py N [bind]"""
import os
import subprocess
import sys
import time


def gpu_numa_node(i):
    import torch
    p = torch.cuda.get_device_properties(i)
    bdf = f"{p.pci_domain_id:04x}:{p.pci_bus_id:02x}:{p.pci_device_id:02x}.0"
    try:
        return int(open(f"/sys/bus/pci/devices/{bdf}/numa_node").read()), bdf
    except Exception:
        return -1, bdf


def node_cpus(node):
    out = []
    for part in open(f"/sys/devices/system/node/node{node}/cpulist").read().strip().split(","):
        a, _, b = part.partition("-")
        out += list(range(int(a), int(b or a) + 1))
    return out


def worker(i, bind, t_start):
    import torch
    node, bdf = gpu_numa_node(i)
    if bind and node >= 0:
        os.sched_setaffinity(0, set(node_cpus(node)) & os.sched_getaffinity(0) or os.sched_getaffinity(0))
    torch.cuda.set_device(i)
    n = 1 << 30
    h_a = torch.empty(n, dtype=torch.uint8).pin_memory()
    h_b = torch.empty(n, dtype=torch.uint8).pin_memory()
    h_a.fill_(1); h_b.fill_(2)
    d_a = torch.empty(n, dtype=torch.uint8, device="cuda")
    d_b = torch.empty(n, dtype=torch.uint8, device="cuda")
    s1, s2 = torch.cuda.Stream(), torch.cuda.Stream()
    torch.cuda.synchronize()
    while time.time() < t_start:
        time.sleep(0.001)
    t0 = time.perf_counter()
    reps = 8
    for _ in range(reps):
        with torch.cuda.stream(s1):
            d_a.copy_(h_a, non_blocking=True)
        with torch.cuda.stream(s2):
            h_b.copy_(d_b, non_blocking=True)
    torch.cuda.synchronize()
    dt = time.perf_counter() - t0
    print(f"gpu {i} {bdf} numa {node} cpus {sorted(os.sched_getaffinity(0))[:3]}..: {reps * n / dt / 1e9:.1f} GB/s per direction (both at once)", flush=True)


if __name__ == "__main__":
    if sys.argv[1] == "worker":
        worker(int(sys.argv[2]), sys.argv[3] == "1", float(sys.argv[4]))
    else:
        N = int(sys.argv[1])
        bind = "1" if len(sys.argv) > 2 and sys.argv[2] == "bind" else "0"
        t_start = time.time() + 25
        ps = [subprocess.Popen([sys.executable, __file__, "worker", str(i), bind, str(t_start)]) for i in range(N)]
        for p in ps:
            p.wait()
