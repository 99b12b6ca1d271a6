#!/usr/bin/env python3
"""Development aid: pinned-copy bandwidth H2D, D2H and both at once (the ceiling of the end-to-end numbers)."""
import torch
n = 1 << 30
h_a = torch.empty(n, dtype=torch.uint8).pin_memory()
h_b = torch.empty(n, dtype=torch.uint8).pin_memory()
d_a = torch.empty(n, dtype=torch.uint8, device="cuda")
d_b = torch.empty(n, dtype=torch.uint8, device="cuda")
s1, s2 = torch.cuda.Stream(), torch.cuda.Stream()


def run(h2d, d2h):
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    s1.wait_stream(torch.cuda.current_stream()); s2.wait_stream(torch.cuda.current_stream())
    if h2d:
        with torch.cuda.stream(s1):
            d_a.copy_(h_a, non_blocking=True)
    if d2h:
        with torch.cuda.stream(s2):
            h_b.copy_(d_b, non_blocking=True)
    torch.cuda.current_stream().wait_stream(s1); torch.cuda.current_stream().wait_stream(s2)
    e1.record()
    torch.cuda.synchronize()
    return e0.elapsed_time(e1)


for name, a, b in (("h2d", 1, 0), ("d2h", 0, 1), ("both", 1, 1)):
    run(a, b)
    ms = min(run(a, b) for _ in range(3))
    print(f"{name}: {ms:.2f} ms -> {(a + b) * n / ms / 1e6:.1f} GB/s total, {n / ms / 1e6:.1f} GB/s per direction")
