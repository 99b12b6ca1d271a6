#!/usr/bin/env python3
"""Stall-reason totals for a line range of a source file in an .ncu-rep.  Usage: ncu_stalls.py rep file_prefix lo hi"""
import csv, subprocess, sys, io, collections
rep, pref, lo, hi = sys.argv[1], sys.argv[2], int(sys.argv[3]), int(sys.argv[4])
out = subprocess.run(["ncu", "-i", rep, "--page", "source", "--csv", "--print-source", "cuda,sass"], capture_output=True, text=True).stdout
tot = collections.Counter(); fname = ""; hdr = None; per_line = collections.defaultdict(collections.Counter)
for r in csv.reader(io.StringIO(out)):
    if not r: continue
    if r[0] == "File Path": fname = r[1].split("/")[-1]; continue
    if r[0] == "Line No": hdr = r; continue
    if hdr is None or len(r) != len(hdr): continue
    try: ln = int(r[0])
    except ValueError: continue
    if not fname.startswith(pref) or not (lo <= ln <= hi): continue
    for i, h in enumerate(hdr):
        if h.startswith("stall_") and "Not Issued" not in h:
            try: v = float(r[i])
            except ValueError: v = 0
            tot[h] += v; per_line[ln][h] += v
s = sum(tot.values()) or 1
print({k: round(100*v/s, 1) for k, v in tot.most_common(8)}, "total", s)
for ln, c in sorted(per_line.items(), key=lambda kv: -sum(kv[1].values()))[:12]:
    print(ln, int(sum(c.values())), {k.replace("stall_", ""): int(v) for k, v in c.most_common(3)})
