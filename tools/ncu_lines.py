#!/usr/bin/env python3
"""Per-CUDA-source-line aggregation of an .ncu-rep (samples, instructions).  Usage: ncu_lines.py rep [topN]"""
import csv, subprocess, sys, io, collections
rep = sys.argv[1]; topn = int(sys.argv[2]) if len(sys.argv) > 2 else 40
import os
out = subprocess.run(["ncu", "-i", rep, "--page", "source", "--csv", "--print-source", "cuda,sass"] + (["-k", "regex:" + os.environ["KERNEL"]] if os.environ.get("KERNEL") else []), capture_output=True, text=True).stdout
agg = collections.OrderedDict()
fname = ""; hdr = None
for r in csv.reader(io.StringIO(out)):
    if not r: continue
    if r[0] == "File Path": fname = r[1].split("/")[-1]; continue
    if r[0] == "Function Name": continue
    if r[0] == "Line No": hdr = r; continue
    if hdr is None or len(r) != len(hdr): continue
    try:
        ln = int(r[0])
    except ValueError:
        continue
    i_s = hdr.index("Warp Stall Sampling (All Samples)"); i_i = hdr.index("Instructions Executed"); i_t = hdr.index("Thread Instructions Executed")
    def f(x):
        try: return float(x)
        except ValueError: return 0.0
    k = (fname, ln)
    a = agg.setdefault(k, [0.0, 0.0, 0.0, r[1].strip()[:100]])
    a[0] += f(r[i_s]); a[1] += f(r[i_i]); a[2] += f(r[i_t])
ts = sum(a[0] for a in agg.values()) or 1; ti = sum(a[1] for a in agg.values()) or 1
print(f"total samples {ts:.0f} warp-inst {ti:.0f}")
mode = sys.argv[3] if len(sys.argv) > 3 else "inst"
key = (lambda kv: kv[1][1]) if mode == "inst" else (lambda kv: kv[1][0])
for (fn, ln), a in sorted(agg.items(), key=key, reverse=True)[:topn]:
    util = a[2] / a[1] / 32 if a[1] else 0
    print(f"{100*a[1]/ti:5.1f}% inst {100*a[0]/ts:5.1f}% samp util {util:4.2f} {fn}:{ln}: {a[3]}")
# ---- region sums (decompress_sm100.cu only)
import os
if os.environ.get("REGIONS"):
    regs = [tuple(map(int, x.split("-"))) for x in os.environ["REGIONS"].split(",")]
    for lo, hi in regs:
        s = sum(a[0] for (fn, ln), a in agg.items() if fn.startswith("decompress") and lo <= ln <= hi)
        i = sum(a[1] for (fn, ln), a in agg.items() if fn.startswith("decompress") and lo <= ln <= hi)
        print(f"lines {lo}-{hi}: {100*s/ts:5.1f}% samp {100*i/ti:5.1f}% inst")
    s = sum(a[0] for (fn, ln), a in agg.items() if not fn.startswith("decompress")); i = sum(a[1] for (fn, ln), a in agg.items() if not fn.startswith("decompress"))
    print(f"other files: {100*s/ts:5.1f}% samp {100*i/ti:5.1f}% inst")
