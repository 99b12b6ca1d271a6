#!/usr/bin/env python3
"""Summarise an .ncu-rep (read here, no GPU): key metrics + top stall source lines."""
import csv, subprocess, sys, io, collections
rep = sys.argv[1]
topn = int(sys.argv[2]) if len(sys.argv) > 2 else 25
raw = subprocess.run(["ncu", "-i", rep, "--page", "raw", "--csv"], capture_output=True, text=True).stdout
rows = list(csv.reader(io.StringIO(raw)))
hdr, units = rows[0], rows[1]
keys = ['gpu__time_duration.sum','dram__bytes_read.sum','dram__bytes_write.sum','sm__warps_active.avg.pct_of_peak_sustained_active','launch__registers_per_thread','launch__occupancy_limit_shared_mem','launch__occupancy_limit_registers','launch__occupancy_limit_warps','smsp__issue_active.avg.pct_of_peak_sustained_active','smsp__inst_executed.sum','smsp__thread_inst_executed_per_inst_executed.ratio','l1tex__data_bank_conflicts_pipe_lsu_mem_shared.sum','l1tex__data_pipe_lsu_wavefronts_mem_shared.sum','smsp__warps_eligible.avg.per_cycle_active','gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed','l1tex__throughput.avg.pct_of_peak_sustained_active','sm__inst_executed_pipe_lsu.avg.pct_of_peak_sustained_active','smsp__inst_executed_pipe_lsu.sum','sm__pipe_alu_cycles_active.avg.pct_of_peak_sustained_active']
for r in rows[2:]:
    name = r[hdr.index('Kernel Name')]
    print("==", name[:80])
    for k in keys:
        if k in hdr:
            i = hdr.index(k); print(f"  {k:72s} {r[i]} {units[i]}")
    for i, h in enumerate(hdr):
        if 'warp_issue_stalled' in h and h.endswith('_per_warp_active.pct'):
            try:
                v = float(r[i])
            except ValueError:
                continue
            if v >= 3: print(f"  stall {h.split('stalled_')[1].split('_per_warp')[0]:28s} {v:.1f}%")
src = subprocess.run(["ncu", "-i", rep, "--page", "source", "--csv"], capture_output=True, text=True).stdout
rows = list(csv.reader(io.StringIO(src)))
# find header row
for hi, r in enumerate(rows):
    if 'Source' in r and any('Samples' in c for c in r):
        break
else:
    print("no source page"); sys.exit(0)
h = rows[hi]
ci = {c: i for i, c in enumerate(h)}
scol = [c for c in h if c.startswith('# Samples') or c == 'Warp Stall Sampling (All Samples)' or 'Sampling' in c]
print("columns:", [c for c in h][:14])
samp = next((c for c in h if 'All Samples' in c or c.startswith('Warp Stall Sampling')), None)
inst = next((c for c in h if c.startswith('Instructions Executed')), None)
print("using", samp, inst)
data = []
for r in rows[hi+1:]:
    if len(r) != len(h): continue
    try:
        s = float(r[ci[samp]]) if samp else 0
        ie = float(r[ci[inst]]) if inst else 0
    except ValueError:
        continue
    data.append((s, ie, r[ci['Source']].strip()[:110], r[ci.get('#', 0)] if '#' in ci else ''))
tot = sum(d[0] for d in data) or 1
toti = sum(d[1] for d in data) or 1
print(f"total samples {tot:.0f}, total inst {toti:.0f}")
for s, ie, txt, ln in sorted(data, reverse=True)[:topn]:
    print(f"  {100*s/tot:5.1f}% samp {100*ie/toti:5.1f}% inst  L{ln}: {txt}")
