#!/bin/bash
# Development aid: compress throughput vs resident CTAs per SM and enabled warp pairs (text, 256 MiB).
mkdir -p gpurun_out; O=gpurun_out/occ_sweep.log; : > $O
for pairs in 1 2 3; do for ctas in 1 2 3 4 5 6; do
  echo -n "pairs=$pairs ctas_per_sm=$ctas " >> $O
  PIM_SNAPPY_PAIRS=$pairs PIM_SNAPPY_CTAS_PER_SM=$ctas timeout 120 python tools/kbench.py --kinds text --mib 256 --reps 2 2>&1 | sed 's/.*"comp_ms": \([0-9.]*\), "comp_GBps": \([0-9.]*\).*/comp_ms=\1 GBps=\2/' >> $O
done; done
cat $O
