#!/bin/bash
# Development aid: compress throughput vs warp pairs per CTA (text, 256 MiB): shared-memory-table pairs alone, then
# with global-table pairs added.
mkdir -p gpurun_out; O=gpurun_out/occ_sweep.log; : > $O
for cfg in "1 0" "2 0" "4 0" "6 0" "7 0" "7 1" "7 2" "7 3" "7 4" "7 5" "7 6" "7 7" "4 4" "0 7"; do set -- $cfg
  echo -n "n_sm=$1 n_gt=$2 " >> $O
  PIM_SNAPPY_NSM=$1 PIM_SNAPPY_NGT=$2 timeout 120 python tools/kbench.py --kinds text --mib 256 --reps 2 2>&1 | sed 's/.*"comp_ms": \([0-9.]*\), "comp_GBps": \([0-9.]*\).*/comp_ms=\1 GBps=\2/' >> $O
done
cat $O
