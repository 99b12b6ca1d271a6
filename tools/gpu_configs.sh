#!/bin/bash
# BASELINE.json configs 1-4 on the GPU box -> gpurun_out/r2_config*.json, r2_blocksweep.jsonl (copied to profiles/ here).
set -u
mkdir -p gpurun_out; O=gpurun_out
timeout 600 python tools/configs.py > $O/r2_config12.jsonl 2> $O/r2_config12.err; echo "configs12 rc=$?"; cut -c1-1500 $O/r2_config12.jsonl
: > $O/r2_blocksweep.jsonl
for bs in 4096 8192 16384 32768 65535 65536; do
  timeout 200 python tools/kbench.py --kinds bin,text --mib 80 --reps 5 --block-size $bs 2>&1 | tee -a $O/r2_blocksweep.jsonl
done
timeout 200 python tools/config4.py > $O/r2_config4.json 2>&1; cat $O/r2_config4.json
