#!/bin/bash
# BASELINE.json config 3 (block-size sweep on the mozilla / spamfile stand-ins) and config 4, device resident.
mkdir -p gpurun_out; O=gpurun_out/blocksweep.log; : > $O
for bs in 4096 8192 16384 32768 65536; do
  timeout 200 python tools/kbench.py --kinds bin,text --mib 80 --reps 5 --block-size $bs 2>&1 | tee -a $O
done
timeout 200 python tools/config4.py > gpurun_out/config4.json 2>&1; cat gpurun_out/config4.json
