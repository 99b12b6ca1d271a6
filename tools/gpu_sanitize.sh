#!/bin/bash
# compute-sanitizer over tools/memcheck_small.py (ONE tool per gpurun call: TOOL=memcheck | racecheck).
# Two passes: the default pair layout, and PIM_SNAPPY_NSM=1 (global-memory hash tables in use).
set -u
mkdir -p gpurun_out
TOOL=${TOOL:-memcheck}
O=gpurun_out/r2_sanitizer_$TOOL.txt
: > $O
SMALL=0; [ "$TOOL" = "racecheck" ] && SMALL=1
timeout 300 python tools/memcheck_small.py > gpurun_out/sanitize_plain.log 2>&1 || { echo "plain run failed"; tail -5 gpurun_out/sanitize_plain.log; exit 1; }
for env in "X=1" "PIM_SNAPPY_NSM=1"; do
  echo "=== compute-sanitizer --tool $TOOL, $env" >> $O
  env $env SANITIZE_SMALL=$SMALL timeout 1500 compute-sanitizer --tool $TOOL --print-limit 20 python tools/memcheck_small.py >> $O 2>&1
  echo "rc=$?" >> $O
done
grep -E "===|ERROR SUMMARY|RACECHECK SUMMARY|memcheck_small|rc=" $O
