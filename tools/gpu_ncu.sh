#!/bin/bash
# ncu full capture of one kernel on a small kbench run.  KARGS = kbench args, KREGEX = kernel regex, SKIP/COUNT.
set -u
mkdir -p gpurun_out
O=gpurun_out
KARGS=${KARGS:---mib 64 --reps 1 --kinds text}
timeout 300 python tools/kbench.py $KARGS > $O/ncu_plain.log 2>&1 &&
timeout 900 ncu --set full --clock-control none --import-source on -k regex:${KREGEX:-snappy_decompress} -s ${SKIP:-1} -c ${COUNT:-1} -f -o $O/${OUT:-prof_dec} python tools/kbench.py $KARGS > $O/ncu_run.log 2>&1
echo "ncu rc=$?"; tail -3 $O/ncu_run.log
