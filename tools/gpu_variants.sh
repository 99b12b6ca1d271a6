#!/bin/bash
# Build-flag sweep on the GPU box: VARIANTS="name|flags;name|flags" -> rebuild, kbench each (development aid).
set -u
mkdir -p gpurun_out
O=gpurun_out/variants.log
: > $O
IFS=';' read -ra VS <<< "${VARIANTS}"
for v in "${VS[@]}"; do
  name="${v%%|*}"; flags="${v#*|}"
  echo "=== $name : $flags" | tee -a $O
  make -s -C pim-compression_b200 clean >/dev/null
  if ! make -s -C pim-compression_b200 all EXTRA="$flags" >> gpurun_out/variants_build.log 2>&1; then echo "build failed" | tee -a $O; continue; fi
  grep -A3 "snappy_compress_kernel\|snappy_decompress_kernel" pim-compression_b200/build.log | grep "Used" | tee -a $O
  for env in ${ENVS:-none}; do
    if [ "$env" != "none" ]; then export $env; echo "--- $env" | tee -a $O; fi
    timeout 300 python tools/kbench.py ${KBENCH_ARGS:---kinds mix,text --mib 512 --reps 3} 2>&1 | tee -a $O
  done
done
make -s -C pim-compression_b200 clean >/dev/null; make -s -C pim-compression_b200 all >/dev/null 2>&1
