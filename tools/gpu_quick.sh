#!/bin/bash
# Quick GPU visit during kernel development: sanity (first mismatch printed), kernel timings.
set -u
mkdir -p gpurun_out
O=gpurun_out
timeout 300 python tests/gpu_sanity.py > $O/sanity.log 2>&1; echo "sanity rc=$?"; tail -15 $O/sanity.log
timeout 300 python tools/kbench.py ${KBENCH_ARGS:-} > $O/kbench.log 2>&1; echo "kbench rc=$?"; cat $O/kbench.log | tail -20
if [ "${RUN_TESTS:-0}" = "1" ]; then
timeout 900 python -m pytest tests -m gpu -x -q > $O/pytest_gpu.log 2>&1; echo "pytest rc=$?"; tail -5 $O/pytest_gpu.log
fi
