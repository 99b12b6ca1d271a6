#!/bin/bash
# One GPU-box visit: parity tests, bench, ncu launch list, ncu full capture of the two codec kernels.
# Usage (from the build container):  gpurun --timeout 1500 -- 'bash tools/gpu_round.sh'
set -u
mkdir -p gpurun_out
O=gpurun_out
SMALL="--streams-per-gpu 1 --stream-bytes 268435456 --steps 1 --warmup 1 --no-e2e --no-cpu-baseline"
nvidia-smi --query-gpu=name,clocks.sm,clocks.max.sm,power.draw --format=csv > $O/smi.txt 2>&1
timeout 900 python -m pytest tests -m gpu -x -q > $O/pytest_gpu.log 2>&1; echo "pytest rc=$?" | tee -a $O/pytest_gpu.log
timeout 600 python bench.py > $O/bench.json 2> $O/bench.err; echo "bench rc=$?"; tail -c 3000 $O/bench.json
timeout 300 python bench.py --impl reference --steps 2 --warmup 1 > $O/bench_ref.json 2> $O/bench_ref.err; echo "ref rc=$?"; cat $O/bench_ref.json
if [ "${SKIP_NCU:-0}" = "0" ]; then
timeout 300 python bench.py $SMALL > $O/plain_small.log 2>&1 &&
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file $O/launches.csv python bench.py $SMALL > $O/ncu_list.log 2>&1
echo "ncu list rc=$?"
timeout 300 python bench.py $SMALL > $O/plain_small2.log 2>&1 &&
timeout 900 ncu --set full --clock-control none --import-source on -k regex:snappy_ -s 2 -c 2 -f -o $O/prof python bench.py $SMALL > $O/ncu_full.log 2>&1
echo "ncu full rc=$?"
fi
ls -la $O
