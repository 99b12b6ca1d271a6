set -u
SMALL="--streams-per-gpu 1 --stream-bytes 268435456 --steps 1 --warmup 1 --no-e2e --no-cpu-baseline"
for v in 0 1024 4096; do
  make -s -C pim-compression_b200 clean >/dev/null; make -s -C pim-compression_b200 all EXTRA="-DPIM_PF_AHEAD=$v" > /dev/null 2>&1
  echo "=== PF_AHEAD=$v"
  timeout 300 ncu --metrics dram__bytes_read.sum,dram__bytes_write.sum,lts__t_sector_hit_rate.pct,gpu__time_duration.sum --clock-control none -k regex:snappy_compress -s 1 -c 1 python bench.py $SMALL 2>&1 | grep -E "dram__bytes|hit_rate|time_duration"
done
make -s -C pim-compression_b200 clean >/dev/null; make -s -C pim-compression_b200 all >/dev/null 2>&1
