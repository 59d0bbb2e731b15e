#!/usr/bin/env bash
set -u
TAG=${1:-s}
OUT=gpurun_out; mkdir -p $OUT
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -k regex:'sort_|rank_mask|expand_' -c 200 --csv --log-file $OUT/${TAG}_sort_launches.csv \
  python bench.py --steps 1 --warmup 3 --no-cpu --emitters 0 > $OUT/${TAG}_ncu.log 2>&1
echo "ncu exit $?"; tail -3 $OUT/${TAG}_ncu.log
