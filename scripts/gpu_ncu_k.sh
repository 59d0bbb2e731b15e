#!/usr/bin/env bash
# ncu launch list restricted to a kernel regex: bash scripts/gpu_ncu_k.sh <tag> <regex>
set -u
TAG=${1:-k}; RE=$2
OUT=gpurun_out; mkdir -p $OUT
timeout 600 ncu --metrics gpu__time_duration.sum,smsp__inst_executed.sum,l1tex__throughput.avg.pct_of_peak_sustained_active,smsp__issue_active.avg.pct_of_peak_sustained_active --clock-control none -k regex:"$RE" -c 12 --csv --log-file $OUT/${TAG}_k.csv \
  python bench.py --steps 3 --warmup 3 --no-cpu --no-extras --emitters 0 > $OUT/${TAG}_k.log 2>&1
echo "ncu exit $?"; tail -14 $OUT/${TAG}_k.csv | cut -c1-300
