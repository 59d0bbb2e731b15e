#!/usr/bin/env bash
# Full ncu capture of selected kernels of the short bench.  bash scripts/gpu_ncu.sh <tag> <kernel-regex> [skip] [count]
set -u
TAG=$1; RE=$2; SKIP=${3:-3}; CNT=${4:-1}
OUT=gpurun_out; mkdir -p $OUT
SHORT="python bench.py --steps 3 --warmup 3 --no-cpu --no-extras --emitters 0"
$SHORT > $OUT/${TAG}_plain.log 2>&1 &&
ncu --set full --clock-control none --import-source on -k regex:$RE -s $SKIP -c $CNT -f -o $OUT/${TAG} $SHORT > $OUT/${TAG}_ncu.log 2>&1
echo "ncu exit $?"; tail -3 $OUT/${TAG}_ncu.log
