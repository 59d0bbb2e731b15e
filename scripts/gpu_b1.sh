#!/usr/bin/env bash
set -u
TAG=${1:-b}
OUT=gpurun_out; mkdir -p $OUT
timeout 600 python bench.py --steps 20 --warmup 5 --no-cpu --emitters 0 > $OUT/${TAG}_bench_nocpu.json 2> $OUT/${TAG}_bench.err
echo "bench exit $?"; cat $OUT/${TAG}_bench_nocpu.json; tail -3 $OUT/${TAG}_bench.err
