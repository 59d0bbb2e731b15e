#!/usr/bin/env bash
# One GPU visit: parity tests, bench, ncu launch list, full captures of the side kernels.  bash scripts/gpu_round.sh <tag>
set -u
TAG=${1:-r}
OUT=gpurun_out
mkdir -p $OUT
timeout 1500 python -m pytest tests -m gpu -q > $OUT/${TAG}_pytest_gpu.log 2>&1
echo "pytest exit $?"; tail -15 $OUT/${TAG}_pytest_gpu.log
timeout 600 python bench.py --steps 20 --warmup 5 --no-cpu > $OUT/${TAG}_bench_nocpu.json 2> $OUT/${TAG}_bench.err
echo "bench exit $?"; cat $OUT/${TAG}_bench_nocpu.json; tail -3 $OUT/${TAG}_bench.err
timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -c 700 --csv --log-file $OUT/${TAG}_launches.csv \
  python bench.py --steps 2 --warmup 3 --no-cpu > $OUT/${TAG}_ncu_launches.log 2>&1
echo "ncu launches exit $?"
timeout 900 ncu --set full --clock-control none --import-source on -k regex:'sort_|pack_instances' -c 14 -o $OUT/${TAG}_side_kernels \
  python bench.py --steps 1 --warmup 3 --no-cpu --emitters 0 > $OUT/${TAG}_ncu_side.log 2>&1
echo "ncu side exit $?"; ls -la $OUT | tail -8
