#!/usr/bin/env bash
# What the driver runs at round end, plus the ncu launch list: bash scripts/gpu_final.sh <tag>
set -u
TAG=${1:-final}
OUT=gpurun_out; mkdir -p $OUT
timeout 1500 python -m pytest tests -m gpu -q --durations=8 > $OUT/${TAG}_pytest_gpu.log 2>&1
echo "pytest exit $?"; tail -14 $OUT/${TAG}_pytest_gpu.log
timeout 300 python -c "import __graft_entry__ as g; g.smoke()" > $OUT/${TAG}_smoke.log 2>&1
echo "smoke exit $?"; tail -3 $OUT/${TAG}_smoke.log
timeout 900 python bench.py --impl reference > $OUT/${TAG}_bench_ref.json 2> $OUT/${TAG}_bench_ref.err
echo "ref exit $?"; cat $OUT/${TAG}_bench_ref.json
timeout 900 python bench.py > $OUT/${TAG}_bench.json 2> $OUT/${TAG}_bench.err
echo "bench exit $?"; cat $OUT/${TAG}_bench.json; tail -3 $OUT/${TAG}_bench.err
timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -c 800 --csv --log-file $OUT/${TAG}_launches.csv \
  python bench.py --steps 2 --warmup 3 --no-cpu > $OUT/${TAG}_ncu_launches.log 2>&1
echo "ncu launches exit $?"
