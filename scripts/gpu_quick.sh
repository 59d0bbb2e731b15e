#!/usr/bin/env bash
# Quick GPU iteration: parity tests + short bench + ncu launch list.  bash scripts/gpu_quick.sh <tag> [full]
set -u
TAG=${1:-q}
OUT=gpurun_out
mkdir -p $OUT
timeout 900 python -m pytest tests -m gpu -x -q > $OUT/${TAG}_pytest_gpu.log 2>&1
echo "pytest exit $?"; tail -15 $OUT/${TAG}_pytest_gpu.log
timeout 600 python bench.py --steps 50 --warmup 5 --no-cpu > $OUT/${TAG}_bench.json 2> $OUT/${TAG}_bench.err
echo "bench exit $?"; cat $OUT/${TAG}_bench.json; tail -3 $OUT/${TAG}_bench.err
SHORT="python bench.py --steps 3 --warmup 3 --no-cpu --no-extras --emitters 0"
$SHORT > $OUT/${TAG}_plain.log 2>&1 &&
ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file $OUT/${TAG}_launches.csv $SHORT > $OUT/${TAG}_ncu_launches.log 2>&1
echo "ncu launches exit $?"
python scripts/ncu_launch_summary.py $OUT/${TAG}_launches.csv
if [ "${2:-}" = "full" ]; then
$SHORT > $OUT/${TAG}_plain2.log 2>&1 &&
ncu --set full --clock-control none --import-source on -k regex:tile_kernel -s 3 -c 1 -f -o $OUT/${TAG}_tile $SHORT > $OUT/${TAG}_ncu_tile.log 2>&1
echo "ncu tile exit $?"
fi
