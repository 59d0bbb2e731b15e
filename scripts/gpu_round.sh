#!/usr/bin/env bash
# One GPU-box session: parity tests, the bench line, the ncu launch list and one full capture of the tile kernel.
# Usage (from the repo root, under gpurun): bash scripts/gpu_round.sh <tag>
set -u
TAG=${1:-r01}
OUT=gpurun_out
mkdir -p $OUT
nvidia-smi --query-gpu=name,clocks.sm,clocks.max.sm,power.draw --format=csv > $OUT/${TAG}_smi.txt 2>&1
timeout 1500 python -m pytest tests -m gpu -x -q > $OUT/${TAG}_pytest_gpu.log 2>&1
echo "pytest exit $?" >> $OUT/${TAG}_pytest_gpu.log
tail -5 $OUT/${TAG}_pytest_gpu.log
timeout 900 python bench.py --steps 20 --warmup 5 > $OUT/${TAG}_bench.json 2> $OUT/${TAG}_bench.err
echo "bench exit $?"; cat $OUT/${TAG}_bench.json
timeout 600 python bench.py --impl reference --steps 3 --warmup 1 > $OUT/${TAG}_bench_ref.json 2> $OUT/${TAG}_bench_ref.err
echo "bench ref exit $?"; cat $OUT/${TAG}_bench_ref.json
SHORT="python bench.py --steps 3 --warmup 3 --no-cpu --no-extras --emitters 0"
$SHORT > $OUT/${TAG}_plain.log 2>&1 &&
ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file $OUT/${TAG}_launches.csv $SHORT > $OUT/${TAG}_ncu_launches.log 2>&1
echo "ncu launches exit $?"
$SHORT > $OUT/${TAG}_plain2.log 2>&1 &&
ncu --set full --clock-control none --import-source on -k regex:tile_kernel -s 3 -c 2 -f -o $OUT/${TAG}_tile $SHORT > $OUT/${TAG}_ncu_tile.log 2>&1
echo "ncu tile exit $?"
ls -la $OUT
