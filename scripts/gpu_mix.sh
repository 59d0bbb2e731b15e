#!/usr/bin/env bash
set -u
TAG=${1:-m}
OUT=gpurun_out; mkdir -p $OUT
timeout 600 python -m pytest tests/test_gpu_extras.py -q -k "billboard or particle_mesh or sorted_lists" > $OUT/${TAG}_pytest.log 2>&1
echo "pytest exit $?"; tail -5 $OUT/${TAG}_pytest.log
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -k regex:'sort_|rank_mask|expand_|billboard' -c 200 --csv --log-file $OUT/${TAG}_sort_launches.csv \
  python bench.py --steps 1 --warmup 3 --no-cpu > $OUT/${TAG}_ncu.log 2>&1
echo "ncu exit $?"; tail -3 $OUT/${TAG}_ncu.log
