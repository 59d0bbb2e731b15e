#!/usr/bin/env bash
# Sort + mask parity tests (small ones) + the bench.  bash scripts/gpu_bench_only.sh <tag>
set -u
TAG=${1:-b}
OUT=gpurun_out; mkdir -p $OUT
timeout 600 python -m pytest tests/test_gpu_extras.py tests/test_gpu_masks.py -q -k "sorted_lists or mask" > $OUT/${TAG}_pytest.log 2>&1
echo "pytest exit $?"; tail -5 $OUT/${TAG}_pytest.log
timeout 600 python bench.py --steps 20 --warmup 5 --no-cpu > $OUT/${TAG}_bench_nocpu.json 2> $OUT/${TAG}_bench.err
echo "bench exit $?"; cat $OUT/${TAG}_bench_nocpu.json; tail -3 $OUT/${TAG}_bench.err
