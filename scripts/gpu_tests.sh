#!/usr/bin/env bash
# GPU parity tests only (all of them, no -x) + a short bench.  bash scripts/gpu_tests.sh <tag>
set -u
TAG=${1:-t}
OUT=gpurun_out
mkdir -p $OUT
timeout 1500 python -m pytest tests -m gpu -q > $OUT/${TAG}_pytest_gpu.log 2>&1
echo "pytest exit $?"; tail -40 $OUT/${TAG}_pytest_gpu.log
timeout 600 python bench.py --steps 20 --warmup 5 --no-cpu > $OUT/${TAG}_bench.json 2> $OUT/${TAG}_bench.err
echo "bench exit $?"; cat $OUT/${TAG}_bench.json; tail -3 $OUT/${TAG}_bench.err
