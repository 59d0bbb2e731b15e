#!/usr/bin/env bash
# Short visit: the side-row tests + the bench.  bash scripts/gpu_quick2.sh <tag>
set -u
TAG=${1:-q}
OUT=gpurun_out; mkdir -p $OUT
timeout 900 python -m pytest tests/test_gpu_masks.py tests/test_gpu_extras.py -q -k "mask or sort" > $OUT/${TAG}_pytest.log 2>&1
echo "pytest exit $?"; tail -8 $OUT/${TAG}_pytest.log
timeout 600 python bench.py --steps 20 --warmup 5 --no-cpu > $OUT/${TAG}_bench_nocpu.json 2> $OUT/${TAG}_bench.err
echo "bench exit $?"; cat $OUT/${TAG}_bench_nocpu.json; tail -3 $OUT/${TAG}_bench.err
