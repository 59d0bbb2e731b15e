#!/usr/bin/env bash
# The round's closing GPU visit:  bash scripts/gpu_final.sh <tag>
#   the no-alloc test alone in a fresh process, then the whole GPU suite, smoke, the four 1-GPU configs, the reference
#   arm, the whole-scene parity gate and the ncu launch list.
set -u
TAG=${1:?tag}; OUT=gpurun_out; mkdir -p $OUT
timeout 600 python -m pytest tests/test_zz_no_alloc.py -m gpu -q > $OUT/${TAG}_pytest_no_alloc_alone.log 2>&1; echo "no_alloc alone exit $?"; tail -3 $OUT/${TAG}_pytest_no_alloc_alone.log
bash scripts/gpu.sh $TAG tests smoke "bench" "bench:--impl reference --steps 5 --warmup 1" "bench:--config C1" "bench:--config C2" "bench:--config C3" \
  "bench:--verify-full --steps 20 --warmup 5 --no-cpu --no-extras" launches
