#!/usr/bin/env bash
# Tile-kernel launch variants on the bench scene: bash scripts/sweep_tile.sh <tag> "<ctas> <persist>" ...
set -u
TAG=${1:?tag}; shift
OUT=gpurun_out; mkdir -p $OUT
: > $OUT/${TAG}_sweep.jsonl
for V in "$@"; do
  set -- $V
  echo "== RN_TILE_CTAS=$1 RN_TILE_PERSIST=$2" | tee -a $OUT/${TAG}_sweep.jsonl
  RN_TILE_CTAS=$1 RN_TILE_PERSIST=$2 timeout 300 python bench.py --steps 40 --warmup 5 --no-cpu --no-extras --emitters 0 --no-parity 2>>$OUT/${TAG}_sweep.err |
    python -c "import sys,json; d=json.loads(sys.stdin.read()); print(json.dumps({'ms_per_step': d['ms_per_step'], 'kernel_ms': d['roofline']['kernel_ms'], 'frac': d['roofline']['frac'], 'step_frac': d['roofline']['step_frac']}))" | tee -a $OUT/${TAG}_sweep.jsonl
done
