#!/usr/bin/env bash
# bash scripts/sweep_tiling.sh <tag> "<tile> <inner> <unit>" ...   tile kernel time of the bench scene per tiling
set -u
TAG=${1:?tag}; shift
OUT=gpurun_out; mkdir -p $OUT
: > $OUT/${TAG}_tiling.jsonl
for V in "$@"; do
  set -- $V
  echo "== tile=$1 inner=$2 unit=$3" | tee -a $OUT/${TAG}_tiling.jsonl
  timeout 300 python bench.py --steps 40 --warmup 5 --no-cpu --no-extras --emitters 0 --no-parity --tile-nodes $1 --inner-nodes $2 --unit-nodes $3 2>>$OUT/${TAG}_tiling.err |
    python -c "import sys,json; d=json.loads(sys.stdin.read()); print(json.dumps({'ms_per_step': d['ms_per_step'], 'kernel_ms': d['roofline']['kernel_ms'], 'frac': d['roofline']['frac'], 'step_frac': d['roofline']['step_frac'], 'e2e_ms': d['e2e']['ms_per_step']}))" | tee -a $OUT/${TAG}_tiling.jsonl
done
