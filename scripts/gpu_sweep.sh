#!/usr/bin/env bash
# Tiling sweep on the bench scene: bash scripts/gpu_sweep.sh <tag>   (tile_nodes inner_nodes unit_nodes per line)
set -u
TAG=${1:-sweep}
OUT=gpurun_out; mkdir -p $OUT
: > $OUT/${TAG}.jsonl
while read -r T I U; do
  [ -z "$T" ] && continue
  echo "== tile $T inner $I unit $U" | tee -a $OUT/${TAG}.jsonl
  timeout 600 python bench.py --steps 30 --warmup 5 --no-cpu --no-extras --emitters 0 --tile-nodes $T --inner-nodes $I --unit-nodes $U 2>>$OUT/${TAG}.err |
    python -c "import sys,json; d=json.loads(sys.stdin.read()); print(json.dumps({'ms_per_step': d['ms_per_step'], 'kernel_ms': d['roofline']['kernel_ms'], 'frac': d['roofline']['frac'], 'value': d['value']}))" | tee -a $OUT/${TAG}.jsonl
done <<CFG
0 0 0
512 128 64
512 128 128
1024 256 64
512 128 32
256 64 64
1024 256 128
CFG
