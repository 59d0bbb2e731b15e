#!/usr/bin/env bash
set -u
TAG=${1:-e}
OUT=gpurun_out; mkdir -p $OUT
: > $OUT/${TAG}_e2e_shards.jsonl
for K in 2 4 6 8 12; do
  timeout 300 python bench.py --steps 20 --warmup 5 --no-cpu --no-extras --emitters 0 --e2e-shards $K 2> $OUT/${TAG}_e2e_$K.err |
    python -c "import sys,json; d=json.loads(sys.stdin.read()); print(json.dumps({'e2e_shards': $K, 'ms_per_step': d['ms_per_step'], 'e2e_ms': d['e2e']['ms_per_step'], 'e2e_masks_ms': d['e2e_masks']['ms_per_step']}))" | tee -a $OUT/${TAG}_e2e_shards.jsonl
done
