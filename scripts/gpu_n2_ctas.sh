#!/usr/bin/env bash
set -u
TAG=${1:-c}
OUT=gpurun_out; mkdir -p $OUT
: > $OUT/${TAG}_ctas.jsonl
for K in 2 4 8 0; do
  timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29511 \
    bench.py --gpus 2 --steps 30 --warmup 5 --no-cpu --no-extras --emitters 0 --nccl-ctas $K 2> $OUT/${TAG}_ctas_$K.err |
    python -c "import sys,json; d=json.loads(sys.stdin.read()); print(json.dumps({'nccl_ctas': $K, 'ms_per_step': d['ms_per_step'], 'value': d['value'], 'tile_ms': d['roofline']['kernel_ms']}))" | tee -a $OUT/${TAG}_ctas.jsonl
  tail -2 $OUT/${TAG}_ctas_$K.err
done
