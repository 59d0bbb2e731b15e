#!/usr/bin/env bash
# Two GPUs: mask tests on GPU 0, then the N=2 bench with both exchange forms.  bash scripts/gpu_n2.sh <tag>
set -u
TAG=${1:-n2}
OUT=gpurun_out; mkdir -p $OUT
timeout 600 python -m pytest tests/test_gpu_masks.py -q > $OUT/${TAG}_pytest_masks.log 2>&1
echo "pytest exit $?"; tail -15 $OUT/${TAG}_pytest_masks.log
for X in masks lists; do
  timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29511 \
    bench.py --gpus 2 --steps 30 --warmup 5 --no-cpu --no-extras --emitters 0 --exchange $X > $OUT/${TAG}_bench_$X.json 2> $OUT/${TAG}_bench_$X.err
  echo "bench $X exit $?"; cat $OUT/${TAG}_bench_$X.json; tail -5 $OUT/${TAG}_bench_$X.err
done
