#!/usr/bin/env bash
# Two GPUs: the N=2 bench (exchange forms given as arguments after the tag).  bash scripts/gpu_n2.sh <tag> masks [lists]
set -u
TAG=${1:-n2}; shift
OUT=gpurun_out; mkdir -p $OUT
for X in "$@"; do
  timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29511 \
    bench.py --gpus 2 --steps 30 --warmup 5 --no-cpu --no-extras --emitters 0 --exchange $X > $OUT/${TAG}_bench_$X.json 2> $OUT/${TAG}_bench_$X.err
  echo "bench $X exit $?"; cat $OUT/${TAG}_bench_$X.json; tail -5 $OUT/${TAG}_bench_$X.err
done
