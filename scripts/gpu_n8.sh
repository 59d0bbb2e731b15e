#!/usr/bin/env bash
set -u
TAG=${1:-n8}; N=${2:-8}; shift; shift
OUT=gpurun_out; mkdir -p $OUT
timeout 420 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29511 \
    bench.py --gpus $N --steps 30 --warmup 5 --no-cpu --no-extras --emitters 0 "$@" > $OUT/${TAG}_bench_n$N.json 2> $OUT/${TAG}_bench_n$N.err
echo "bench exit $?"; cat $OUT/${TAG}_bench_n$N.json; tail -4 $OUT/${TAG}_bench_n$N.err
