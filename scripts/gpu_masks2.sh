#!/usr/bin/env bash
set -u
TAG=${1:-m2}
OUT=gpurun_out; mkdir -p $OUT
timeout 600 python -m pytest tests/test_gpu_masks.py -q > $OUT/${TAG}_pytest.log 2>&1
echo "pytest exit $?"; tail -5 $OUT/${TAG}_pytest.log
timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29511 \
    bench.py --gpus 2 --steps 30 --warmup 5 --no-cpu --emitters 0 > $OUT/${TAG}_bench_n2.json 2> $OUT/${TAG}_bench_n2.err
echo "bench exit $?"; cat $OUT/${TAG}_bench_n2.json; tail -4 $OUT/${TAG}_bench_n2.err
