#!/usr/bin/env bash
# Last check of the round on a 2-GPU box: GPU tests, N=1 bench, N=2 bench with the defaults.
set -u
TAG=${1:-last}
OUT=gpurun_out; mkdir -p $OUT
timeout 600 python -m pytest tests -m gpu -q > $OUT/${TAG}_pytest.log 2>&1
echo "pytest exit $?"; tail -3 $OUT/${TAG}_pytest.log
timeout 300 python bench.py --steps 20 --warmup 5 --no-cpu --no-extras --emitters 0 > $OUT/${TAG}_bench_n1.json 2> $OUT/${TAG}_bench_n1.err
echo "bench n1 exit $?"; cut -c1-200 $OUT/${TAG}_bench_n1.json
timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29514 \
    bench.py --gpus 2 --steps 30 --warmup 5 --no-cpu > $OUT/${TAG}_bench_n2.json 2> $OUT/${TAG}_bench_n2.err
echo "bench n2 exit $?"; cut -c1-200 $OUT/${TAG}_bench_n2.json; grep -i "multicast\|Traceback" $OUT/${TAG}_bench_n2.err | tail -3
