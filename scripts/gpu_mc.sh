#!/usr/bin/env bash
# NVLS multicast exchange: probe, then the N-GPU bench with both transports.  bash scripts/gpu_mc.sh <tag> <N>
set -u
TAG=${1:-mc}; N=${2:-2}
OUT=gpurun_out; mkdir -p $OUT
timeout 120 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29512 scripts/probe_symm.py > $OUT/${TAG}_probe.log 2>&1
echo "probe exit $?"; grep -v "^W\|Warning" $OUT/${TAG}_probe.log | tail -8 | cut -c1-300
for X in multicast nccl; do
  timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29513 \
    bench.py --gpus $N --steps 30 --warmup 5 --no-cpu --no-extras --emitters 0 --exchange-transport $X > $OUT/${TAG}_bench_$X.json 2> $OUT/${TAG}_bench_$X.err
  echo "bench $X exit $?"; cut -c1-400 $OUT/${TAG}_bench_$X.json; grep -i "multicast\|error\|Traceback" $OUT/${TAG}_bench_$X.err | tail -5 | cut -c1-300
done
