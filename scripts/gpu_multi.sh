#!/usr/bin/env bash
# bash scripts/gpu_multi.sh <tag> <n_gpus> "<bench args>" ...   one torchrun bench per argument string -> gpurun_out/<tag>_n<N>_<i>.json
set -u
TAG=$1; N=$2; shift 2
OUT=gpurun_out; mkdir -p $OUT
i=0
for ARGS in "$@"; do
  echo "=== N=$N $ARGS"
  timeout 1200 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port $((29500 + i)) \
    bench.py --gpus $N $ARGS > $OUT/${TAG}_n${N}_$i.json 2> $OUT/${TAG}_n${N}_$i.err
  echo "exit $?"; tail -c 600 $OUT/${TAG}_n${N}_$i.err | tail -4
  python - <<PY
import json
try:
    d = json.loads(open("$OUT/${TAG}_n${N}_$i.json").read().strip().splitlines()[-1])
    print(d["config"].get("name"), "n_gpus", d["n_gpus"], "value", round(d["value"], 1), d["unit"], "ms", round(d["ms_per_step"], 4), "scaling", d["scaling"],
          "e2e", round(d["e2e"]["value"], 1), round(d["e2e"]["ms_per_step"], 4), "exchange", d.get("exchange"))
except Exception as e:
    print("no json:", e)
PY
  i=$((i + 1))
done
