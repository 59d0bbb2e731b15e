#!/usr/bin/env bash
# Full ncu capture of the side kernels (one launch each where possible): bash scripts/gpu_side_ncu.sh <tag>
set -u
TAG=${1:-side}
OUT=gpurun_out; mkdir -p $OUT
timeout 330 ncu --set full --clock-control none --import-source on \
  -k regex:'sort_scatter|sort_hist|rank_mask|pack_instances|expand_emit|billboard|particles_mesh' \
  -c 28 -f -o $OUT/${TAG}_side_kernels python bench.py --steps 1 --warmup 3 --no-cpu --e2e-shards 1 --extras-reps 1 > $OUT/${TAG}_ncu_side.log 2>&1
echo "ncu side exit $?"; tail -2 $OUT/${TAG}_ncu_side.log
