#!/usr/bin/env bash
# bash scripts/diag_tile.sh : the tile kernel's experiment knobs against the priorities scene
OUT=gpurun_out; mkdir -p $OUT
: > $OUT/diag_tile.log
for V in "" "RN_TILE_DEBUG=1" "RN_TILE_DEBUG=2" "RN_TILE_DEBUG=4" "RN_TILE_DEBUG=7" "RN_TILE_CTAS=5" "RN_TILE_PERSIST=0" "RN_TILE_PERSIST=1"; do
  env $V timeout 300 python scripts/diag_tile.py >> $OUT/diag_tile.log 2>&1
done
cat $OUT/diag_tile.log
