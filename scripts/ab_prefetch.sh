#!/usr/bin/env bash
# A/B of the tile kernel's spine-row prefetch (RN_TILE_PREFETCH_SPINE=0 / 1) + the scene and edit parity tests
set -u
OUT=gpurun_out; mkdir -p $OUT; TAG=${1:-ab}
ARGS="--steps 50 --warmup 5 --no-cpu --no-extras --emitters 0"
RN_TILE_PREFETCH_SPINE=0 timeout 300 python bench.py $ARGS > $OUT/${TAG}_off.json 2> $OUT/${TAG}_off.err; echo "off exit $?"
timeout 300 python bench.py $ARGS > $OUT/${TAG}_on.json 2> $OUT/${TAG}_on.err; echo "on exit $?"
timeout 300 python bench.py $ARGS --config C2 > $OUT/${TAG}_on_C2.json 2> $OUT/${TAG}_on_C2.err; echo "C2 exit $?"
for f in off on on_C2; do python - $OUT/${TAG}_$f.json <<'PY'
import json, sys
d = json.loads(open(sys.argv[1]).read().strip().splitlines()[-1])
r = d["roofline"]
print(sys.argv[1], "ms", round(d["ms_per_step"], 4), "kernel_ms", round(r["kernel_ms"], 4), "frac", round(r["frac"], 4), "step_frac", round(r["step_frac"], 4),
      "e2e_ms", round(d["e2e"]["ms_per_step"], 4), "parity", d["parity"]["lists_equal"], d["parity"]["world_trs_bit_exact"], d["parity"]["matrices_within_1e-5"])
PY
done
timeout 900 python -m pytest tests/test_gpu_scene.py tests/test_gpu_edits.py -m gpu -q -x > $OUT/${TAG}_pytest.log 2>&1; echo "pytest exit $?"; tail -3 $OUT/${TAG}_pytest.log
