#!/usr/bin/env bash
# The GPU test files that scripts/ab_prefetch.sh does not run + smoke + a short bench:  bash scripts/gpu_rest.sh <tag>
set -u
OUT=gpurun_out; mkdir -p $OUT; TAG=${1:-rest}
timeout 240 python -m pytest tests/test_zz_no_alloc.py tests/test_gpu_masks.py tests/test_gpu_pipeline.py tests/test_gpu_group.py tests/test_gpu_occlusion.py \
  tests/test_gpu_particles.py tests/test_gpu_extras.py tests/test_gpu_shim.py -m gpu -q > $OUT/${TAG}_pytest.log 2>&1; echo "pytest exit $?"; tail -3 $OUT/${TAG}_pytest.log
timeout 100 python -c "import __graft_entry__ as g; g.smoke()" > $OUT/${TAG}_smoke.log 2>&1; echo "smoke exit $?"; tail -2 $OUT/${TAG}_smoke.log
timeout 200 python bench.py --steps 50 --warmup 5 --no-cpu --no-extras > $OUT/${TAG}_bench.json 2> $OUT/${TAG}_bench.err; echo "bench exit $?"
python - $OUT/${TAG}_bench.json <<'PY'
import json, sys
d = json.loads(open(sys.argv[1]).read().strip().splitlines()[-1])
r = d["roofline"]
print("ms", round(d["ms_per_step"], 4), "kernel_ms", round(r["kernel_ms"], 4), "frac", round(r["frac"], 4), "step_frac", round(r["step_frac"], 4),
      "e2e_ms", round(d["e2e"]["ms_per_step"], 4), "parity", d["parity"]["lists_equal"], d["parity"]["world_trs_bit_exact"], d["parity"]["matrices_within_1e-5"])
PY
