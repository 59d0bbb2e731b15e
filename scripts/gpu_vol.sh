#!/usr/bin/env bash
set -u
TAG=${1:-v}
OUT=gpurun_out; mkdir -p $OUT
timeout 900 python -m pytest tests -m gpu -q --durations=5 > $OUT/${TAG}_pytest.log 2>&1
echo "pytest exit $?"; tail -25 $OUT/${TAG}_pytest.log
