#!/bin/bash
# GPU tests, then every build under build/variants on the bench workload.  Usage: scripts/gpu_x2.sh <tag> [bench args]
TAG=$1; shift; OUT=gpurun_out/$TAG; mkdir -p $OUT
timeout 900 python -m pytest tests -m gpu -x -q > $OUT/pytest.log 2>&1; echo "pytest exit $?"; tail -5 $OUT/pytest.log
bash scripts/gpu_variants.sh $TAG "$@"
