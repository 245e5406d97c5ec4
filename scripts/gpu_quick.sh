#!/bin/bash
# GPU tests + one bench line.  Usage: scripts/gpu_quick.sh <tag> [bench args]
TAG=${1:-quick}; shift
OUT=gpurun_out/$TAG; mkdir -p $OUT
python -m pytest tests -m gpu -x -q > $OUT/pytest.log 2>&1; echo "pytest exit $?"; tail -15 $OUT/pytest.log
python bench.py "$@" > $OUT/bench.json 2> $OUT/bench.err; echo "bench exit $?"; tail -3 $OUT/bench.err
cat $OUT/bench.json
