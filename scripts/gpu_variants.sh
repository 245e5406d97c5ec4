#!/bin/bash
# Bench several builds of the library (same sources, different -D tuning flags) in one gpurun call.
# Usage: scripts/gpu_variants.sh <tag> [bench args]   -- variants are the .so files under build/variants/
TAG=$1; shift
OUT=gpurun_out/$TAG; mkdir -p $OUT
for so in build/variants/*.so; do
  name=$(basename $so .so)
  B200JPEG_LIB=$PWD/$so python bench.py "$@" > $OUT/$name.json 2> $OUT/$name.err || echo "$name FAILED"
  python - $OUT/$name.json $name <<'PY'
import json,sys
try:
    d=json.load(open(sys.argv[1])); print(sys.argv[2], round(d['value']), round(d['event_ms_per_step'],3), {k:round(v['ms'],3) for k,v in d['kernels'].items()})
except Exception as e: print(sys.argv[2], 'no result', e)
PY
done
