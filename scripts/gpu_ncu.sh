#!/bin/bash
# ncu --set full on selected kernels.  Usage: scripts/gpu_ncu.sh <tag> <kernel regex> <count> [bench args]
TAG=$1; RE=$2; CNT=$3; shift 3
OUT=gpurun_out/$TAG; mkdir -p $OUT
python bench.py "$@" > $OUT/plain.log 2>&1 &&
ncu --set full --clock-control none --import-source on -k regex:"$RE" -c $CNT -o $OUT/prof python bench.py "$@" > $OUT/ncu.log 2>&1
echo "ncu exit $?"; tail -3 $OUT/ncu.log
