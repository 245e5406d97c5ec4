#!/bin/bash
# One gpurun call: GPU tests, bench, ncu launch list, ncu full capture.  Usage: scripts/gpu_profile.sh <tag>
# (run as: gpurun --timeout 1500 -- 'bash scripts/gpu_profile.sh r01b')
TAG=${1:-run}
OUT=gpurun_out/$TAG
mkdir -p $OUT
nvidia-smi --query-gpu=name,clocks.sm,clocks.max.sm,power.draw --format=csv > $OUT/smi.txt 2>&1
python -m pytest tests -m gpu -x -q > $OUT/pytest.log 2>&1; echo "pytest exit $?" | tee -a $OUT/pytest.log
tail -5 $OUT/pytest.log
python bench.py > $OUT/bench.json 2> $OUT/bench.err; echo "bench exit $?"
cat $OUT/bench.json
python bench.py --steps 2 --warmup 1 > $OUT/plain.log 2>&1 &&
ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file $OUT/launches.csv \
    python bench.py --steps 2 --warmup 1 > $OUT/ncu_launches.log 2>&1
echo "ncu launches exit $?"
python bench.py --steps 1 --warmup 1 > $OUT/plain2.log 2>&1 &&
ncu --set full --clock-control none --import-source on -k regex:'k_pixel|k_entropy|k_unstuff_(count|scatter)|k_build_luts' -c 10 \
    -o $OUT/prof python bench.py --steps 1 --warmup 1 > $OUT/ncu_full.log 2>&1
echo "ncu full exit $?"
ls -la $OUT
