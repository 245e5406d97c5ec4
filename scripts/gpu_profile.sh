#!/bin/bash
# One gpurun call: GPU tests, bench lines of every BASELINE workload, latency of single images, ncu launch list + full
# captures.  Usage: scripts/gpu_profile.sh <tag> [phases]   phases: any of t b l n x (default all but x)
# (compute-sanitizer is closed on this pool by its operators: no sanitizer phase)
#   gpurun --timeout 2400 -- 'bash scripts/gpu_profile.sh r02a'
TAG=${1:-run}
PH=${2:-tbln}
OUT=gpurun_out/$TAG
mkdir -p $OUT
nvidia-smi --query-gpu=name,clocks.sm,clocks.max.sm,power.draw --format=csv > $OUT/smi.txt 2>&1
nproc >> $OUT/smi.txt; free -g | head -2 >> $OUT/smi.txt
if [[ $PH == *t* ]]; then
  timeout 1500 python -m pytest tests -m gpu -q -p no:cacheprovider > $OUT/pytest.log 2>&1; echo "pytest exit $?" | tee -a $OUT/pytest.log
  tail -15 $OUT/pytest.log
fi
if [[ $PH == *b* ]]; then
  for W in C2 C3 C4-gray C4-422 C4-444 C5; do
    timeout 600 python bench.py --workload $W > $OUT/bench_$W.json 2> $OUT/bench_$W.err; echo "bench $W exit $?"; cut -c1-600 $OUT/bench_$W.json
  done
  timeout 300 python bench.py --workload C2 --entropy sync > $OUT/bench_C2_sync.json 2> $OUT/bench_C2_sync.err; echo "bench C2 sync exit $?"; cut -c1-400 $OUT/bench_C2_sync.json
  for W in C4-gray C4-422 C4-444; do
    timeout 300 python bench.py --workload $W --quality 50 > $OUT/bench_${W}_q50.json 2> $OUT/bench_${W}_q50.err; echo "bench $W q50 exit $?"
  done
fi
if [[ $PH == *l* ]]; then
  timeout 300 python scripts/latency.py > $OUT/latency.jsonl 2> $OUT/latency.err; echo "latency exit $?"; cat $OUT/latency.jsonl
fi
if [[ $PH == *n* ]]; then
  timeout 300 python bench.py --steps 2 --warmup 1 > $OUT/plain.log 2>&1 &&
  timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -c 600 --csv --log-file $OUT/launches.csv \
      python bench.py --steps 2 --warmup 1 --skip-parity > $OUT/ncu_launches.log 2>&1
  echo "ncu launches exit $?"
  timeout 600 ncu --set full --clock-control none --import-source on -k regex:'k_pixel|k_entropy_seg|k_unstuff_(count|scatter)' -c 8 \
      -o $OUT/prof_C2 python bench.py --steps 1 --warmup 1 --skip-parity > $OUT/ncu_full_C2.log 2>&1
  echo "ncu full C2 exit $?"
  timeout 600 ncu --set full --clock-control none --import-source on -k regex:'k_xsync|k_entropy_seg|k_xunits|k_pixel' -c 8 \
      -o $OUT/prof_C3 python bench.py --workload C3 --steps 1 --warmup 1 --skip-parity > $OUT/ncu_full_C3.log 2>&1
  echo "ncu full C3 exit $?"
  timeout 600 ncu --set full --clock-control none --import-source on -k regex:'k_xsync|k_entropy_seg' -c 8 \
      -o $OUT/prof_C2sync python bench.py --entropy sync --steps 1 --warmup 1 --skip-parity > $OUT/ncu_full_C2sync.log 2>&1
  echo "ncu full C2 sync exit $?"
  for P in C2 C3 C2sync; do
    ncu -i $OUT/prof_$P.ncu-rep --page raw --csv > $OUT/raw_$P.csv 2>/dev/null
  done
fi
if [[ $PH == *x* ]]; then
  B200JPEG_LIB=build/variants/tuning.so timeout 900 python scripts/dev/xsweep.py C2 C3 C5 > $OUT/xsweep.jsonl 2> $OUT/xsweep.err
  echo "xsweep exit $?"; tail -3 $OUT/xsweep.err
fi
ls -la $OUT
