#!/bin/bash
# Bench lines only, at N GPUs of one box (charged N x): scripts/gpu_multi_bench.sh <tag> <N> <workload>...
#   gpurun --gpus 8 --timeout 900 -- 'bash scripts/gpu_multi_bench.sh r02n 8 C2 C5'
TAG=$1; N=$2; shift 2
OUT=gpurun_out/$TAG; mkdir -p $OUT
for W in "$@"; do
  timeout 400 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port $((29500 + N)) \
    bench.py --gpus $N --workload $W > $OUT/bench_${W}_n$N.json 2> $OUT/bench_${W}_n$N.err; echo "bench $W n$N exit $?"; cut -c1-200 $OUT/bench_${W}_n$N.json
done
