#!/bin/bash
# Multi-GPU bench as the driver launches it, plus the reference arm.  Usage: scripts/gpu_multi.sh <tag> <ngpu>
TAG=$1; N=$2; OUT=gpurun_out/$TAG; mkdir -p $OUT
nvidia-smi -L > $OUT/smi.txt
python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29517 bench.py --gpus $N --steps 5 --warmup 3 > $OUT/bench_n$N.json 2> $OUT/bench_n$N.err; echo "bench N=$N exit $?"
tail -2 $OUT/bench_n$N.err; cat $OUT/bench_n$N.json | cut -c1-400
python bench.py --impl reference --gpus 1 --steps 2 --warmup 1 > $OUT/ref.json 2> $OUT/ref.err; echo "ref exit $?"; cat $OUT/ref.json | cut -c1-600
