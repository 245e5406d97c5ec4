#!/bin/bash
# One multi-GPU gpurun call (charged N x): PCIe probe, the one-process multi-device entry point, bench lines at N GPUs.
#   gpurun --gpus 8 --timeout 1200 -- 'bash scripts/gpu_multi.sh r02m 8'
TAG=${1:-multi}; N=${2:-8}
OUT=gpurun_out/$TAG; mkdir -p $OUT
nvidia-smi --query-gpu=index,name --format=csv > $OUT/smi.txt; nproc >> $OUT/smi.txt; free -g | head -2 >> $OUT/smi.txt
nvidia-smi topo -m > $OUT/topo.txt 2>&1
timeout 200 build/pcie_probe $N > $OUT/pcie_probe.jsonl 2> $OUT/pcie_probe.err; echo "probe exit $?"; cat $OUT/pcie_probe.jsonl
timeout 300 python scripts/dev/multi_e2e.py $N > $OUT/multi_e2e.jsonl 2> $OUT/multi_e2e.err; echo "multi_e2e exit $?"; cat $OUT/multi_e2e.jsonl; tail -2 $OUT/multi_e2e.err
run() { W=$1; G=$2; timeout 400 python -m torch.distributed.run --nnodes=1 --nproc-per-node $G --master-addr 127.0.0.1 --master-port $((29500 + G)) \
          bench.py --gpus $G --workload $W > $OUT/bench_${W}_n$G.json 2> $OUT/bench_${W}_n$G.err; echo "bench $W n$G exit $?"; cut -c1-260 $OUT/bench_${W}_n$G.json; }
run C2 $N
for G in 2 4 8; do [ $G -le $N ] && run C5 $G; done
ls $OUT
