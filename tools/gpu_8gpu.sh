#!/bin/bash
# r3n: 8-GPU torchrun of both workloads (the driver's scaling run, confirmed ahead of round end)
set -u
OUT=gpurun_out; mkdir -p $OUT
N=$(nvidia-smi -L | wc -l); echo "gpus=$N"
for W in datagen rollout; do
  EXTRA=""; [ $W = rollout ] && EXTRA="--workload rollout"
  timeout 400 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 2951$((RANDOM%10)) bench.py --gpus $N $EXTRA > $OUT/bench_${W}_${N}gpu_r3n.json 2> $OUT/bench_${W}_${N}gpu_r3n.err; echo "$W rc=$?"; cut -c1-600 $OUT/bench_${W}_${N}gpu_r3n.json; tail -n 3 $OUT/bench_${W}_${N}gpu_r3n.err
done
