#!/bin/bash
# configs[2]: 1 M envs over 8 GPUs (131 072 per GPU), datagen workload
set -u
OUT=gpurun_out; mkdir -p $OUT
N=$(nvidia-smi -L | wc -l); echo "gpus=$N"
timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29531 bench.py --gpus $N --envs-per-gpu 131072 --steps 400 --warmup 20 > $OUT/bench_cfg2_${N}gpu_r3o.json 2> $OUT/bench_cfg2_${N}gpu_r3o.err; echo "rc=$?"; cut -c1-700 $OUT/bench_cfg2_${N}gpu_r3o.json; tail -n 3 $OUT/bench_cfg2_${N}gpu_r3o.err
