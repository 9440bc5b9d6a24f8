#!/bin/bash
# r1u: full GPU suite, rollout bench, 2-GPU torchrun of both workloads
set -u
OUT=gpurun_out; mkdir -p $OUT
timeout 900 python -m pytest tests -m gpu -x -q > $OUT/pytest_gpu_r1u.log 2>&1; echo "pytest rc=$?"; tail -6 $OUT/pytest_gpu_r1u.log
timeout 300 python bench.py --workload rollout > $OUT/bench_rollout_r1u.json 2> $OUT/bench_rollout_r1u.err; echo "rollout rc=$?"; cat $OUT/bench_rollout_r1u.json; tail -3 $OUT/bench_rollout_r1u.err
if [ "$(nvidia-smi -L | wc -l)" -ge 2 ]; then
timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29511 bench.py --gpus 2 --steps 1000 --warmup 50 > $OUT/bench_2gpu_r1u.json 2> $OUT/bench_2gpu_r1u.err; echo "2gpu rc=$?"; cat $OUT/bench_2gpu_r1u.json | cut -c1-400; tail -3 $OUT/bench_2gpu_r1u.err
timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29512 bench.py --gpus 2 --workload rollout > $OUT/bench_rollout_2gpu_r1u.json 2> $OUT/bench_rollout_2gpu_r1u.err; echo "2gpu rollout rc=$?"; cat $OUT/bench_rollout_2gpu_r1u.json | cut -c1-400; tail -3 $OUT/bench_rollout_2gpu_r1u.err
timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29513 bench.py --impl reference --gpus 2 --steps 3 --warmup 1 > $OUT/bench_ref_2gpu_r1u.json 2>> $OUT/bench_2gpu_r1u.err; echo "2gpu ref rc=$?"; cat $OUT/bench_ref_2gpu_r1u.json | cut -c1-300
fi
