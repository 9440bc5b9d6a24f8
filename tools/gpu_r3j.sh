#!/bin/bash
# r3j: device-side weight refresh, cross-stream ordering after early-return steps; RL loop bench
set -u
OUT=gpurun_out; mkdir -p $OUT
timeout 900 python -m pytest tests -m gpu -x -q > $OUT/pytest_gpu_r3j.log 2>&1; echo "suite rc=$?"; tail -n 4 $OUT/pytest_gpu_r3j.log
timeout 300 python tools/rl_bench.py > $OUT/rl_r3j.json 2> $OUT/rl_r3j.err; echo "rl rc=$?"; cat $OUT/rl_r3j.json; tail -n 2 $OUT/rl_r3j.err
