#!/bin/bash
# r3b: RL-step tests, then the whole GPU suite
set -u
OUT=gpurun_out; mkdir -p $OUT
timeout 600 python -m pytest tests/test_rl.py -m gpu -x -q > $OUT/pytest_rl_r3b.log 2>&1; echo "rl rc=$?"; tail -n 30 $OUT/pytest_rl_r3b.log
timeout 900 python -m pytest tests -m gpu -q > $OUT/pytest_gpu_r3b.log 2>&1; echo "all rc=$?"; tail -n 8 $OUT/pytest_gpu_r3b.log
