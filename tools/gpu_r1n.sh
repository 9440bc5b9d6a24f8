#!/bin/bash
set -u
OUT=gpurun_out; mkdir -p $OUT
timeout 900 python -m pytest tests/test_policy.py -m gpu -x -q > $OUT/pytest_gpu_r1n.log 2>&1; echo "pytest rc=$?"; tail -5 $OUT/pytest_gpu_r1n.log
timeout 300 python tools/policy_bench.py --envs 8192 > $OUT/policy_r1n_8192.json 2> $OUT/policy_r1n.err; echo "policy rc=$?"; cat $OUT/policy_r1n_8192.json; tail -3 $OUT/policy_r1n.err
