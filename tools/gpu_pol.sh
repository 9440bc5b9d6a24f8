#!/bin/bash
# quick policy iteration: parity tests + per-layer timings.  Usage: bash tools/gpu_pol.sh <tag>
set -u
TAG=$1; OUT=gpurun_out; mkdir -p $OUT
timeout 900 python -m pytest tests/test_policy.py -m gpu -x -q > $OUT/pytest_pol_$TAG.log 2>&1; echo "pytest rc=$?"; tail -5 $OUT/pytest_pol_$TAG.log
timeout 300 python tools/policy_bench.py --envs 8192 > $OUT/policy_${TAG}_8192.json 2> $OUT/policy_$TAG.err; echo "policy rc=$?"; cat $OUT/policy_${TAG}_8192.json; tail -3 $OUT/policy_$TAG.err
