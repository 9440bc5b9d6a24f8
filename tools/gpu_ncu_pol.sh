#!/bin/bash
# ncu --set full capture of the policy kernels.  Usage: bash tools/gpu_ncu_pol.sh <tag> <kernel-regex> <count>
set -u
TAG=$1; KRE=$2; CNT=${3:-1}; OUT=gpurun_out; mkdir -p $OUT
timeout 300 python tools/policy_bench.py --envs 8192 --iters 3 --rollout-steps 2 > $OUT/plain_pol_$TAG.log 2>&1 &&
timeout 600 ncu --set full --clock-control none --import-source on -k regex:"$KRE" -s 6 -c $CNT -f -o $OUT/prof_policy_$TAG python tools/policy_bench.py --envs 8192 --iters 3 --rollout-steps 2 > $OUT/ncu_policy_full_$TAG.log 2>&1
echo "ncu full rc=$?"
