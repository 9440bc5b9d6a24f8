#!/bin/bash
# ncu --set full capture of one kernel of bench.py.  Usage: bash tools/gpu_ncu.sh <tag> <kernel-regex> [env assignments...]
set -u
TAG=$1; KRE=$2; shift 2
OUT=gpurun_out
mkdir -p $OUT
CMD="python bench.py --steps 40 --warmup 5 --no-cpu-baseline"
env "$@" timeout 300 $CMD > $OUT/plain_$TAG.log 2>&1 &&
env "$@" timeout 900 ncu --set full --clock-control none --import-source on -k regex:$KRE -s 10 -c 3 -o $OUT/prof_$TAG $CMD > $OUT/ncu_$TAG.log 2>&1
echo "ncu rc=$?"; tail -3 $OUT/ncu_$TAG.log
