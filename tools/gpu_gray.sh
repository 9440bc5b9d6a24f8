#!/bin/bash
set -u
TAG=$1; OUT=gpurun_out; mkdir -p $OUT
timeout 900 python -m pytest tests/test_policy.py -m gpu -x -q > $OUT/pytest_gray_$TAG.log 2>&1; echo "pytest rc=$?"; tail -12 $OUT/pytest_gray_$TAG.log
timeout 300 python tools/gray_bench.py > $OUT/gray_$TAG.json 2> $OUT/gray_$TAG.err; echo "gray rc=$?"; cat $OUT/gray_$TAG.json; tail -3 $OUT/gray_$TAG.err
