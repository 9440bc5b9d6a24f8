#!/bin/bash
# r3h: early-return host step: parity tests, then the default bench
set -u
OUT=gpurun_out; mkdir -p $OUT
timeout 600 python -m pytest tests/test_env_gpu.py -m gpu -x -q -k "step_host" > $OUT/pytest_host_r3h.log 2>&1; echo "host tests rc=$?"; tail -n 5 $OUT/pytest_host_r3h.log
timeout 600 python bench.py > $OUT/bench_r3h.json 2> $OUT/bench_r3h.err; echo "bench rc=$?"; cat $OUT/bench_r3h.json; tail -n 3 $OUT/bench_r3h.err
