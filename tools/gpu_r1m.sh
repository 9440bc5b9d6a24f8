#!/bin/bash
# r1m: zero-copy host step + Toeplitz conv1: tests, bench, policy bench (both conv1 paths), launch list
set -u
OUT=gpurun_out; mkdir -p $OUT
timeout 900 python -m pytest tests -m gpu -x -q > $OUT/pytest_gpu_r1m.log 2>&1; echo "pytest rc=$?"; tail -15 $OUT/pytest_gpu_r1m.log
timeout 600 python bench.py --no-cpu-baseline --steps 500 > $OUT/bench_r1m.json 2> $OUT/bench_r1m.err; echo "bench rc=$?"; cat $OUT/bench_r1m.json; tail -3 $OUT/bench_r1m.err


timeout 300 python tools/policy_bench.py --envs 8192 > $OUT/policy_r1m_8192.json 2> $OUT/policy_r1m.err; echo "policy rc=$?"; cat $OUT/policy_r1m_8192.json; tail -3 $OUT/policy_r1m.err

timeout 600 ncu --metrics gpu__time_duration.sum,sm__pipe_tensor_cycles_active.avg.pct_of_peak_sustained_active,dram__bytes_read.sum,dram__bytes_write.sum --clock-control none -k regex:"layer_kernel|conv1_kernel" -s 24 -c 6 --csv --log-file $OUT/policy_launches_r1m.csv python tools/policy_bench.py --envs 8192 --iters 5 --rollout-steps 4 > $OUT/ncu_policy_r1m.log 2>&1
echo "ncu policy rc=$?"
timeout 600 ncu --set full --clock-control none --import-source on -k regex:"conv1_kernel|layer_kernel" -s 6 -c 6 -o $OUT/prof_policy_r1m python tools/policy_bench.py --envs 8192 --iters 3 --rollout-steps 2 > $OUT/ncu_policy_full_r1m.log 2>&1
echo "ncu full rc=$?"
