#!/bin/bash
# full suite + memcheck of the policy kernels + policy bench + rollout bench + launch list of the rollout
set -u
OUT=gpurun_out; mkdir -p $OUT
timeout 900 python -m pytest tests -m gpu -x -q > $OUT/pytest_gpu_r2a.log 2>&1; echo "pytest rc=$?"; tail -4 $OUT/pytest_gpu_r2a.log
timeout 900 compute-sanitizer --tool memcheck --error-exitcode 9 python -m pytest tests/test_policy.py -m gpu -x -q -k "37 or in_the_loop" > $OUT/memcheck_policy_r2a.log 2>&1; echo "memcheck rc=$?"; tail -4 $OUT/memcheck_policy_r2a.log
timeout 300 python tools/policy_bench.py --envs 8192 > $OUT/policy_r2a_8192.json 2> $OUT/policy_r2a.err; cat $OUT/policy_r2a_8192.json
timeout 300 python bench.py --workload rollout > $OUT/bench_rollout_r2a.json 2> $OUT/bench_rollout_r2a.err; echo "rollout rc=$?"; cut -c1-300 $OUT/bench_rollout_r2a.json
timeout 600 ncu --metrics gpu__time_duration.sum,sm__pipe_tensor_cycles_active.avg.pct_of_peak_sustained_active,dram__bytes_read.sum,dram__bytes_write.sum --clock-control none -s 40 -c 28 --csv --log-file $OUT/rollout_launches_r2a.csv python bench.py --workload rollout --steps 20 --warmup 3 > $OUT/ncu_rollout_r2a.log 2>&1; echo "ncu list rc=$?"
timeout 600 ncu --set full --clock-control none --import-source on -k regex:"conv1_kernel|conv_ws_kernel|layer_kernel" -s 12 -c 6 -f -o $OUT/prof_policy_r2a python tools/policy_bench.py --envs 8192 --iters 3 --rollout-steps 2 > $OUT/ncu_policy_full_r2a.log 2>&1; echo "ncu full rc=$?"
