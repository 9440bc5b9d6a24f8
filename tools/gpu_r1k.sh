#!/bin/bash
# r1k: re-validate HEAD (tests, bench, reference arm), first policy/rollout numbers + their ncu launch list
set -u
OUT=gpurun_out; mkdir -p $OUT
timeout 900 python -m pytest tests -m gpu -x -q > $OUT/pytest_gpu_r1k.log 2>&1; echo "pytest rc=$?"; tail -4 $OUT/pytest_gpu_r1k.log
timeout 600 python bench.py > $OUT/bench_r1k.json 2> $OUT/bench_r1k.err; echo "bench rc=$?"; cat $OUT/bench_r1k.json
timeout 300 python tools/policy_bench.py --envs 8192 > $OUT/policy_r1k_8192.json 2> $OUT/policy_r1k.err; echo "policy rc=$?"; cat $OUT/policy_r1k_8192.json
timeout 300 python tools/policy_bench.py --envs 4096 > $OUT/policy_r1k_4096.json 2>> $OUT/policy_r1k.err; cat $OUT/policy_r1k_4096.json
timeout 600 ncu --metrics gpu__time_duration.sum,sm__pipe_tensor_cycles_active.avg.pct_of_peak_sustained_active,dram__bytes_read.sum,dram__bytes_write.sum --clock-control none -k regex:layer_kernel -s 24 -c 12 --csv --log-file $OUT/policy_launches_r1k.csv python tools/policy_bench.py --envs 8192 --iters 5 --rollout-steps 4 > $OUT/ncu_policy_r1k.log 2>&1
echo "ncu policy rc=$?"; tail -3 $OUT/policy_r1k.err
