#!/bin/bash
# One gpurun call, assembled from arguments (replaces the per-experiment gpu_r*.sh scripts of round 1).
#   bash tools/gpu_round.sh TAG [part ...]
# parts: tests | tests:<pytest -k expr> | bench | driver | long | ref | stress | stress8 | ncu:<kernel regex> | policy | rl | rl8 | scale8
# Everything is written to gpurun_out/ with TAG in the name.  ncu parts run only after the same command exited 0 plainly.
set -u
TAG=${1:-r2}; shift || true
OUT=gpurun_out
mkdir -p $OUT
nvidia-smi --query-gpu=index,name,clocks.sm,clocks.max.sm,power.draw --format=csv > $OUT/smi_$TAG.txt 2>&1
for PART in "$@"; do
  case $PART in
    tests)   timeout 2400 python -m pytest tests -m gpu -x -q > $OUT/pytest_gpu_$TAG.log 2>&1; echo "pytest rc=$?" | tee -a $OUT/pytest_gpu_$TAG.log; tail -15 $OUT/pytest_gpu_$TAG.log ;;
    tests:*) timeout 1800 python -m pytest tests -m gpu -x -q -k "${PART#tests:}" > $OUT/pytest_k_$TAG.log 2>&1; echo "pytest -k rc=$?" | tee -a $OUT/pytest_k_$TAG.log; tail -15 $OUT/pytest_k_$TAG.log ;;
    bench)   timeout 900 python bench.py > $OUT/bench_$TAG.json 2> $OUT/bench_$TAG.err; echo "bench rc=$?"; cat $OUT/bench_$TAG.json; tail -3 $OUT/bench_$TAG.err ;;
    driver)  timeout 900 python bench.py --gpus 1 --steps 20 --warmup 5 > $OUT/bench_driver_$TAG.json 2> $OUT/bench_driver_$TAG.err; echo "driver-style bench rc=$?"; cat $OUT/bench_driver_$TAG.json ;;
    long)    timeout 900 python bench.py --steps 400 --warmup 5 --no-rollout --no-cpu-baseline > $OUT/bench_long_$TAG.json 2> $OUT/bench_long_$TAG.err; echo "long bench rc=$?"; cat $OUT/bench_long_$TAG.json ;;
    ref)     timeout 600 python bench.py --impl reference --steps 20 --warmup 5 > $OUT/bench_ref_$TAG.json 2> $OUT/bench_ref_$TAG.err; echo "bench ref rc=$?"; cat $OUT/bench_ref_$TAG.json ;;
    stress)  timeout 1200 python tools/stress_sweep.py --envs 384 --horizon 10000 --out $OUT/stress_sweep_N1.json 2> $OUT/stress_$TAG.err; echo "stress rc=$?" ;;
    stress8) for N in 2 4 8; do timeout 1200 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 2953$N tools/stress_sweep.py --envs 384 --horizon 10000 --out $OUT/stress_sweep_N$N.json 2> $OUT/stress_N${N}_$TAG.err; echo "stress N=$N rc=$?"; done ;;
    scale8)  for N in 2 4 8; do timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 2954$N bench.py --gpus $N --steps 20 --warmup 5 > $OUT/bench_${N}gpu_$TAG.json 2> $OUT/bench_${N}gpu_$TAG.err; echo "bench N=$N rc=$?"; cat $OUT/bench_${N}gpu_$TAG.json; done ;;
    policy)  timeout 600 python tools/policy_bench.py > $OUT/policy_$TAG.json 2> $OUT/policy_$TAG.err; echo "policy rc=$?"; cat $OUT/policy_$TAG.json ;;
    rl8)     for N in 2 8; do timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 2955$N tools/rl_bench.py > $OUT/rl_${N}gpu_$TAG.json 2> $OUT/rl_${N}gpu_$TAG.err; echo "rl N=$N rc=$?"; cat $OUT/rl_${N}gpu_$TAG.json; done ;;
    rl)      timeout 900 python tools/rl_bench.py > $OUT/rl_$TAG.json 2> $OUT/rl_$TAG.err; echo "rl rc=$?"; cat $OUT/rl_$TAG.json ;;
    ncu:*)   KRE=${PART#ncu:}
             CMD="python bench.py --steps 2 --warmup 3 --chunk 16 --no-cpu-baseline --no-rollout"
             case $KRE in conv*|layer*|gray*) CMD="python tools/policy_bench.py --iters 3 --rollout-steps 4" ;; esac
             timeout 300 $CMD > $OUT/plain_$TAG.log 2>&1 &&
             timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file $OUT/launches_$TAG.csv $CMD > $OUT/ncu_list_$TAG.log 2>&1
             echo "ncu list rc=$?"
             timeout 300 $CMD > $OUT/plain2_$TAG.log 2>&1 &&
             timeout 1200 ncu --set full --clock-control none --import-source on -k regex:$KRE -s 10 -c 3 -o $OUT/prof_$TAG $CMD > $OUT/ncu_full_$TAG.log 2>&1
             echo "ncu full rc=$?" ;;
    *) echo "unknown part $PART" ;;
  esac
done
ls -la $OUT | tail -30
