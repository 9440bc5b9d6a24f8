#!/bin/bash
# r3k: e2e prologue: actions fetched with the state, system fence by one thread of the last warp; then 2-GPU torchrun of every arm
set -u
OUT=gpurun_out; mkdir -p $OUT
timeout 600 python -m pytest tests/test_env_gpu.py -m gpu -x -q > $OUT/pytest_env_r3k.log 2>&1; echo "env tests rc=$?"; tail -n 2 $OUT/pytest_env_r3k.log
for i in 1 2; do timeout 300 python bench.py --steps 3000 --warmup 100 --no-cpu-baseline 2>/dev/null | python -c "
import sys,json
d=json.loads(sys.stdin.read()); e=d['e2e']; print('4096', round(d['value']/1e6,2),'M/s frac',round(d['roofline']['frac'],4), 'e2e', round(e['value']/1e6,2), 'full_wait', round(e['full_wait']['value']/1e6,2), 'piped', round(e['pipelined']['value']/1e6,2))"; done | tee $OUT/e2e_r3k.txt
if [ "$(nvidia-smi -L | wc -l)" -ge 2 ]; then
N=2
timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29521 bench.py --gpus $N > $OUT/bench_${N}gpu_r3k.json 2> $OUT/bench_${N}gpu_r3k.err; echo "2gpu rc=$?"; cut -c1-300 $OUT/bench_${N}gpu_r3k.json; tail -n 2 $OUT/bench_${N}gpu_r3k.err
timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29522 bench.py --gpus $N --workload rollout > $OUT/bench_rollout_${N}gpu_r3k.json 2>> $OUT/bench_${N}gpu_r3k.err; echo "2gpu rollout rc=$?"; cut -c1-300 $OUT/bench_rollout_${N}gpu_r3k.json
timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29523 bench.py --impl reference --gpus $N --steps 3 --warmup 1 > $OUT/bench_ref_${N}gpu_r3k.json 2>> $OUT/bench_${N}gpu_r3k.err; echo "2gpu ref rc=$?"; cut -c1-200 $OUT/bench_ref_${N}gpu_r3k.json
fi
