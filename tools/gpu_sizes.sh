#!/bin/bash
# the step kernel at configs[2] / configs[3] per-GPU sizes
set -u
OUT=gpurun_out; mkdir -p $OUT
for N in 8192 65536 131072; do
  timeout 600 python bench.py --envs-per-gpu $N --steps 200 --warmup 10 --no-cpu-baseline > $OUT/bench_size_$N.json 2> $OUT/bench_size_$N.err; echo "rc=$? n=$N"
  python -c "
import json; d=json.load(open('$OUT/bench_size_$N.json')); print($N, 'value %.1fM ms %.4f achieved %.0f GB/s frac %.3f e2e %.1fM' % (d['value']/1e6, d['ms_per_step'], d['roofline']['whole_step']['achieved'], d['roofline']['whole_step']['frac'], d['e2e']['value']/1e6))"
done
