#!/bin/bash
# r3d: A/B of intra-CTA dynamic unit claiming (AHK_UNITS=dynamic) against round-robin, same box
set -u
OUT=gpurun_out; mkdir -p $OUT
AHK_UNITS=dynamic timeout 600 python -m pytest tests/test_env_gpu.py -m gpu -x -q > $OUT/pytest_dyn_r3d.log 2>&1; echo "parity(dynamic) rc=$?"; tail -n 3 $OUT/pytest_dyn_r3d.log
for E in 4096 16384 65536; do
for M in static dynamic static dynamic; do
  AHK_UNITS=$M timeout 300 python bench.py --envs-per-gpu $E --steps 1500 --warmup 50 --no-cpu-baseline 2>/dev/null | python -c "
import sys,json
d=json.loads(sys.stdin.read()); print('$E $M', round(d['value']/1e6,2),'M/s', round(d['ms_per_step']*1e3,2),'us frac',round(d['roofline']['frac'],4), 'e2e', round(d['e2e']['value']/1e6,2))"
done; done | tee $OUT/units_ab_r3d.txt
