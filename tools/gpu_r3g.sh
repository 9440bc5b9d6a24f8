#!/bin/bash
# r3g: deferred-wait launches as the default: whole GPU suite (twice: timing-dependent hazards), then the size sweep
set -u
OUT=gpurun_out; mkdir -p $OUT
timeout 900 python -m pytest tests -m gpu -x -q > $OUT/pytest_gpu_r3g.log 2>&1; echo "suite rc=$?"; tail -n 3 $OUT/pytest_gpu_r3g.log
for i in 1 2 3; do timeout 600 python -m pytest tests/test_env_gpu.py -m gpu -x -q -k "overlap or full_size or lockstep" 2>&1 | tail -n 1; done
run() { L=$1; shift
  env "$@" timeout 300 python bench.py --envs-per-gpu $E --steps 3000 --warmup 200 --no-cpu-baseline 2>/dev/null | python -c "
import sys,json
d=json.loads(sys.stdin.read()); print('$E $L', round(d['value']/1e6,2),'M/s', round(d['ms_per_step']*1e3,2),'us frac',round(d['roofline']['frac'],4), 'e2e', round(d['e2e']['value']/1e6,2), 'piped', round(d['e2e'].get('pipelined',{}).get('value',0)/1e6,2))"
}
for E in 4096 8192; do
  run "default" A=1
  run "pdl=1" AHK_PDL=1
  run "pdl=0" AHK_PDL=0
done | tee $OUT/pdl3_sweep_r3g.txt
