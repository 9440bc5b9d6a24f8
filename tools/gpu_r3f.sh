#!/bin/bash
# r3f: where griddepcontrol.wait sits — AHK_PDL=1 before the first frame store, 3 at the end of the kernel, 2 nowhere (potential only)
set -u
OUT=gpurun_out; mkdir -p $OUT
run() { L=$1; shift
  env "$@" timeout 300 python bench.py --envs-per-gpu $E --steps 3000 --warmup 200 --no-cpu-baseline 2>/dev/null | python -c "
import sys,json
d=json.loads(sys.stdin.read()); print('$E $L', round(d['value']/1e6,2),'M/s', round(d['ms_per_step']*1e3,2),'us frac',round(d['roofline']['frac'],4), 'e2e', round(d['e2e']['value']/1e6,2))"
}
for E in 4096 8192 16384; do
  run "pdl=1" AHK_PDL=1
  run "pdl=3" AHK_PDL=3
  run "pdl=2" AHK_PDL=2
  run "pdl=3 balance=0" AHK_PDL=3 AHK_BALANCE=0
  run "pdl=1" AHK_PDL=1
  run "pdl=3" AHK_PDL=3
done | tee $OUT/pdl_wait_ab_r3f.txt
