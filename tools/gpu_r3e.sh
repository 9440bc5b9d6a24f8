#!/bin/bash
# r3e: A/B of the SM-keyed adaptive partition under overlapped launches (AHK_SMKEY), with dynamic unit claiming on
set -u
OUT=gpurun_out; mkdir -p $OUT
timeout 600 python -m pytest tests/test_env_gpu.py -m gpu -x -q > $OUT/pytest_smkey_r3e.log 2>&1; echo "parity rc=$?"; tail -n 3 $OUT/pytest_smkey_r3e.log
run() { # label, env assignments...
  L=$1; shift
  env "$@" timeout 300 python bench.py --envs-per-gpu $E --steps 3000 --warmup 200 --no-cpu-baseline 2>/dev/null | python -c "
import sys,json
d=json.loads(sys.stdin.read()); print('$E $L', round(d['value']/1e6,2),'M/s', round(d['ms_per_step']*1e3,2),'us frac',round(d['roofline']['frac'],4), 'e2e', round(d['e2e']['value']/1e6,2))"
}
for E in 4096 8192 16384 65536; do
  run "smkey=1" AHK_SMKEY=1
  run "smkey=0" AHK_SMKEY=0
  run "smkey=1 pdl=1" AHK_SMKEY=1 AHK_PDL=1
  run "smkey=1 pdl=0" AHK_SMKEY=1 AHK_PDL=0
  run "balance=0 pdl=1" AHK_BALANCE=0 AHK_PDL=1
done | tee $OUT/smkey_ab_r3e.txt
python tools/phase_trace.py > $OUT/phase_trace_r3e.json; python tools/overlap_trace.py > $OUT/overlap_trace_r3e.json
