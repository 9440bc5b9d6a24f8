set -u
run() { L=$1; shift
  env "$@" timeout 300 python bench.py --envs-per-gpu $E --steps 4000 --warmup 400 --no-cpu-baseline 2>/dev/null | python -c "
import sys,json
d=json.loads(sys.stdin.read()); print('$E $L', round(d['value']/1e6,2),'M/s', round(d['ms_per_step']*1e3,2),'us frac',round(d['roofline']['frac'],4))"
}
for E in 4096 8192; do
for A in 1 2 4 8 16 64; do run "adapt_every=$A" AHK_ADAPT_EVERY=$A; done
run "balance=0" AHK_BALANCE=0
done | tee gpurun_out/adapt_ab_r3s.txt
