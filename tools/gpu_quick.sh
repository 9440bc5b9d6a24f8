#!/bin/bash
# GPU parity tests + short A/B bench lines.  Usage: bash tools/gpu_quick.sh <tag>
TAG=$1; OUT=gpurun_out; mkdir -p $OUT
timeout 600 python -m pytest tests -m gpu -x -q > $OUT/pytest_gpu_$TAG.log 2>&1; echo "pytest rc=$?"; tail -4 $OUT/pytest_gpu_$TAG.log
timeout 300 python bench.py --no-cpu-baseline > $OUT/bench_$TAG.json 2> $OUT/bench_$TAG.err; tail -3 $OUT/bench_$TAG.err
AHK_PATH=split timeout 300 python bench.py --no-cpu-baseline > $OUT/bench_${TAG}_split.json 2>> $OUT/bench_$TAG.err
python - <<PY
import json
for f in ("bench_$TAG.json", "bench_${TAG}_split.json"):
    try:
        d = json.load(open("$OUT/" + f)); r = d["roofline"]
        print(f, "value %.1fM  ms/step %.4f  kernel %.4f ms  env %.4f ms  achieved %.0f GB/s  e2e %.1fM" % (d["value"]/1e6, d["ms_per_step"], r["kernel_ms"], r["env_kernel_ms"], r["achieved"], d["e2e"]["value"]/1e6))
    except Exception as e:
        print(f, "unreadable", e)
PY
