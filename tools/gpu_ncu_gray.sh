#!/bin/bash
# ncu --set full of the 'gray-diff' network's kernels (one forward pass: input, conv1 tile kernel, conv2-4 pipelines, dense, out)
set -u
TAG=$1; OUT=gpurun_out; mkdir -p $OUT
timeout 300 python tools/gray_bench.py > $OUT/plain_gray_$TAG.log 2>&1 &&
timeout 900 ncu --set full --clock-control none --import-source on -k regex:"gray_diff_kernel|layer_kernel|conv_ws_kernel" -s 14 -c 7 -f -o $OUT/prof_gray_$TAG python tools/gray_bench.py > $OUT/ncu_gray_$TAG.log 2>&1
echo "ncu rc=$?"; tail -n 2 $OUT/ncu_gray_$TAG.log
