#!/bin/bash
# A/B builds of the CUDA library with build-time overrides:  bash tools/build_variant.sh NAME -DAHK_C1T_NO_MMA | -DAHK_C1T_NO_CONVERT  (timing experiments: conv1t_kernel with one role removed)
# -> tools/ab/libahk_NAME.so (git-ignored; travels to the GPU box), for `python tools/policy_bench.py --lib tools/ab/libahk_NAME.so`
set -e
NAME=$1; shift
cd "$(dirname "$0")/../deep-learning-air-hockey-robot_b200/csrc"
mkdir -p ../../tools/ab
nvcc -gencode arch=compute_100a,code=sm_100a -O3 -std=c++17 -lineinfo -Xcompiler -fPIC -cudart static "$@" -shared \
     -o ../../tools/ab/libahk_$NAME.so ahk_api.cu ahk_kernels.cu ahk_step.cu ahk_raster.cu ahk_rl.cu ahk_train.cu ahk_policy.cu 2>&1 | grep -v "warning #177\|Remark\|^$" || true
ls -la ../../tools/ab/libahk_$NAME.so
