#!/bin/bash
# usage: tools/build_lib.sh <out.so> [extra nvcc flags...]   -- a tuning build of the CUDA library
set -e
cd "$(dirname "$0")/.."
out=$1; shift
/usr/local/cuda/bin/nvcc -gencode arch=compute_100a,code=sm_100a -lineinfo -O3 -std=c++17 --expt-relaxed-constexpr \
  --extended-lambda -Xcompiler -fPIC -shared "$@" -o $out \
  accel_sim_framework_distributed_b200/csrc/ti_api.cu accel_sim_framework_distributed_b200/csrc/ti_host.cpp
