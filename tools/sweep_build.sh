#!/bin/bash
# Tuning builds of the CUDA library: chunk size / CTA size / staging capacity variants.
# usage: tools/sweep_build.sh "16384 256 1024 2" "8192 256 512 4" ...   -> build/sweep/libti_<chunk>_<threads>_<stage>_<minblocks>.so
set -e
cd "$(dirname "$0")/.."
mkdir -p build/sweep
for v in "$@"; do
  set -- $v
  out=build/sweep/libti_$1_$2_$3_$4_${5:-2}${TI_TAG}.so
  /usr/local/cuda/bin/nvcc -gencode arch=compute_100a,code=sm_100a -lineinfo -O3 -std=c++17 --expt-relaxed-constexpr \
    --extended-lambda -Xcompiler -fPIC -shared -DTI_CHUNK=$1 -DTI_THREADS=$2 -DTI_TXN_STAGE=$3 -DTI_MIN_BLOCKS=$4 -DTI_TEXT_BUFS=${5:-2} $TI_EXTRA -o $out \
    accel_sim_framework_distributed_b200/csrc/ti_api.cu accel_sim_framework_distributed_b200/csrc/ti_host.cpp &
done
wait
ls -la build/sweep
