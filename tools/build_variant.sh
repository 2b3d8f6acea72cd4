#!/bin/bash
# Builds another copy of libfmq_b200.so with extra -D flags, for A/B runs through FMQ_LIB_PATH.
#   tools/build_variant.sh NAME -DFMQ_IW_WARPS=8 -DFMQ_IW_STAGES=2 ...   ->  build/variants/libfmq_NAME.so
set -e
cd "$(dirname "$0")/.."
name=$1; shift
mkdir -p build/variants
nvcc -gencode arch=compute_100a,code=sm_100a -O3 -lineinfo -std=c++17 --fmad=false -diag-suppress 128 \
     -Xcompiler -fPIC -shared "$@" -o build/variants/libfmq_$name.so monitor_systam_flink_quickstart_b200/csrc/fmq_engine.cu
echo built build/variants/libfmq_$name.so
