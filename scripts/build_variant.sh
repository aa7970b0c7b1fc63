#!/bin/bash
# usage: scripts/build_variant.sh NAME [extra nvcc flags, e.g. -DHM_STAMP -DHM_EXPERIMENT=2]  -> variants/lib_NAME.so
# A/B harness for kernel experiments: builds the in-tree sources with extra flags; on the GPU box
# scripts/ab.sh copies one variant at a time over hivemall_b200/libhivemall_b200.so and runs bench.py.
set -e
name=$1; shift
cd "$(dirname "$0")/../hivemall_b200/csrc"
mkdir -p ../../variants
nvcc -gencode arch=compute_100a,code=sm_100a -O3 -lineinfo -std=c++17 -fmad=false -Xcompiler -fPIC,-O2 -shared -cudart shared \
  -Xptxas -v -I ../../include "$@" -o ../../variants/lib_$name.so hm_options.cpp hm_table.cpp hm_hash.cu hm_sort.cu hm_linear.cu hm_fm.cu hm_learner.cu 2>&1 \
  | grep -A2 "${VARIANT_GREP:-linear_update_kernel}" | grep "spill\|registers" | sed "s/^/$name /" || true
