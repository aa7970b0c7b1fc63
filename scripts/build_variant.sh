#!/bin/bash
# usage: scripts/build_variant.sh NAME path/to/hm_linear_variant.cu [extra nvcc flags]   -> variants/lib_NAME.so
# (A/B harness for kernel experiments: the GPU box copies one variant over hivemall_b200/libhivemall_b200.so per run)
set -e
name=$1; src=$2; shift 2
cd "$(dirname "$0")/../hivemall_b200/csrc"
cp "$src" ./_var_$name.cu
nvcc -gencode arch=compute_100a,code=sm_100a -O3 -lineinfo -std=c++17 -fmad=false -Xcompiler -fPIC,-O2 -shared -cudart shared \
  -Xptxas -v -I ../../include "$@" -o ../../variants/lib_$name.so hm_options.cpp hm_hash.cu hm_sort.cu _var_$name.cu hm_fm.cu hm_learner.cu 2>&1 \
  | grep -A2 "linear_update_kernel\|linear_forward_kernel" | grep "spill\|registers" | sed "s/^/$name /"
rm -f ./_var_$name.cu
