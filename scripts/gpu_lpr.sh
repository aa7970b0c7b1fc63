#!/bin/bash
# On the GPU box: forward kernel, lanes per row (HM_FWD_LPR) with hot features on, then the hot/PDL/linear tests.
mkdir -p gpurun_out
B="--steps 8 --warmup 3 --no-e2e --no-cpu --no-sweep"
run() { echo -n "$1: "; shift; env "$@" timeout 300 python bench.py $B 2>&1 | python scripts/benchline.py; }
for rep in 1 2; do
  run "auto" HM_X=1
  run "lpr32" HM_FWD_LPR=32
  run "lpr16" HM_FWD_LPR=16
  [ -f variants/lib_head.so ] && { cp hivemall_b200/libhivemall_b200.so /tmp/lib_base.so; cp variants/lib_head.so hivemall_b200/libhivemall_b200.so; run "head" HM_X=1; cp /tmp/lib_base.so hivemall_b200/libhivemall_b200.so; }
done 2>&1 | tee gpurun_out/lpr_ab3.txt
timeout 900 python -m pytest tests/test_gpu_hot.py tests/test_gpu_pdl.py tests/test_gpu_linear.py tests/test_gpu_fullsize.py -x -q 2>&1 | tail -4
