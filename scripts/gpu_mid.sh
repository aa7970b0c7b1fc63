#!/bin/bash
# On the GPU box: the -m gpu suite, then the mini-batch sweep of the in-tree build against variants/lib_*.so.
mkdir -p gpurun_out
if [ -z "$NO_TESTS" ]; then timeout 1500 python -m pytest tests -m gpu -x -q > gpurun_out/pytest.log 2>&1; echo "pytest rc=$?"; tail -3 gpurun_out/pytest.log; fi
B="--steps 6 --warmup 2 --no-e2e --no-cpu --no-sweep"
cp hivemall_b200/libhivemall_b200.so /tmp/lib_base.so
for mb in ${MBS:-512 1024 2048 4096 8192 12288 16384 65536}; do
  echo -n "mb=$mb base: "; timeout 300 python bench.py $B --mini-batch $mb 2>&1 | python scripts/benchline.py
  for v in "$@"; do
    cp variants/lib_$v.so hivemall_b200/libhivemall_b200.so
    echo -n "mb=$mb $v: "; timeout 300 python bench.py $B --mini-batch $mb 2>&1 | python scripts/benchline.py
    cp /tmp/lib_base.so hivemall_b200/libhivemall_b200.so
  done
done 2>&1 | tee gpurun_out/mid_variants2.txt
