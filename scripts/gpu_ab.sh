#!/bin/bash
# On the GPU box: A/B of the in-tree build against variants/lib_*.so (bench.py device-resident arm only).
cp hivemall_b200/libhivemall_b200.so /tmp/lib_base.so
B="--steps 8 --warmup 3 --no-e2e --no-cpu --no-sweep ${AB_ARGS}"
for rep in 1 2; do
  echo -n "base: "; timeout 300 python bench.py $B 2>&1 | python scripts/benchline.py
  for v in "$@"; do
    cp variants/lib_$v.so hivemall_b200/libhivemall_b200.so
    echo -n "$v: "; timeout 300 python bench.py $B 2>&1 | python scripts/benchline.py
    cp /tmp/lib_base.so hivemall_b200/libhivemall_b200.so
  done
done
