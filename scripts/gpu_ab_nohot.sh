#!/bin/bash
# A/B with hot features off (isolates the plain forward path)
cp hivemall_b200/libhivemall_b200.so /tmp/lib_base.so
B="--steps 8 --warmup 3 --no-e2e --no-cpu --no-sweep"
for rep in 1 2; do
  echo -n "base nohot: "; HM_HOT_FEATURES=0 timeout 300 python bench.py $B 2>&1 | python scripts/benchline.py
  for v in "$@"; do
    cp variants/lib_$v.so hivemall_b200/libhivemall_b200.so
    echo -n "$v nohot: "; HM_HOT_FEATURES=0 timeout 300 python bench.py $B 2>&1 | python scripts/benchline.py
    cp /tmp/lib_base.so hivemall_b200/libhivemall_b200.so
  done
done
