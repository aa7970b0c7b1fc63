#!/bin/bash
# usage (on the GPU box): [AB_ARGS="--workload fm16"] scripts/ab.sh v0 v1 ...   -- bench each variants/lib_NAME.so in turn (twice, interleaved)
for rep in 1 2; do for v in "$@"; do
  cp variants/lib_$v.so hivemall_b200/libhivemall_b200.so
  echo -n "$v: "; timeout 300 python bench.py --steps 12 --warmup 3 --no-e2e --no-cpu ${AB_ARGS} 2>&1 | python scripts/benchline.py
done; done
