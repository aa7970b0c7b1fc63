#!/bin/bash
mkdir -p gpurun_out
timeout 900 python -m pytest tests -m gpu -x -q > gpurun_out/pytest.log 2>&1; echo "pytest rc=$?"; tail -3 gpurun_out/pytest.log
B="--steps 3 --warmup 1 --no-e2e --no-cpu --no-sweep"
for mb in 8192 1024 128 16384; do
  echo -n "mb $mb pdl on:  "; timeout 300 python bench.py $B --mini-batch $mb 2>&1 | python scripts/benchline.py
  echo -n "mb $mb pdl off: "; HM_PDL=0 timeout 300 python bench.py $B --mini-batch $mb 2>&1 | python scripts/benchline.py
done
echo -n "mb 65536: "; timeout 300 python bench.py --steps 8 --warmup 3 --no-e2e --no-cpu --no-sweep 2>&1 | python scripts/benchline.py
