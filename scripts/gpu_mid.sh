#!/bin/bash
# On the GPU box: mid-size mini-batches -- lanes per row of the forward kernel.
mkdir -p gpurun_out
B="--steps 8 --warmup 3 --no-e2e --no-cpu --no-sweep"
for mb in 4096 8192 16384 32768; do
  for lpr in 32 16; do
    echo -n "mb=$mb lpr=$lpr: "; HM_FWD_LPR=$lpr timeout 300 python bench.py $B --mini-batch $mb 2>&1 | python scripts/benchline.py
  done
done 2>&1 | tee gpurun_out/mid_lpr.txt
