#!/bin/bash
# On the GPU box: one `ncu --set full` capture per hot kernel of bench.py's default workload (after the program
# has run clean without ncu).  Reports land in gpurun_out/ and are read back with `ncu -i ... --page raw --csv`.
mkdir -p gpurun_out
B="--steps 1 --warmup 1 --no-e2e --no-cpu --no-sweep"
timeout 300 python bench.py $B > /dev/null 2>&1 || { echo "bench failed without ncu"; exit 1; }
for k in sort_scatter linear_forward linear_update sort_hist; do
  timeout 600 ncu --set full --clock-control none --import-source on -k regex:${k}_kernel --launch-skip 9 --launch-count 3 \
    -f -o gpurun_out/r02_${k} python bench.py $B > gpurun_out/ncu_${k}.log 2>&1
  echo "$k rc=$?"; ls -la gpurun_out/r02_${k}.ncu-rep 2>/dev/null
done
