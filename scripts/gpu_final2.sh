#!/bin/bash
# On the GPU box: round-end evidence in one call -- suite, smoke, reference arm, default bench line, launch list,
# `ncu --set full` captures of the four hot kernels, secondary workloads.
bash scripts/gpu_final.sh
timeout 300 ncu -k regex:"sort_|linear_|hot_|seg_|scan_|tile_|row_" --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/launches.csv \
  python bench.py --steps 1 --warmup 1 --no-e2e --no-cpu --no-sweep > gpurun_out/ncu.log 2>&1; echo "ncu launch list rc=$?"
bash scripts/gpu_ncu.sh
bash scripts/gpu_workloads.sh
