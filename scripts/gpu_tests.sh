#!/bin/bash
# On the GPU box: the -m gpu suite (optionally only the files given), then a short device-resident bench.
mkdir -p gpurun_out
timeout 1500 python -m pytest ${@:-tests} -m gpu -x -q > gpurun_out/pytest.log 2>&1; echo "pytest rc=$?" | tee -a gpurun_out/pytest.log
tail -25 gpurun_out/pytest.log
echo -n "bench: "; timeout 300 python bench.py --steps 8 --warmup 3 --no-e2e --no-cpu --no-sweep 2>&1 | python scripts/benchline.py
