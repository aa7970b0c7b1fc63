#!/bin/bash
# On a multi-GPU box: usage scripts/gpu_multi.sh N  -- 2-GPU tests, smoke() with its MIX schedule, bench.py on N GPUs
# (mix_check + secondary), concurrent host-to-device bandwidth.  Everything lands in gpurun_out/.
N=${1:-2}
mkdir -p gpurun_out
nvidia-smi -L | head -9
timeout 600 python -m pytest tests/test_gpu_multi.py -m gpu -x -q 2>&1 | tail -3
timeout 600 python -c "import __graft_entry__ as g; g.smoke()" 2>&1 | tail -2
timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29517 \
  bench.py --gpus $N --steps ${STEPS:-12} --warmup 3 > gpurun_out/bench_${N}gpu.json 2> gpurun_out/bench_${N}gpu.err
echo "bench rc=$?"; tail -c 400 gpurun_out/bench_${N}gpu.err; python scripts/benchline.py < gpurun_out/bench_${N}gpu.json
timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29518 \
  scripts/micro/h2d_concurrent.py --out gpurun_out/h2d_concurrent_${N}.json 2>&1 | tail -2
