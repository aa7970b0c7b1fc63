#!/bin/bash
# On the GPU box: full -m gpu suite, then A/B of the in-tree build against variants/lib_*.so, an ncu launch
# list and the update kernel's per-CTA timeline.  Everything lands in gpurun_out/.
mkdir -p gpurun_out
timeout 1200 python -m pytest tests -m gpu -x -q > gpurun_out/pytest.log 2>&1; echo "pytest rc=$?" | tee -a gpurun_out/pytest.log
tail -5 gpurun_out/pytest.log
cp hivemall_b200/libhivemall_b200.so /tmp/lib_base.so
B="--steps 8 --warmup 3 --no-e2e --no-cpu --no-sweep"
echo -n "base: "; timeout 300 python bench.py $B 2>&1 | python scripts/benchline.py
for v in "$@"; do
  cp variants/lib_$v.so hivemall_b200/libhivemall_b200.so
  echo -n "$v: "; timeout 300 python bench.py $B 2>&1 | python scripts/benchline.py
done
cp /tmp/lib_base.so hivemall_b200/libhivemall_b200.so
echo -n "hot off: "; HM_HOT_FEATURES=0 timeout 300 python bench.py $B 2>&1 | python scripts/benchline.py
echo -n "base again: "; timeout 300 python bench.py $B 2>&1 | python scripts/benchline.py
timeout 300 ncu -k regex:"sort_|linear_|hot_|seg_|scan_|tile_" --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/launches.csv \
  python bench.py --steps 1 --warmup 1 --no-e2e --no-cpu --no-sweep > gpurun_out/ncu.log 2>&1; echo "ncu rc=$?"
if [ -f variants/lib_stamp.so ]; then
  cp variants/lib_stamp.so hivemall_b200/libhivemall_b200.so
  timeout 120 env PYTHONPATH=. python scripts/upd_timeline.py ftrl 2>&1 | tee gpurun_out/timeline.txt
  cp /tmp/lib_base.so hivemall_b200/libhivemall_b200.so
fi
if [ -n "$FULL_BENCH" ]; then
  timeout 900 python bench.py > gpurun_out/bench_full.json 2> gpurun_out/bench_full.err; echo "full bench rc=$?"; tail -c 600 gpurun_out/bench_full.err
fi
