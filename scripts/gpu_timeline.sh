#!/bin/bash
# On the GPU box: per-CTA timeline of the last mini-batch and the sequence of the last block's mini-batches (-DHM_STAMP build).
mkdir -p gpurun_out
cp hivemall_b200/libhivemall_b200.so /tmp/lib_base.so
cp variants/lib_stamp.so hivemall_b200/libhivemall_b200.so
for mb in 128 2048 8192 65536; do timeout 120 env PYTHONPATH=. python scripts/mb_sequence.py ftrl $mb 2>&1 | tail -14; done | tee gpurun_out/mb_sequence.txt
cp /tmp/lib_base.so hivemall_b200/libhivemall_b200.so
