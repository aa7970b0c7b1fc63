#!/usr/bin/env python
"""Host-to-device copy bandwidth with 1..N ranks copying at the same time (one process per GPU).

Why: bench.py's end-to-end arm scales 1.00 / 0.98 / 0.53 / 0.40 over 1 / 2 / 4 / 8 GPUs while the
device-resident arm scales 0.96.  This measures the only thing the end-to-end arm adds -- pinned
host memory -> HBM copies of the slice size the library uses (170 MiB), from buffers allocated by
the library's own hm_host_alloc -- with every rank copying concurrently, nothing else running.
If the box cannot deliver more than ~170 GB/s in total, the end-to-end ceiling is the host's, not
the library's.

launch: python -m torch.distributed.run --nnodes=1 --nproc-per-node N --master-addr 127.0.0.1 \
        --master-port 29511 scripts/micro/h2d_concurrent.py [--out profiles/r02_h2d_concurrent.json]
Rank 0 prints / writes one JSON object: per-rank and total GB/s for k = 1, 2, 4, ... N active ranks.
"""
import argparse
import json
import os
import sys

import torch
import torch.distributed as dist

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)
import hivemall_b200 as hm  # noqa: E402


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--out", default="")
    ap.add_argument("--mib", type=int, default=170)
    ap.add_argument("--reps", type=int, default=24)
    ap.add_argument("--no-numa", action="store_true")
    args = ap.parse_args()
    world = int(os.environ.get("WORLD_SIZE", "1"))
    rank = int(os.environ.get("RANK", "0"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    cpus = None if args.no_numa else hm.bind_host_to_gpu(local)
    if world > 1:
        dist.init_process_group("nccl", device_id=dev)
    n = args.mib << 20
    addr, view = hm.host_alloc(n)
    h = torch.frombuffer(view, dtype=torch.uint8)
    h.fill_(rank + 1)
    d = torch.empty(n, dtype=torch.uint8, device=dev)
    s = torch.cuda.Stream()

    def timed(active: bool) -> float:
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()
        gbps = 0.0
        if active:
            with torch.cuda.stream(s):
                for _ in range(2):
                    d.copy_(h, non_blocking=True)
                e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
                e0.record()
                for _ in range(args.reps):
                    d.copy_(h, non_blocking=True)
                e1.record()
                s.synchronize()
            gbps = n * args.reps / (e0.elapsed_time(e1) * 1e-3) / 1e9
        if world > 1:
            dist.barrier()
        return gbps

    results = []
    k = 1
    ks = []
    while k < world:
        ks.append(k)
        k *= 2
    ks.append(world)
    for k in ks:
        g = timed(rank < k)
        t = torch.tensor([g], device=dev, dtype=torch.float64)
        allv = [torch.zeros_like(t) for _ in range(world)]
        if world > 1:
            dist.all_gather(allv, t)
        else:
            allv = [t]
        per = [float(x) for x in allv][:k]
        results.append({"active_ranks": k, "per_rank_GBps": [round(x, 2) for x in per], "total_GBps": round(sum(per), 2),
                        "min_GBps": round(min(per), 2)})
    if rank == 0:
        out = {"what": "pinned host (hm_host_alloc) -> HBM, cudaMemcpyAsync of %d MiB x %d per rank, k ranks copying "
                       "concurrently, CUDA-event timed per rank" % (args.mib, args.reps),
               "world": world, "host_cpus_visible": os.cpu_count(),
               "host_binding": "NVML affinity" if cpus else "none", "results": results}
        print(json.dumps(out))
        if args.out:
            with open(os.path.join(ROOT, args.out), "w") as f:
                json.dump(out, f, indent=1)
    hm.host_free(addr)
    if world > 1:
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
