#!/usr/bin/env python
"""bench.py -- examples/sec of the training hot path on B200, with roofline and CPU baseline.

Workload (BASELINE.json configs[1]): train_classifier, logistic loss + FTRL, 2^24 hashed
dims, Criteo-shaped synthetic rows (39 features/row), 100M rows per GPU, -mini_batch 65536
(BASELINE.json leaves the mini-batch size open; it is stated in `config`).

A "step" is one pass of the hot path over one ingest block of ROWS_PER_STEP rows (64
mini-batches): segmented sort-by-feature of the block, then per mini-batch the forward
and the fused reduce+update kernels.  Inputs are far larger than L2 (1.3 GB per step).

  value     rows/s with the block already resident in HBM (hm_udtf_process_device)
  e2e       rows/s through the host-buffer C-ABI call (hm_udtf_process on pinned host
            memory: H2D inside the timed region, plus a D2H read of the last mini-batch's
            margins every step)
  roofline  dominant kernel (largest share of device time), per-launch algorithmic bytes /
            CUDA-event duration on the learner's stream, against MEASURED_PEAKS.json
  cpu_baseline  the oracle (oracle/hm_oracle.c, OpenMP) on a bounded sample of the same
            workload on the host cores; kind "port": the reference mount holds no code

`--impl reference` times the CPU arm alone.  No published reference number exists
(BASELINE.md), so vs_baseline is null.
"""
from __future__ import annotations

import argparse
import json
import os
import statistics
import subprocess
import sys
import threading
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

MINI_BATCH = 65536
ROWS_PER_STEP = 1 << 22
TOTAL_ROWS = 24 * ROWS_PER_STEP          # 100,663,296 rows >= the 100M of configs[1]
SEED = 1


def _wl(fn, options, label, metric, dims=1 << 24, gen="criteo", nnz=39, regression=False, mix_every=1, total_rows=TOTAL_ROWS):
    return {"fn": fn, "options": options, "label": label, "metric": metric, "dims": dims, "gen": gen, "nnz": nnz,
            "regression": regression, "mix_every": mix_every, "total_rows": total_rows}


# The default is BASELINE.json configs[1]; the others are secondary measurements
# (`python bench.py --workload NAME`): "adagrad" is the optimizer north_star's target sentence names,
# "fm16" configs[2], "adam26" configs[3] and "fm64" configs[4] (both meant for --gpus 2/4/8).
WORKLOADS = {
    "ftrl": _wl("train_classifier", f"-loss logloss -opt ftrl -mini_batch {MINI_BATCH} -dims {1 << 24} -alpha 0.05 -lambda1 0.0001",
                "configs[1]: train_classifier logistic+FTRL, 2^24 hashed dims, Criteo-shaped synthetic (39 feats/row), "
                "100M rows per GPU",
                "examples/sec (train_classifier logloss+FTRL, 2^24 hashed dims, Criteo-shaped)"),
    "adagrad": _wl("train_classifier", f"-loss logloss -opt adagrad -mini_batch {MINI_BATCH} -dims {1 << 24} -eta0 0.05",
                   "north_star target: train_classifier logistic+AdaGrad, 2^24 hashed dims, Criteo-shaped synthetic "
                   "(39 feats/row), 100M rows per GPU",
                   "examples/sec (train_classifier logloss+AdaGrad, 2^24 hashed dims, Criteo-shaped)"),
    "fm16": _wl("train_fm", f"-factors 16 -classification -opt adagrad -mini_batch {MINI_BATCH} -dims {1 << 24} -eta0 0.05",
                "configs[2]: train_fm k=16 AdaGrad, 2^24 hashed dims, Criteo-shaped synthetic (39 feats/row)",
                "examples/sec (train_fm k=16 AdaGrad, 2^24 hashed dims, Criteo-shaped)"),
    "adam26": _wl("train_regressor", f"-loss squared -opt adam -mini_batch {MINI_BATCH} -dims {1 << 26} -eta0 0.001",
                  "configs[3]: train_regressor squared loss + Adam, 2^26 hashed dims, Criteo-shaped synthetic "
                  "(39 feats/row; nnz/row is not fixed by BASELINE.json), 100M rows per GPU (1B rows = 10 GPUs' worth)",
                  "examples/sec (train_regressor squared+Adam, 2^26 hashed dims, Criteo-shaped)",
                  dims=1 << 26, regression=True),
    "fm64": _wl("train_fm", f"-factors 64 -classification -opt adagrad -mini_batch {MINI_BATCH} -dims {1 << 24} -eta0 0.05",
                "configs[4]: train_fm k=64 AdaGrad, 2^24 hashed dims, KDD12-shaped synthetic (11 feats/row), "
                "16.8M rows per GPU (134M rows over 8 GPUs), replicas averaged once per pass over them",
                "examples/sec (train_fm k=64 AdaGrad, 2^24 hashed dims, KDD12-shaped)",
                gen="kdd12", nnz=11, mix_every=4, total_rows=4 * ROWS_PER_STEP),
}


def make_rows(wl: dict, n_rows: int, row0: int, device="cpu"):
    """Rows [row0, row0 + n_rows) of the workload's synthetic stream."""
    from hivemall_b200 import synth
    if wl["gen"] == "kdd12":
        return synth.kdd12_like(n_rows, dims=wl["dims"], seed=SEED, device=device, row0=row0)
    return synth.criteo_like(n_rows, dims=wl["dims"], seed=SEED, device=device, row0=row0, regression=wl["regression"])


def parse_args():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=24)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--rows-per-step", type=int, default=ROWS_PER_STEP)
    ap.add_argument("--total-rows", type=int, default=0, help="0 = the workload's own")
    ap.add_argument("--cpu-sample-rows", type=int, default=0, help="0 = sized for ~15 s")
    ap.add_argument("--workload", default="ftrl", choices=sorted(WORKLOADS))
    ap.add_argument("--no-e2e", action="store_true")
    ap.add_argument("--e2e-format", default="split", choices=["split", "csr"],
                    help="host format of the end-to-end arm for Criteo-shaped rows")
    ap.add_argument("--no-cpu", action="store_true")
    ap.add_argument("--no-numa", action="store_true", help="do not bind ranks to their GPU's NUMA node (N > 1)")
    return ap.parse_args()


# ------------------------------------------------------------------------- helpers

class ClockSampler:
    """nvidia-smi clocks / throttle reasons sampled every 200 ms while a region runs."""
    Q = ("clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.hw_slowdown,"
         "clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,"
         "clocks_event_reasons.sw_power_cap")

    def __init__(self, index: int):
        self.index = index
        self.proc = None
        self.lines: list[str] = []

    def __enter__(self):
        try:
            self.proc = subprocess.Popen(
                ["nvidia-smi", f"--id={self.index}", f"--query-gpu={self.Q}", "--format=csv,noheader,nounits",
                 "-lms", "200"], stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
            self.thread = threading.Thread(target=self._read, daemon=True)
            self.thread.start()
        except OSError:
            self.proc = None
        return self

    def _read(self):
        for line in self.proc.stdout:
            self.lines.append(line.strip())

    def __exit__(self, *exc):
        if self.proc:
            time.sleep(0.25)
            self.proc.terminate()
            try:
                self.proc.wait(timeout=2)
            except subprocess.TimeoutExpired:
                self.proc.kill()

    def summary(self) -> dict:
        sm, mx, reasons = [], [], set()
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        for ln in self.lines:
            parts = [p.strip() for p in ln.split(",")]
            if len(parts) < 7:
                continue
            try:
                sm.append(float(parts[0]))
                mx.append(float(parts[1]))
            except ValueError:
                continue
            for nm, v in zip(names, parts[3:7]):
                if v.lower().startswith("active"):
                    reasons.add(nm)
        if not sm:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": [], "samples": 0}
        return {"sm_mhz": statistics.median(sm), "sm_max_mhz": max(mx), "reasons": sorted(reasons),
                "samples": len(sm)}


def jsonable(o, path="out"):
    """Plain-Python copy of the result dict (numpy / torch scalars -> float / int)."""
    if isinstance(o, dict):
        return {k: jsonable(v, f"{path}.{k}") for k, v in o.items()}
    if isinstance(o, (list, tuple)):
        return [jsonable(v, f"{path}[]") for v in o]
    if o is None or isinstance(o, (bool, int, float, str)):
        return o
    if hasattr(o, "item"):
        print(f"bench.py: {path} was a {type(o).__name__}", file=sys.stderr)
        return o.item()
    return str(o)


def measured_peaks() -> tuple[float, str]:
    path = os.path.join(ROOT, "MEASURED_PEAKS.json")
    try:
        with open(path) as f:
            return float(json.load(f)["hbm_gbs"]), "measured (MEASURED_PEAKS.json hbm_gbs)"
    except Exception:
        return 6650.0, "fallback (B200_PROFILING.md 6.65 TB/s)"


_CPU_DATA = {}


def _cpu_rows(rows: int, workload: str):
    """First `rows` rows of the synthetic stream as host arrays (generated once per process)."""
    import numpy as np
    wl = WORKLOADS[workload]
    d = _CPU_DATA.setdefault(workload, {"rows": 0})
    if d["rows"] < rows:
        pieces = [make_rows(wl, min(1 << 20, rows - r0), r0) for r0 in range(0, rows, 1 << 20)]
        d.update(rows=rows,
                 feature=np.concatenate([p.feature.numpy() for p in pieces]),
                 value=np.concatenate([p.value.numpy() for p in pieces]),
                 label=np.concatenate([p.label.numpy() for p in pieces]))
    n = wl["nnz"]
    return (np.arange(rows + 1, dtype=np.int64) * n, d["feature"][: rows * n], d["value"][: rows * n], d["label"][:rows])


def cpu_baseline(rows: int, threads: int, workload: str = "ftrl") -> dict:
    """Oracle (OpenMP for the linear learner, scalar for train_fm) on the first `rows` rows of
    the same synthetic stream; only the training call is timed."""
    sys.path.insert(0, os.path.join(ROOT, "tests"))
    import hivemall_b200 as hm
    import _oracle
    fn, options = WORKLOADS[workload]["fn"], WORKLOADS[workload]["options"]
    cfg = _oracle.config_from(hm.parse_options(fn, options))
    rowptr, feature, value, label = _cpu_rows(rows, workload)
    if fn == "train_fm":
        threads = 1
        t0 = time.perf_counter()
        _oracle.train_fm(cfg, rowptr, feature, value, label)
        dt = time.perf_counter() - t0
        what = "hmo_train_fm (scalar)"
    else:
        t0 = time.perf_counter()
        _oracle.train_linear_mt(cfg, rowptr, feature, value, label, threads)
        dt = time.perf_counter() - t0
        what = f"hmo_train_linear_mt ({threads} OpenMP threads)"
    return {"value": rows / dt, "unit": "examples/s", "cores": threads, "kind": "port",
            "sample": f"first {rows} rows of the same stream, {options}, oracle/hm_oracle.c {what}, {dt:.1f} s; "
                      f"kind=port because the reference mount holds no code",
            "seconds": dt}


def run_reference_arm(args) -> None:
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    threads = os.cpu_count() or 1
    # each step is a bounded sample of the stream sized for about 3 s of CPU work (so that the
    # one-off allocation cost of a call does not dominate), whole run within a few minutes
    probe = cpu_baseline(4 * MINI_BATCH, threads, args.workload)
    budget = max(0.5, min(3.0, 150.0 / max(1, args.steps + args.warmup)))
    rows = int(max(4 * MINI_BATCH, min(probe["value"] * budget, 512 * MINI_BATCH)))
    rows = rows // MINI_BATCH * MINI_BATCH
    times = []
    for i in range(args.warmup + args.steps):
        r = cpu_baseline(rows, threads, args.workload)
        if i >= args.warmup:
            times.append(r["seconds"])
    dt = sum(times) / len(times)
    value = rows / dt
    out = {
        "impl": "reference", "metric": WORKLOADS[args.workload]["metric"],
        "value": value, "unit": "examples/s", "n_gpus": args.gpus, "steps": args.steps, "warmup": args.warmup,
        "ms_per_step": dt * 1e3, "higher_is_better": True, "scaling": "weak", "vs_baseline": None,
        "dtype": "f32 (fp64 sums)", "data": "synthetic",
        "config": {"workload": WORKLOADS[args.workload]["label"],
                   "options": WORKLOADS[args.workload]["options"], "mini_batch": MINI_BATCH, "rows_per_step": rows,
                   "note": "CPU arm; each step is a bounded sample of the stream"},
        "cpu_baseline": {"value": value, "unit": "examples/s", "cores": r["cores"], "kind": "port",
                         "sample": f"{rows} rows per step, {r['sample']}"},
        "e2e": {"value": value, "unit": "examples/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "gpu_launches": 0,
    }
    print(json.dumps(out))


# ------------------------------------------------------------------------- GPU arm

def main():
    args = parse_args()
    if args.impl == "reference":
        run_reference_arm(args)
        return

    import numpy as np
    import torch
    import torch.distributed as dist
    import hivemall_b200 as hm
    from hivemall_b200 import synth, mix

    world = int(os.environ.get("WORLD_SIZE", "1"))
    rank = int(os.environ.get("RANK", "0"))
    local_rank = int(os.environ.get("LOCAL_RANK", "0"))
    if not torch.cuda.is_available():
        raise SystemExit("bench.py needs a CUDA device: hivemall_b200 has no CPU fallback")
    torch.cuda.set_device(local_rank)
    dev = torch.device("cuda", local_rank)
    # one process per GPU: run next to it and allocate the pinned host buffers in its NUMA node
    host_cpus = hm.bind_host_to_gpu(local_rank) if world > 1 and not args.no_numa else None
    if world > 1:
        dist.init_process_group("nccl", device_id=dev)

    wl = WORKLOADS[args.workload]
    nnz_per_row = wl["nnz"]
    total_rows = args.total_rows or wl["total_rows"]
    K, W = args.steps, args.warmup
    rps = args.rows_per_step // MINI_BATCH * MINI_BATCH
    n_blocks = max(1, min(total_rows // rps, K + W))
    # ---- synthetic rows of this rank's shard, resident in HBM
    row_base = rank * total_rows
    blocks = []
    for b in range(n_blocks):
        parts = [make_rows(wl, min(1 << 20, rps - r), row_base + b * rps + r, device=dev) for r in range(0, rps, 1 << 20)]
        feature = torch.cat([p.feature for p in parts])
        value = torch.cat([p.value for p in parts])
        label = torch.cat([p.label for p in parts])
        rowptr = torch.arange(rps + 1, device=dev, dtype=torch.int64) * nnz_per_row
        blocks.append((rowptr, feature, value, label))
        del parts
    nnz_step = rps * nnz_per_row
    torch.cuda.synchronize()

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    # ---- device-resident arm ------------------------------------------------------
    fn, options, wl_label = wl["fn"], wl["options"], wl["label"]
    mix_every = wl["mix_every"]
    u = hm.UDTF(options, local_rank, function_name=fn)
    mixer = mix.ReplicaMixer(u, world) if world > 1 else None
    stream = torch.cuda.ExternalStream(u.stream, device=dev)

    def step_device(i):
        rowptr, feature, value, label = blocks[i % n_blocks]
        u.process_device(rps, nnz_step, rowptr, feature, value, label)
        if mixer and (i + 1) % mix_every == 0:
            mixer.mix()     # MIX averaging at a step boundary (a step = 64 mini-batches)

    for i in range(W):
        step_device(i)
    u.sync()
    barrier()
    launches0 = u.stats().kernel_launches
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    with ClockSampler(local_rank) as clocks:
        with torch.cuda.stream(stream):
            e0.record()
        for i in range(W, W + K):
            step_device(i)
        with torch.cuda.stream(stream):
            e1.record()
        u.sync()
        barrier()
    ms = e0.elapsed_time(e1)
    launches = u.stats().kernel_launches - launches0
    if world > 1:
        tms = torch.tensor([ms], device=dev)
        dist.all_reduce(tms, op=dist.ReduceOp.MAX)
        ms = float(tms)
    rows_per_s = world * rps * K / (ms / 1e3)

    # ---- same steps again with per-kernel CUDA events (roofline of the dominant kernel)
    u.set_profiling(True)
    p0, p1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    with torch.cuda.stream(stream):
        p0.record()
    for i in range(W + K, W + 2 * K):
        step_device(i)
    with torch.cuda.stream(stream):
        p1.record()
    u.sync()
    u.set_profiling(False)
    ms_prof = p0.elapsed_time(p1)
    kt = {k: v for k, v in u.kernel_times().items() if v["launches"] > 0}
    total_kernel_ms = sum(v["ms"] for v in kt.values())
    dom = max(kt, key=lambda k: kt[k]["ms"])
    d = kt[dom]
    peak, peak_src = measured_peaks()
    achieved = (d["bytes"] / d["launches"]) / (d["ms"] / d["launches"] * 1e-3) / 1e9 if d["ms"] > 0 else 0.0
    try:
        with open(os.path.join(ROOT, "profiles", "traffic.json")) as f:
            tj = json.load(f)
        traffic, traffic_src = tj.get(dom), tj.get("capture")
    except Exception:
        traffic, traffic_src = None, None
    roofline = {
        "bound": "hbm", "kernel": dom, "achieved": achieved, "peak": peak, "unit": "GB/s",
        "frac": achieved / peak, "traffic": traffic, "traffic_source": traffic_src, "peak_source": peak_src,
        "launches": d["launches"], "avg_us": d["ms"] / d["launches"] * 1e3,
        "algorithmic_bytes_per_launch": d["bytes"] / d["launches"],
        "share_of_kernel_time": d["ms"] / total_kernel_ms if total_kernel_ms else None,
        "kernels": {k: {"launches": v["launches"], "ms": round(v["ms"], 3),
                        "GBps": round(v["bytes"] / (v["ms"] * 1e-3) / 1e9, 1) if v["ms"] > 0 and v["bytes"] > 0 else None}
                    for k, v in sorted(kt.items(), key=lambda kv: -kv[1]["ms"])},
        "profiled_ms_per_step": ms_prof / K,
        "note": "per-kernel times come from a second pass of the same K steps with a CUDA-event pair around every "
                "launch and the sort/train stream overlap switched off (each kernel timed alone); `value` comes "
                "from the plain pass with the overlap on",
    }
    # whole-pipeline view: algorithmic bytes of every kernel over the plain timed region
    pipeline_bytes = sum(v["bytes"] for v in kt.values()) / K
    roofline["pipeline_GBps"] = pipeline_bytes / (ms / K * 1e-3) / 1e9

    # ---- end-to-end arm: host buffers through the C-ABI --------------------------------
    # Formats: "csr" = hm_udtf_process with a value per element; "index_only" = hm_udtf_process(value = NULL)
    # for rows whose values are all 1.0 (KDD12-shaped); "split" = hm_udtf_process_split for Criteo-shaped rows
    # (13 quantitative columns with values + 26 index-only columns), merged on the device: the same rows in
    # the same element order, without the implied 1.0s on the bus.
    def pin(t):
        addr, view = hm.host_alloc(t.numel() * t.element_size())
        pinned = torch.frombuffer(view, dtype=t.dtype)
        pinned.copy_(t)
        return addr, pinned

    def run_e2e(fmt):
        n_host = min(n_blocks, 4)
        n_q = 13
        host, addrs = [], []
        for b in range(n_host):
            rowptr, feature, value, label = blocks[b]
            if fmt == "split":
                f2, v2 = feature.view(rps, nnz_per_row), value.view(rps, nnz_per_row)
                steps = torch.arange(rps + 1, device=dev, dtype=torch.int64)
                parts = [steps * n_q, f2[:, :n_q].contiguous().view(-1), v2[:, :n_q].contiguous().view(-1),
                         steps * (nnz_per_row - n_q), f2[:, n_q:].contiguous().view(-1), label]
            elif fmt == "index_only":
                parts = [rowptr, feature, label]
            else:
                parts = [rowptr, feature, value, label]
            arrs = [pin(t) for t in parts]
            del parts
            addrs += [a for a, _ in arrs]
            host.append(arrs)
        torch.cuda.synchronize()
        v = hm.UDTF(options, local_rank, function_name=fn)
        mixer2 = mix.ReplicaMixer(v, world) if world > 1 else None
        margins = np.empty(MINI_BATCH, dtype=np.float32)
        losses = []
        pending = []     # (margins, labels) of the previous step: its loss is worked out on the host while the
                         # device is busy with the current one (the blocking D2H read itself stays in every step)

        def host_loss():
            if pending:
                m, t = pending.pop()
                if wl["regression"]:
                    losses.append(float(np.mean(0.5 * (m - t) ** 2)))
                else:
                    losses.append(float(np.mean(np.logaddexp(0.0, -(t * 2 - 1) * m))))

        def step_host(i):
            arrs = host[i % n_host]
            lbt = arrs[-1][1]
            if fmt == "split":
                v.process_split_pointers(rps, *[a for a, _ in arrs])
            elif fmt == "index_only":
                v.process_pointers(rps, arrs[0][0], arrs[1][0], None, arrs[2][0])
            else:
                v.process_pointers(rps, *[a for a, _ in arrs])
            if mixer2 and (i + 1) % mix_every == 0:
                mixer2.mix()
            host_loss()
            m = v.last_margins(margins)                      # D2H read of the step's result, waits
            pending.append((m.copy(), lbt[rps - len(m):].numpy()))

        for i in range(W):
            step_host(i)
        barrier()
        t0 = time.perf_counter()
        for i in range(W, W + K):
            step_host(i)
        host_loss()
        v.sync()
        barrier()
        dt = time.perf_counter() - t0
        if world > 1:
            tdt = torch.tensor([dt], device=dev)
            dist.all_reduce(tdt, op=dist.ReduceOp.MAX)
            dt = float(tdt)
        if fmt == "split":
            h2d = 2 * (rps + 1) * 8 + rps * n_q * 8 + rps * (nnz_per_row - n_q) * 4 + rps * 4
        else:
            h2d = (rps + 1) * 8 + nnz_step * (4 if fmt == "index_only" else 8) + rps * 4
        v.destroy()
        for a in addrs:
            hm.host_free(a)
        return {"value": world * rps * K / dt, "ms_per_step": dt / K * 1e3, "h2d_bytes_per_step": h2d,
                "pcie_GBps": h2d * K / dt / 1e9, "last_minibatch_loss": losses[-1]}

    e2e = None
    if not args.no_e2e:
        fmt = "index_only" if wl["gen"] == "kdd12" else ("csr" if args.e2e_format == "csr" else "split")
        main_run = run_e2e(fmt)
        e2e = {"value": main_run["value"], "unit": "examples/s", "h2d_bytes_per_step": main_run["h2d_bytes_per_step"],
               "d2h_bytes_per_step": MINI_BATCH * 4, "ms_per_step": main_run["ms_per_step"],
               "last_minibatch_loss": main_run["last_minibatch_loss"],
               "loss": "squared/2" if wl["regression"] else "logloss", "pcie_GBps": main_run["pcie_GBps"],
               "timing": "host wall clock around K calls, each ending in a blocking D2H read",
               "host_binding": (f"rank bound to {len(host_cpus)} CPUs next to its GPU (NVML affinity)" if host_cpus
                                else "none"),
               "format": {"split": "hm_udtf_process_split: 13 quantitative columns (index, value) + 26 index-only "
                                   "columns per row, merged on the device (the implied 1.0s stay off the bus)",
                          "index_only": "hm_udtf_process with value = NULL (every value is 1.0)",
                          "csr": "hm_udtf_process: plain CSR, one f32 value per element"}[fmt]}
        if fmt == "split":      # the same rows as plain CSR, for comparison
            plain = run_e2e("csr")
            e2e["plain_csr"] = {"value": plain["value"], "ms_per_step": plain["ms_per_step"],
                                "h2d_bytes_per_step": plain["h2d_bytes_per_step"], "pcie_GBps": plain["pcie_GBps"],
                                "last_minibatch_loss": plain["last_minibatch_loss"]}

    # ---- CPU baseline (rank 0, N=1 only) -------------------------------------------
    cpu = None
    if rank == 0 and world == 1 and not args.no_cpu:
        threads = os.cpu_count() or 1
        rows = args.cpu_sample_rows
        if rows <= 0:
            probe = cpu_baseline(2 * MINI_BATCH, threads, args.workload)
            rows = int(max(2 * MINI_BATCH, min(probe["value"] * 15.0, 512 * MINI_BATCH)))
        rows = max(MINI_BATCH, rows // MINI_BATCH * MINI_BATCH)
        cpu = cpu_baseline(rows, threads, args.workload)
        cpu.pop("seconds", None)

    if rank == 0:
        out = {
            "metric": wl["metric"],
            "value": rows_per_s, "unit": "examples/s", "n_gpus": world, "steps": K, "warmup": W,
            "ms_per_step": ms / K, "higher_is_better": True, "scaling": "weak", "vs_baseline": None,
            "dtype": "f32 (fp64 sums)", "data": "synthetic",
            "nnz_per_sec": rows_per_s * nnz_per_row,
            "config": {
                "workload": wl_label,
                "options": options, "mini_batch": MINI_BATCH, "rows_per_step": rps,
                "mini_batches_per_step": rps // MINI_BATCH, "rows_resident_per_gpu": n_blocks * rps,
                "l2": f"inputs larger than L2 ({(nnz_step * 8 + rps * 12) / 1e9:.2f} GB streamed per step)",
                "mix": (f"all-reduce of every replica array every {mix_every} step(s) of 64 mini-batches"
                        if world > 1 else "none (1 GPU)"),
                "parallelism": f"dp{world}: row shards, full replica per GPU",
            },
            "roofline": roofline,
            "cpu_baseline": cpu,
            "e2e": e2e,
            "gpu_launches": int(launches),
            "clocks": clocks.summary(),
        }
        print(json.dumps(jsonable(out)))
    u.destroy()
    if world > 1:
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
