#!/usr/bin/env python
"""bench.py -- examples/sec of the training hot path on B200, with roofline and CPU baseline.

Workload (BASELINE.json configs[1]): train_classifier, logistic loss + FTRL, 2^24 hashed
dims, Criteo-shaped synthetic rows (39 features/row), 100M rows per GPU, -mini_batch 65536
(BASELINE.json leaves the mini-batch size open; it is stated in `config`, `--mini-batch`
changes it, and the N=1 line carries a sweep over 128 / 1024 / 8192 / 65536).

A "step" is one pass of the hot path over one ingest call of ROWS_PER_STEP rows (64
mini-batches of 65536): hot-feature selection and segmented sort-by-feature of its blocks, then
per mini-batch the forward and the fused reduce+update kernels.  Inputs are far larger than
L2 (1.3 GB per step).

  value     rows/s with the rows already resident in HBM (hm_udtf_process_device)
  e2e       rows/s through the host-buffer C-ABI call (hm_udtf_process* on pinned host
            memory: H2D inside the timed region, plus a D2H read of the last mini-batch's
            margins every step)
  roofline  dominant kernel (largest share of device time), per-launch algorithmic bytes /
            CUDA-event duration on the learner's stream, against MEASURED_PEAKS.json; and
            pipeline_frac = rows x bytes_per_example / step time / peak
  cpu_baseline  the oracle (oracle/hm_oracle.c, OpenMP) on a bounded sample of the same
            workload on the host cores; kind "port": the reference mount holds no code
  N > 1     mix_check (replicas identical after the last MIX all-reduce, its cost timed with
            its own events) and `secondary`: short runs of BASELINE.json's two multi-GPU
            configs (configs[3] adam26, configs[4] fm64)

`--impl reference` times the CPU arm alone, in a process that loads oracle/ only.  No
published reference number exists, so vs_baseline is null.
"""
from __future__ import annotations

import argparse
import importlib.util
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


def _wl(fn, options, label, metric, dims=1 << 24, gen="criteo", nnz=39, regression=False, mix_every=1,
        total_rows=TOTAL_ROWS, state_words=2, factors=0):
    return {"fn": fn, "options": options, "label": label, "metric": metric, "dims": dims, "gen": gen, "nnz": nnz,
            "regression": regression, "mix_every": mix_every, "total_rows": total_rows, "state_words": state_words,
            "factors": factors}


# The default is BASELINE.json configs[1]; the others are secondary measurements
# (`python bench.py --workload NAME`): "adagrad" is the optimizer north_star's target sentence names,
# "fm16" configs[2], "adam26" configs[3] and "fm64" configs[4] (both also run, briefly, under --gpus 2/4/8).
# state_words = fp32 words per feature that SURVEY.md 8(d)'s update term reads and writes (S_r = S_w):
# FTRL is counted with the survey's provisional (z, n, w) = 3 although the replica stores (z, n) only.
# The option strings carry a literal {mb}: the mini-batch size is an argument (--mini-batch).
WORKLOADS = {
    "ftrl": _wl("train_classifier", "-loss logloss -opt ftrl -mini_batch {mb} -dims 16777216 -alpha 0.05 -lambda1 0.0001",
                "configs[1]: train_classifier logistic+FTRL, 2^24 hashed dims, Criteo-shaped synthetic (39 feats/row), "
                "100M rows per GPU",
                "examples/sec (train_classifier logloss+FTRL, 2^24 hashed dims, Criteo-shaped)", state_words=3),
    "adagrad": _wl("train_classifier", "-loss logloss -opt adagrad -mini_batch {mb} -dims 16777216 -eta0 0.05",
                   "north_star target: train_classifier logistic+AdaGrad, 2^24 hashed dims, Criteo-shaped synthetic "
                   "(39 feats/row), 100M rows per GPU",
                   "examples/sec (train_classifier logloss+AdaGrad, 2^24 hashed dims, Criteo-shaped)", state_words=2),
    "fm16": _wl("train_fm", "-factors 16 -classification -opt adagrad -mini_batch {mb} -dims 16777216 -eta0 0.05",
                "configs[2]: train_fm k=16 AdaGrad, 2^24 hashed dims, Criteo-shaped synthetic (39 feats/row)",
                "examples/sec (train_fm k=16 AdaGrad, 2^24 hashed dims, Criteo-shaped)", state_words=2, factors=16),
    "adam26": _wl("train_regressor", "-loss squared -opt adam -mini_batch {mb} -dims 67108864 -eta0 0.001",
                  "configs[3]: train_regressor squared loss + Adam, 2^26 hashed dims, Criteo-shaped synthetic "
                  "(39 feats/row; nnz/row is not fixed by BASELINE.json), 100M rows per GPU (1B rows = 10 GPUs' worth)",
                  "examples/sec (train_regressor squared+Adam, 2^26 hashed dims, Criteo-shaped)",
                  dims=1 << 26, regression=True, state_words=3),
    "fm64": _wl("train_fm", "-factors 64 -classification -opt adagrad -mini_batch {mb} -dims 16777216 -eta0 0.05",
                "configs[4]: train_fm k=64 AdaGrad, 2^24 hashed dims, KDD12-shaped synthetic (11 feats/row), "
                "16.8M rows per GPU (134M rows over 8 GPUs), replicas averaged once per pass over them",
                "examples/sec (train_fm k=64 AdaGrad, 2^24 hashed dims, KDD12-shaped)",
                gen="kdd12", nnz=11, mix_every=4, total_rows=4 * ROWS_PER_STEP, state_words=2, factors=64),
    # not a BASELINE.json config: one of the online learners on the same kernels, for the record
    "pa1": _wl("train_pa1", "-mini_batch {mb} -dims 16777216 -c 0.1",
               "train_pa1 (passive-aggressive PA-I), 2^24 hashed dims, Criteo-shaped synthetic (39 feats/row)",
               "examples/sec (train_pa1, 2^24 hashed dims, Criteo-shaped)", state_words=1),
    # ... and one of the learners that keep a variance per feature: two update launches per mini-batch
    "arow": _wl("train_arow", "-mini_batch {mb} -dims 16777216 -r 1",
                "train_arow (adaptive regularization of weights, diagonal), 2^24 hashed dims, Criteo-shaped synthetic (39 feats/row)",
                "examples/sec (train_arow, 2^24 hashed dims, Criteo-shaped)", state_words=2),
}


def options_of(wl: dict, mini_batch: int = MINI_BATCH) -> str:
    return wl["options"].format(mb=mini_batch)


def bytes_per_example(wl: dict) -> dict:
    """ingest + forward + update, all fp32/int32.  n = nnz/row, S = state words read and
    written per touched feature, k = FM factors (k' = k: V carries AdaGrad state per element)."""
    n, s, k = wl["nnz"], wl["state_words"], wl["factors"]
    ingest = n * 8 + 8
    forward = n * 4 * (1 + k)
    update = n * 4 * (2 * s) * (1 + k)
    return {"bytes": ingest + forward + update, "ingest": ingest, "forward": forward, "update": update,
            "formula": "n*(4+4)+4+4  +  n*4*(1+k)  +  n*4*(S_r+S_w)*(1+k')",
            "n": n, "S_r": s, "S_w": s, "k": k,
            "note": ("FTRL counted as the survey's provisional (z, n, w); the replica stores (z, n) only "
                     "(S = 2 would give 1100 B)") if wl is WORKLOADS["ftrl"] else None}


def _synth():
    """hivemall_b200/synth.py loaded by path: pure torch, and importing it this way does not import the
    package (the CPU arm's process must not load the product library)."""
    mod = sys.modules.get("_hm_synth")
    if mod is None:
        spec = importlib.util.spec_from_file_location("_hm_synth", os.path.join(ROOT, "hivemall_b200", "synth.py"))
        mod = importlib.util.module_from_spec(spec)
        sys.modules["_hm_synth"] = mod
        spec.loader.exec_module(mod)
    return mod


def make_rows(wl: dict, n_rows: int, row0: int, device="cpu"):
    """Rows [row0, row0 + n_rows) of the workload's synthetic stream."""
    synth = _synth()
    if wl["gen"] == "kdd12":
        return synth.kdd12_like(n_rows, dims=wl["dims"], seed=SEED, device=device, row0=row0)
    return synth.criteo_like(n_rows, dims=wl["dims"], seed=SEED, device=device, row0=row0, regression=wl["regression"])


def config_of(wl: dict, mini_batch: int, rps: int, world: int, steps: int, warmup: int, total_rows: int) -> dict:
    """The `config` object: a function of the arguments only, so that both arms print the same one."""
    n_blocks = max(1, min(total_rows // rps, steps + warmup))
    return {
        "workload": wl["label"],
        "options": options_of(wl, mini_batch), "mini_batch": mini_batch, "rows_per_step": rps,
        "mini_batches_per_step": rps // mini_batch, "rows_resident_per_gpu": n_blocks * rps,
        "l2": f"inputs larger than L2 ({(rps * wl['nnz'] * 8 + rps * 12) / 1e9:.2f} GB streamed per step)",
        "mix": (f"all-reduce of every replica array every {wl['mix_every']} step(s) of {rps // mini_batch} mini-batches"
                if world > 1 else "none (1 GPU)"),
        "parallelism": f"dp{world}: row shards, full replica per GPU",
    }


def parse_args():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=24)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--rows-per-step", type=int, default=ROWS_PER_STEP)
    ap.add_argument("--mini-batch", type=int, default=MINI_BATCH)
    ap.add_argument("--total-rows", type=int, default=0, help="0 = the workload's own")
    ap.add_argument("--cpu-sample-rows", type=int, default=0, help="0 = sized for ~15 s")
    ap.add_argument("--workload", default="ftrl", choices=sorted(WORKLOADS))
    ap.add_argument("--no-e2e", action="store_true")
    ap.add_argument("--e2e-format", default="columns", choices=["columns", "split", "csr", "strings"],
                    help="host format of the end-to-end arm for Criteo-shaped rows")
    ap.add_argument("--no-cpu", action="store_true")
    ap.add_argument("--no-sweep", action="store_true", help="skip the mini-batch sweep of the N=1 line")
    ap.add_argument("--no-secondary", action="store_true", help="skip configs[3]/[4] on the N>1 line")
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


# ------------------------------------------------------------------------- CPU arm (oracle only)

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


def cpu_baseline(rows: int, threads: int, workload: str = "ftrl", mini_batch: int = MINI_BATCH) -> dict:
    """Oracle (OpenMP for the linear learner, scalar for train_fm) on the first `rows` rows of
    the same synthetic stream; only the training call is timed.  Loads oracle/ and nothing of
    the product: the option string is parsed by tests/_oracle.config_from_options."""
    sys.path.insert(0, os.path.join(ROOT, "tests"))
    import _oracle
    wl = WORKLOADS[workload]
    fn, options = wl["fn"], options_of(wl, mini_batch)
    cfg = _oracle.config_from_options(fn, options)
    rowptr, feature, value, label = _cpu_rows(rows, workload)
    if fn == "train_fm":
        threads = 1
        t0 = time.perf_counter()
        _oracle.train_fm(cfg, rowptr, feature, value, label)
        dt = time.perf_counter() - t0
        what = "hmo_train_fm (scalar)"
    elif 9 <= cfg.loss <= 14:
        # the variance-keeping learners: the oracle's loop for them is single-threaded (hmo_train_linear_mt falls back to it)
        threads = 1
        t0 = time.perf_counter()
        _oracle.train_linear_mt(cfg, rowptr, feature, value, label, 1)
        dt = time.perf_counter() - t0
        what = "train_cov (scalar)"
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
    wl = WORKLOADS[args.workload]
    mb = args.mini_batch
    threads = os.cpu_count() or 1
    rps = args.rows_per_step // mb * mb
    total_rows = args.total_rows or wl["total_rows"]
    # Each step is the GPU arm's step: rows_per_step rows of the same stream (the first rows_per_step rows,
    # a fresh model every step).  train_fm's oracle is scalar and would take minutes per step: its steps are a
    # bounded sample, and the line says so.
    rows = rps
    bounded = None
    if wl["fn"] == "train_fm":
        probe = cpu_baseline(min(rps, 2 * mb), threads, args.workload, mb)
        budget = max(0.5, min(3.0, 150.0 / max(1, args.steps + args.warmup)))
        rows = int(max(mb, min(probe["value"] * budget, rps))) // mb * mb
        bounded = f"train_fm oracle is scalar: {rows} of the step's {rps} rows are timed"
    times = []
    r = None
    for i in range(args.warmup + args.steps):
        r = cpu_baseline(rows, threads, args.workload, mb)
        if i >= args.warmup:
            times.append(r["seconds"])
    dt = sum(times) / len(times)
    value = rows / dt
    out = {
        "impl": "reference", "metric": wl["metric"],
        "value": value, "unit": "examples/s", "n_gpus": args.gpus, "steps": args.steps, "warmup": args.warmup,
        "ms_per_step": dt * 1e3 * (rps / rows), "higher_is_better": True, "scaling": "weak", "vs_baseline": None,
        "dtype": "f32 (fp64 sums)", "data": "synthetic",
        "config": config_of(wl, mb, rps, max(1, args.gpus), args.steps, args.warmup, total_rows),
        "cpu_baseline": {"value": value, "unit": "examples/s", "cores": r["cores"], "kind": "port",
                         "sample": f"{rows} rows per step ({bounded or 'the whole step'}), every step the first rows of "
                                   f"the stream with a fresh model; {r['sample']}"},
        "e2e": {"value": value, "unit": "examples/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "gpu_launches": 0,
        "loaded": "oracle/_build/libhm_oracle.so only (tests/_oracle.py parses the option string; hivemall_b200 is "
                  "not imported by this arm)",
    }
    print(json.dumps(out))


# ------------------------------------------------------------------------- GPU arm

class DeviceArm:
    """One workload resident in HBM on this rank, stepped through hm_udtf_process_device."""

    def __init__(self, wl, mini_batch, rps, steps, warmup, total_rows, world, rank, local_rank):
        import torch
        import hivemall_b200 as hm
        from hivemall_b200 import mix
        self.torch, self.hm = torch, hm
        self.wl, self.mb, self.rps, self.K, self.W = wl, mini_batch, rps, steps, warmup
        self.world, self.rank, self.local_rank = world, rank, local_rank
        self.dev = torch.device("cuda", local_rank)
        self.n_blocks = max(1, min(total_rows // rps, steps + warmup))
        row_base = rank * total_rows
        self.blocks = []
        for b in range(self.n_blocks):
            parts = [make_rows(wl, min(1 << 20, rps - r), row_base + b * rps + r, device=self.dev)
                     for r in range(0, rps, 1 << 20)]
            feature = torch.cat([p.feature for p in parts])
            value = torch.cat([p.value for p in parts])
            label = torch.cat([p.label for p in parts])
            rowptr = torch.arange(rps + 1, device=self.dev, dtype=torch.int64) * wl["nnz"]
            self.blocks.append((rowptr, feature, value, label))
            del parts
        self.nnz_step = rps * wl["nnz"]
        torch.cuda.synchronize()
        self.options = options_of(wl, mini_batch)
        self.u = hm.UDTF(self.options, local_rank, function_name=wl["fn"])
        self.mixer = mix.ReplicaMixer(self.u, world) if world > 1 else None
        self.stream = torch.cuda.ExternalStream(self.u.stream, device=self.dev)

    def step(self, i, timed_mix=False):
        rowptr, feature, value, label = self.blocks[i % self.n_blocks]
        self.u.process_device(self.rps, self.nnz_step, rowptr, feature, value, label)
        if self.mixer and (i + 1) % self.wl["mix_every"] == 0:
            self.mixer.mix(timed=timed_mix)     # MIX averaging at a step boundary

    def barrier(self):
        if self.world > 1:
            import torch.distributed as dist
            dist.barrier()
        self.torch.cuda.synchronize()

    def run(self, clock_index=None):
        """W untimed steps, then K timed ones (CUDA events on the learner's stream, max over ranks)."""
        torch = self.torch
        for i in range(self.W):
            self.step(i)
        self.u.sync()
        self.barrier()
        launches0 = self.u.stats().kernel_launches
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        sampler = ClockSampler(clock_index) if clock_index is not None else None
        if sampler:
            sampler.__enter__()
        with torch.cuda.stream(self.stream):
            e0.record()
        for i in range(self.W, self.W + self.K):
            self.step(i, timed_mix=True)
        with torch.cuda.stream(self.stream):
            e1.record()
        self.u.sync()
        self.barrier()
        if sampler:
            sampler.__exit__()
        ms = e0.elapsed_time(e1)
        launches = self.u.stats().kernel_launches - launches0
        if self.world > 1:
            import torch.distributed as dist
            tms = torch.tensor([ms], device=self.dev)
            dist.all_reduce(tms, op=dist.ReduceOp.MAX)
            ms = float(tms)
        out = {"value": self.world * self.rps * self.K / (ms / 1e3), "ms_per_step": ms / self.K, "ms": ms,
               "launches": int(launches), "clocks": sampler.summary() if sampler else None}
        if self.mixer:
            out["mix"] = self.mix_report()
        return out

    def mix_report(self):
        """Cost of the MIX all-reduce (its own event pairs) and whether it left the replicas identical:
        a 64-bit checksum of every replica array, MIN- and MAX-reduced over the ranks."""
        import torch.distributed as dist
        torch = self.torch
        m = self.mixer
        mix_ms = m.timed_ms()
        sums = []
        with torch.cuda.stream(self.stream):
            for t in m.tensors:
                sums.append(t.view(torch.int32).sum(dtype=torch.int64))
            for t in m.byte_maps:
                sums.append(t.sum(dtype=torch.int64))
            lo = torch.stack(sums)
            hi = lo.clone()
            dist.all_reduce(lo, op=dist.ReduceOp.MIN)
            dist.all_reduce(hi, op=dist.ReduceOp.MAX)
            same = bool(torch.equal(lo, hi))
        return {"replicas_identical": same, "mixes": m.mixes, "timed_mixes": len(mix_ms),
                "mix_ms_per_step": (sum(mix_ms) / self.K) if mix_ms else None,
                "mix_ms_each": (sum(mix_ms) / len(mix_ms)) if mix_ms else None,
                "bytes_per_mix": m.bytes_per_mix,
                "checksum": "sum of every replica array's words as int64, all-reduced MIN and MAX over the ranks "
                            "after the last mix"}

    def profile(self):
        """The same K steps again with a CUDA-event pair around every launch, streams serialised."""
        torch = self.torch
        self.u.set_profiling(True)
        p0, p1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        with torch.cuda.stream(self.stream):
            p0.record()
        for i in range(self.W + self.K, self.W + 2 * self.K):
            self.step(i)
        with torch.cuda.stream(self.stream):
            p1.record()
        self.u.sync()
        self.u.set_profiling(False)
        kt = {k: v for k, v in self.u.kernel_times().items() if v["launches"] > 0}
        return kt, p0.elapsed_time(p1) / self.K

    def close(self):
        self.u.destroy()
        self.blocks = []
        self.torch.cuda.empty_cache()


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
    mb = args.mini_batch
    nnz_per_row = wl["nnz"]
    total_rows = args.total_rows or wl["total_rows"]
    K, W = args.steps, args.warmup
    rps = args.rows_per_step // mb * mb
    fn, options = wl["fn"], options_of(wl, mb)

    # ---- device-resident arm ------------------------------------------------------
    arm = DeviceArm(wl, mb, rps, K, W, total_rows, world, rank, local_rank)
    n_blocks, blocks, nnz_step = arm.n_blocks, arm.blocks, arm.nnz_step

    def barrier():
        arm.barrier()

    main_run = arm.run(clock_index=local_rank)
    ms, rows_per_s, launches = main_run["ms"], main_run["value"], main_run["launches"]
    hot_feats = [int(x) for x in arm.u.hot_features()]

    # ---- same steps again with per-kernel CUDA events (roofline of the dominant kernel)
    kt, ms_prof_step = arm.profile()
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
    bpe = bytes_per_example(wl)
    pipeline_gbps = rps * bpe["bytes"] / (ms / K * 1e-3) / 1e9 / 1.0
    roofline = {
        "bound": "hbm", "kernel": dom, "achieved": achieved, "peak": peak, "unit": "GB/s",
        "frac": achieved / peak, "traffic": traffic, "traffic_source": traffic_src, "peak_source": peak_src,
        "launches": d["launches"], "avg_us": d["ms"] / d["launches"] * 1e3,
        "algorithmic_bytes_per_launch": d["bytes"] / d["launches"],
        "share_of_kernel_time": d["ms"] / total_kernel_ms if total_kernel_ms else None,
        "kernels": {k: {"launches": v["launches"], "ms": round(v["ms"], 3),
                        "GBps": round(v["bytes"] / (v["ms"] * 1e-3) / 1e9, 1) if v["ms"] > 0 and v["bytes"] > 0 else None}
                    for k, v in sorted(kt.items(), key=lambda kv: -kv[1]["ms"])},
        "profiled_ms_per_step": ms_prof_step,
        "note": "per-kernel times come from a second pass of the same K steps with a CUDA-event pair around every "
                "launch and the sort/train stream overlap switched off (each kernel timed alone); `value` comes "
                "from the plain pass with the overlap on",
        # whole-pipeline view, as SURVEY.md 8(d) defines it: the task's algorithmic bytes (not the
        # implementation's traffic: the sort's passes are not counted) over the plain timed region, per GPU
        "bytes_per_example": bpe,
        "pipeline_GBps": pipeline_gbps,
        "pipeline_frac": pipeline_gbps / peak,
    }

    # ---- end-to-end arm: host buffers through the C-ABI --------------------------------
    # Formats: "columns" = hm_udtf_process_columns for Criteo-shaped rows (the 13 quantitative indices given once,
    # a [rows][13] value matrix and a [rows][26] index matrix: no repeated index, no implied 1.0, no row pointer
    # on the bus); "csr" = hm_udtf_process with a value per element; "index_only" = hm_udtf_process(value = NULL)
    # for rows whose values are all 1.0 (KDD12-shaped); "split" = hm_udtf_process_split for Criteo-shaped rows
    # (13 quantitative columns with values + 26 index-only columns), merged on the device: the same rows in
    # the same element order, without the implied 1.0s on the bus; "strings" = hm_udtf_process_strings, rows as a
    # Hive UDTF receives them (array<string> of "name:value" / "name"), hashed on the device.
    def pin(t):
        addr, view = hm.host_alloc(t.numel() * t.element_size())
        pinned = torch.frombuffer(view, dtype=t.dtype)
        pinned.copy_(t)
        return addr, pinned

    def string_rows(block_index, n_rows):
        """Rows of a block as feature strings: quantitative "q<j>:<value>" (%.6g), categorical "c<j>=<index>".
        Returns (rowptr, bytes, offsets, label) host tensors."""
        rowptr, feature, value, label = blocks[block_index]
        f = feature[: n_rows * nnz_per_row].view(n_rows, nnz_per_row).cpu().numpy()
        v = value[: n_rows * nnz_per_row].view(n_rows, nnz_per_row).cpu().numpy()
        n_q = 13 if wl["gen"] != "kdd12" else 0
        cols = []
        for j in range(nnz_per_row):
            if j < n_q:
                cols.append(np.char.add(f"q{j}:", np.char.mod("%.6g", v[:, j])))
            else:
                cols.append(np.char.add(f"c{j}=", f[:, j].astype(np.int64).astype(str)))
        strs = np.stack(cols, axis=1).reshape(-1)
        enc = np.char.encode(strs, "ascii")
        lens = np.char.str_len(enc).astype(np.int64)
        offsets = np.zeros(len(enc) + 1, dtype=np.int64)
        offsets[1:] = np.cumsum(lens)
        data = np.frombuffer(b"".join(enc.tolist()), dtype=np.uint8).copy()
        rp = np.arange(n_rows + 1, dtype=np.int64) * nnz_per_row
        return (torch.from_numpy(rp), torch.from_numpy(data), torch.from_numpy(offsets), label[:n_rows].cpu())

    def run_e2e(fmt):
        n_host = min(n_blocks, 4)
        n_q = 13
        rows_step = rps
        if fmt == "strings":
            n_host = min(n_blocks, 2)
            rows_step = min(rps, 1 << 20)      # (building 160 M Python strings per block is the bench's cost, not the path's)
        host, addrs = [], []
        for b in range(n_host):
            rowptr, feature, value, label = blocks[b]
            if fmt == "columns":
                f2, v2 = feature.view(rps, nnz_per_row), value.view(rps, nnz_per_row)
                parts = [f2[0, :n_q].contiguous(), v2[:, :n_q].contiguous().view(-1), f2[:, n_q:].contiguous().view(-1), label]
            elif fmt == "split":
                f2, v2 = feature.view(rps, nnz_per_row), value.view(rps, nnz_per_row)
                steps = torch.arange(rps + 1, device=dev, dtype=torch.int64)
                parts = [steps * n_q, f2[:, :n_q].contiguous().view(-1), v2[:, :n_q].contiguous().view(-1),
                         steps * (nnz_per_row - n_q), f2[:, n_q:].contiguous().view(-1), label]
            elif fmt == "index_only":
                parts = [rowptr, feature, label]
            elif fmt == "strings":
                parts = list(string_rows(b, rows_step))
            else:
                parts = [rowptr, feature, value, label]
            arrs = [pin(t) for t in parts]
            del parts
            addrs += [a for a, _ in arrs]
            host.append(arrs)
        torch.cuda.synchronize()
        v = hm.UDTF(options, local_rank, function_name=fn)
        mixer2 = mix.ReplicaMixer(v, world) if world > 1 else None
        margins = np.empty(mb, dtype=np.float32)
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
            if fmt == "columns":
                v.process_columns_pointers(rows_step, n_q, arrs[0][0], arrs[1][0], nnz_per_row - n_q, arrs[2][0], arrs[3][0])
            elif fmt == "split":
                v.process_split_pointers(rows_step, *[a for a, _ in arrs])
            elif fmt == "index_only":
                v.process_pointers(rows_step, arrs[0][0], arrs[1][0], None, arrs[2][0])
            elif fmt == "strings":
                v.process_strings_pointers(rows_step, *[a for a, _ in arrs])
            else:
                v.process_pointers(rows_step, *[a for a, _ in arrs])
            if mixer2 and (i + 1) % wl["mix_every"] == 0:
                mixer2.mix()
            host_loss()
            m = v.last_margins(margins)                      # D2H read of the step's result, waits
            pending.append((m.copy(), lbt[rows_step - len(m):].numpy()))

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
        if fmt == "columns":
            h2d = n_q * 4 + rps * n_q * 4 + rps * (nnz_per_row - n_q) * 4 + rps * 4
        elif fmt == "split":
            h2d = 2 * (rps + 1) * 8 + rps * n_q * 8 + rps * (nnz_per_row - n_q) * 4 + rps * 4
        elif fmt == "strings":
            h2d = sum(t.numel() * t.element_size() for _, t in host[0])
        else:
            h2d = (rps + 1) * 8 + nnz_step * (4 if fmt == "index_only" else 8) + rps * 4
        v.destroy()
        for a in addrs:
            hm.host_free(a)
        return {"value": world * rows_step * K / dt, "ms_per_step": dt / K * 1e3, "h2d_bytes_per_step": h2d,
                "pcie_GBps": h2d * K / dt / 1e9, "last_minibatch_loss": losses[-1], "rows_per_step": rows_step}

    FORMATS = {"columns": "hm_udtf_process_columns: the 13 quantitative indices once, a [rows][13] value matrix, a [rows][26] "
                          "index matrix; expanded into CSR on the device (160 bytes per row on the bus)",
               "split": "hm_udtf_process_split: 13 quantitative columns (index, value) + 26 index-only "
                        "columns per row, merged on the device (the implied 1.0s stay off the bus)",
               "index_only": "hm_udtf_process with value = NULL (every value is 1.0)",
               "csr": "hm_udtf_process: plain CSR, one f32 value per element",
               "strings": "hm_udtf_process_strings: rows as feature strings (\"q3:0.25\", \"c17=90210\"), hashed "
                          "with MurmurHash3 on the device straight into the chunk's CSR"}
    e2e = None
    if not args.no_e2e:
        fmt = args.e2e_format
        if fmt != "strings":
            fmt = "index_only" if wl["gen"] == "kdd12" else fmt
        first = run_e2e(fmt)
        e2e = {"value": first["value"], "unit": "examples/s", "h2d_bytes_per_step": first["h2d_bytes_per_step"],
               "d2h_bytes_per_step": mb * 4, "ms_per_step": first["ms_per_step"],
               "rows_per_step": first["rows_per_step"],
               "last_minibatch_loss": first["last_minibatch_loss"],
               "loss": "squared/2" if wl["regression"] else "logloss", "pcie_GBps": first["pcie_GBps"],
               "timing": "host wall clock around K calls, each ending in a blocking D2H read",
               "host_binding": (f"rank bound to {len(host_cpus)} CPUs next to its GPU (NVML affinity)" if host_cpus
                                else "none"),
               "format": FORMATS[fmt]}
        if fmt == "columns":      # the same rows in the other host formats, for comparison
            sp = run_e2e("split")
            e2e["split"] = {"value": sp["value"], "ms_per_step": sp["ms_per_step"], "h2d_bytes_per_step": sp["h2d_bytes_per_step"],
                            "pcie_GBps": sp["pcie_GBps"], "last_minibatch_loss": sp["last_minibatch_loss"], "format": FORMATS["split"]}
            plain = run_e2e("csr")
            e2e["plain_csr"] = {"value": plain["value"], "ms_per_step": plain["ms_per_step"],
                                "h2d_bytes_per_step": plain["h2d_bytes_per_step"], "pcie_GBps": plain["pcie_GBps"],
                                "last_minibatch_loss": plain["last_minibatch_loss"]}
            if world == 1 and hasattr(hm.UDTF, "process_strings_pointers"):
                s = run_e2e("strings")
                e2e["strings"] = {"value": s["value"], "ms_per_step": s["ms_per_step"], "rows_per_step": s["rows_per_step"],
                                  "h2d_bytes_per_step": s["h2d_bytes_per_step"], "pcie_GBps": s["pcie_GBps"],
                                  "last_minibatch_loss": s["last_minibatch_loss"], "format": FORMATS["strings"]}

    arm.close()
    del blocks

    # ---- N = 1: the same workload at other mini-batch sizes, and configs[0] ----------------
    sweep = None
    if world == 1 and not args.no_sweep:
        sweep = {"workload": wl["label"], "rows_per_step": rps, "steps": 3, "warmup": 1,
                 "note": "device-resident value at other -mini_batch sizes (BASELINE.json names only 128, for configs[0])",
                 "examples_per_s": {str(mb): rows_per_s}}
        for m2 in (128, 1024, 8192, 65536):
            if m2 == mb:
                continue
            a2 = DeviceArm(wl, m2, rps // m2 * m2, 3, 1, 4 * rps, 1, 0, local_rank)
            sweep["examples_per_s"][str(m2)] = a2.run()["value"]
            a2.close()
        sweep["examples_per_s"] = dict(sorted(sweep["examples_per_s"].items(), key=lambda kv: int(kv[0])))
        # configs[0]: a9a-shaped, 32,768 rows, -mini_batch 128, logloss + AdaGrad, pushed from host memory
        data = synth.a9a_like(32768, seed=0)
        rp0, f0, v0, y0 = data.numpy()
        u0 = hm.UDTF("-loss logloss -opt AdaGrad -mini_batch 128 -dims 124", local_rank, function_name="train_classifier")
        times0 = []
        for i in range(6):
            t0 = time.perf_counter()
            u0.process_batch(rp0, f0, v0, y0)
            u0.sync()
            times0.append(time.perf_counter() - t0)
        u0.destroy()
        sweep["configs0_a9a"] = {"options": "-loss logloss -opt AdaGrad -mini_batch 128 -dims 124", "rows": 32768,
                                 "examples_per_s": 32768 / min(times0[1:]), "ms_per_pass": min(times0[1:]) * 1e3,
                                 "timing": "host wall clock around one hm_udtf_process call of the 32,768 rows from "
                                           "pageable host memory + sync, best of 5 after one warm-up"}

    # ---- N > 1: BASELINE.json's two multi-GPU configs, briefly ------------------------------
    secondary = None
    if world > 1 and not args.no_secondary and args.workload == "ftrl":
        secondary = {}
        for name in ("adam26", "fm64"):
            w2 = WORKLOADS[name]
            k2 = 2 * w2["mix_every"]
            a2 = DeviceArm(w2, MINI_BATCH, ROWS_PER_STEP, k2, w2["mix_every"], w2["total_rows"], world, rank, local_rank)
            r2 = a2.run()
            a2.close()
            secondary[name] = {"workload": w2["label"], "options": options_of(w2), "value": r2["value"],
                               "unit": "examples/s", "ms_per_step": r2["ms_per_step"], "steps": k2,
                               "warmup": w2["mix_every"], "rows_per_step": ROWS_PER_STEP, "mix": r2.get("mix")}

    # ---- CPU baseline (rank 0, N=1 only): a separate process, so that it maps oracle/ only -------------
    cpu = None
    if rank == 0 and world == 1 and not args.no_cpu:
        threads = os.cpu_count() or 1
        rows = args.cpu_sample_rows
        if rows <= 0:
            probe = cpu_baseline(2 * mb, threads, args.workload, mb)
            rows = int(max(2 * mb, min(probe["value"] * 15.0, 512 * mb)))
        rows = max(mb, rows // mb * mb)
        cpu = cpu_baseline(rows, threads, args.workload, mb)
        cpu.pop("seconds", None)

    if rank == 0:
        out = {
            "metric": wl["metric"],
            "value": rows_per_s, "unit": "examples/s", "n_gpus": world, "steps": K, "warmup": W,
            "ms_per_step": ms / K, "higher_is_better": True, "scaling": "weak", "vs_baseline": None,
            "dtype": "f32 (fp64 sums)", "data": "synthetic",
            "nnz_per_sec": rows_per_s * nnz_per_row,
            "config": config_of(wl, mb, rps, world, K, W, total_rows),
            "hot_features": {"count": len(hot_feats),
                             "note": "features picked per ingest block by their frequency in a sample of its rows: reduced by "
                                     "the forward kernel, not sorted (the choice affects speed only)"},
            "roofline": roofline,
            "cpu_baseline": cpu,
            "e2e": e2e,
            "gpu_launches": int(launches),
            "clocks": main_run["clocks"],
        }
        if main_run.get("mix"):
            out["mix_check"] = main_run["mix"]
        if sweep:
            out["mini_batch_sweep"] = sweep
        if secondary:
            out["secondary"] = secondary
        print(json.dumps(jsonable(out)))
    if world > 1:
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
