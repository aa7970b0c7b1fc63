"""MIX-style model averaging across GPUs, as NCCL all-reduce (sum, then scale by 1/P).

BASELINE.json's north_star: "the reference's MIX parameter averaging replaced by NCCL
allreduce over NVLink at the reference's mixing interval", each GPU holding a full model
replica and a contiguous shard of the rows.  The reference's MIX protocol itself cannot be
read (the mount has no source, README.md:20-22): the interval is the -mix_interval option
(mini-batches between averagings; 0 = once per epoch/step boundary chosen by the caller),
and EVERY replica array (weights and optimizer slots, V and its accumulators for train_fm)
is averaged so that the replicas stay identical after a mix.

torch.distributed is plumbing here (process group, NCCL communicator); the arrays are the
learner's own device memory, wrapped zero-copy.
"""
from __future__ import annotations

from typing import Sequence


def shard_rows(n_rows: int, rank: int, world: int, multiple_of: int = 1) -> tuple[int, int]:
    """Contiguous row shard [lo, hi) of rank `rank`: rows are split as evenly as possible
    in units of `multiple_of` rows (use the mini-batch size so that only the last rank can
    end on a short mini-batch)."""
    if world < 1 or not (0 <= rank < world):
        raise ValueError(f"bad rank/world {rank}/{world}")
    units = (n_rows + multiple_of - 1) // multiple_of
    lo_u = units * rank // world
    hi_u = units * (rank + 1) // world
    return min(n_rows, lo_u * multiple_of), min(n_rows, hi_u * multiple_of)


def mix_due(mini_batches_done: int, interval: int) -> bool:
    """True when an averaging is due after `mini_batches_done` local mini-batches."""
    return interval > 0 and mini_batches_done > 0 and mini_batches_done % interval == 0


def average_tensors(tensors: Sequence, world: int, group=None) -> None:
    """In-place average of each tensor over the process group (sum all-reduce, then 1/P).
    Works on any backend (nccl on GPUs; gloo in the CPU tests)."""
    import torch.distributed as dist
    if world == 1:
        return
    for t in tensors:
        dist.all_reduce(t, op=dist.ReduceOp.SUM, group=group)
        t.mul_(1.0 / world)


class ReplicaMixer:
    """Averages one learner's replica across ranks on the learner's own CUDA stream."""

    def __init__(self, udtf, world: int, group=None):
        import torch
        self.udtf = udtf
        self.world = world
        self.group = group
        self.stream = torch.cuda.ExternalStream(udtf.stream, device=f"cuda:{udtf.device}")
        dev = f"cuda:{udtf.device}"
        arrays = udtf.arrays()
        # float arrays are averaged; width 0 marks a byte map ("touched") that is OR-ed (max)
        self.tensors = [torch.as_tensor(udtf.array_view(name), device=dev) for name, a in arrays.items() if a[2] > 0]
        self.byte_maps = [torch.as_tensor(udtf.array_view(name), device=dev).view(torch.uint8)
                          for name, a in arrays.items() if a[2] == 0]
        self.bytes_per_mix = sum(t.numel() * 4 for t in self.tensors) + sum(t.numel() for t in self.byte_maps)
        self.mixes = 0
        self._events = []      # (start, end) CUDA events of the timed mixes

    def mix(self, timed: bool = False) -> None:
        """All-reduce every replica array, then scale by 1/P with the library's own kernel.
        Ordered after all training already queued on the learner's stream.  timed: bracket the
        exchange (all-reduce + scale) with its own CUDA event pair on that stream (timed_ms)."""
        import torch
        import torch.distributed as dist
        if self.world == 1:
            return
        ev = None
        if timed:
            ev = (torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True))
            with torch.cuda.stream(self.stream):
                ev[0].record()
        with torch.cuda.stream(self.stream):
            for t in self.tensors:
                dist.all_reduce(t, op=dist.ReduceOp.SUM, group=self.group)
            for t in self.byte_maps:
                dist.all_reduce(t, op=dist.ReduceOp.MAX, group=self.group)
        self.udtf.scale_replica(1.0 / self.world)
        if ev:
            with torch.cuda.stream(self.stream):
                ev[1].record()
            self._events.append(ev)
        self.mixes += 1

    def timed_ms(self) -> list:
        """Durations (ms) of the mixes run with timed=True so far; waits for them, then forgets them."""
        out = []
        for a, b in self._events:
            b.synchronize()
            out.append(a.elapsed_time(b))
        self._events = []
        return out
