"""The MIX schedule shared by tests/test_gpu_multi.py and __graft_entry__.smoke(): two ranks, one GPU each,
train their own row shard and average over NCCL every INTERVAL mini-batches; the replicas must end
identical, and equal to a CPU-oracle replay of the same schedule within the weight tolerance.
Test infrastructure (it runs the oracle)."""
import os
import socket
import sys

import numpy as np

import hivemall_b200 as hm
from hivemall_b200 import mix, synth

import _oracle
from _util import assert_weights_close

OPTIONS = "-opt adagrad -mini_batch 256 -dims 65536"
ROWS, INTERVAL, WORLD = 8192, 4, 2


def _free_port() -> int:
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    p = s.getsockname()[1]
    s.close()
    return p


def _rank_main(rank, world, port, out_dir):
    import torch
    import torch.distributed as dist
    sys.path.insert(0, os.path.dirname(__file__))
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
    torch.cuda.set_device(rank)
    dist.init_process_group("nccl", rank=rank, world_size=world, device_id=torch.device("cuda", rank))
    cfg = hm.parse_options("train_classifier", OPTIONS)
    rp, f, v, y = synth.criteo_like(ROWS, dims=65536, seed=11).numpy()
    lo, hi = mix.shard_rows(ROWS, rank, world, cfg.mini_batch)
    u = hm.train_classifier(OPTIONS, device=rank)
    mixer = mix.ReplicaMixer(u, world)
    step_rows = INTERVAL * cfg.mini_batch
    for r0 in range(lo, hi, step_rows):
        r1 = min(hi, r0 + step_rows)
        u.process_batch(rp[r0:r1 + 1], f, v, y[r0:r1])
        mixer.mix()
    u.sync()
    np.save(os.path.join(out_dir, f"w{rank}.npy"), u.array_to_host("w"))
    np.save(os.path.join(out_dir, f"s{rank}.npy"), u.array_to_host("state"))
    u.destroy()
    dist.destroy_process_group()


def run_and_check(out_dir: str) -> int:
    """Spawn the two ranks, then replay the schedule on the CPU oracle.  Returns the number of non-zero weights."""
    import torch.multiprocessing as mp
    mp.start_processes(_rank_main, args=(WORLD, _free_port(), out_dir), nprocs=WORLD, join=True, start_method="spawn")
    w0, w1 = np.load(os.path.join(out_dir, "w0.npy")), np.load(os.path.join(out_dir, "w1.npy"))
    s0, s1 = np.load(os.path.join(out_dir, "s0.npy")), np.load(os.path.join(out_dir, "s1.npy"))
    assert np.array_equal(w0, w1) and np.array_equal(s0, s1), "replicas differ after a mix"
    cfg = _oracle.config_from(hm.parse_options("train_classifier", OPTIONS))
    rp, f, v, y = synth.criteo_like(ROWS, dims=65536, seed=11).numpy()
    shards = [mix.shard_rows(ROWS, r, WORLD, cfg.mini_batch) for r in range(WORLD)]
    W = [np.zeros(65536, np.float32) for _ in range(WORLD)]
    S = [np.zeros(65536, np.float32) for _ in range(WORLD)]
    T = [0] * WORLD
    step_rows = INTERVAL * cfg.mini_batch
    n_steps = (shards[0][1] - shards[0][0] + step_rows - 1) // step_rows
    for k in range(n_steps):
        for r, (lo, hi) in enumerate(shards):
            a = lo + k * step_rows
            b = min(hi, a + step_rows)
            res = _oracle.train_linear(cfg, rp[a:b + 1], f, v, y[a:b], w=W[r], s1=S[r], t=T[r])
            W[r], S[r], T[r] = res.w, res.s1, res.t
        mw = ((W[0] + W[1]) * np.float32(0.5)).astype(np.float32)
        ms = ((S[0] + S[1]) * np.float32(0.5)).astype(np.float32)
        W = [mw.copy() for _ in range(WORLD)]
        S = [ms.copy() for _ in range(WORLD)]
    assert_weights_close(w0, W[0])
    assert_weights_close(s0, S[0], "state")
    nz = int(np.count_nonzero(w0))
    assert nz > 1000
    return nz
