"""Host-side logic of the N>1 path on CPU: row sharding, mixing schedule and the
sum-then-scale averaging, with two real processes over gloo (127.0.0.1).

The compute inside each rank is the CPU oracle (this is a test; the product has no CPU
path): what is checked is that two ranks that train on their own shard and average at the
mixing interval end with identical replicas, equal to a single-process replay of the same
schedule."""
import os
import socket
import sys

import numpy as np
import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

import hivemall_b200 as hm
from hivemall_b200 import mix, synth

import _oracle

OPTIONS = "-opt adagrad -mini_batch 64 -dims 4096"
ROWS, INTERVAL, WORLD = 1024, 4, 2


def test_shard_rows_partitions_in_minibatch_units():
    for n, world, mb in [(1000, 3, 64), (64, 4, 64), (0, 2, 8), (12345, 8, 128), (10, 2, 1)]:
        cuts = [mix.shard_rows(n, r, world, mb) for r in range(world)]
        assert cuts[0][0] == 0 and cuts[-1][1] == n
        for (a, b), (c, d) in zip(cuts, cuts[1:]):
            assert b == c and a <= b
        for a, b in cuts[:-1]:
            assert (b - a) % mb == 0          # only the last rank may end on a short mini-batch
    with pytest.raises(ValueError):
        mix.shard_rows(10, 2, 2)


def test_mix_due():
    assert [mix.mix_due(k, 4) for k in range(9)] == [False, False, False, False, True, False, False, False, True]
    assert not mix.mix_due(5, 0)


def _free_port() -> int:
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    p = s.getsockname()[1]
    s.close()
    return p


def _train_segment(cfg, data, lo, hi, w, s1, t):
    rp, f, v, y = data
    res = _oracle.train_linear(cfg, rp[lo:hi + 1], f, v, y[lo:hi], w=w, s1=s1, t=t)
    return res.w, res.s1, res.t


def _rank_main(rank, world, port, out_dir):
    sys.path.insert(0, os.path.dirname(__file__))
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
    dist.init_process_group("gloo", rank=rank, world_size=world)
    cfg = _oracle.config_from(hm.parse_options("train_classifier", OPTIONS))
    data = synth.criteo_like(ROWS, dims=4096, seed=11).numpy()
    lo, hi = mix.shard_rows(ROWS, rank, world, cfg.mini_batch)
    w = np.zeros(4096, np.float32); s1 = np.zeros(4096, np.float32); t = 0
    step_rows = INTERVAL * cfg.mini_batch
    done = 0
    for r0 in range(lo, hi, step_rows):
        w, s1, t = _train_segment(cfg, data, r0, min(hi, r0 + step_rows), w, s1, t)
        done += INTERVAL
        if mix.mix_due(done, INTERVAL):
            tw, ts = torch.from_numpy(w), torch.from_numpy(s1)
            mix.average_tensors([tw, ts], world)          # in place: sum all-reduce, then 1/P
    np.save(os.path.join(out_dir, f"w{rank}.npy"), w)
    np.save(os.path.join(out_dir, f"s{rank}.npy"), s1)
    dist.destroy_process_group()


def test_two_ranks_average_like_a_single_process_replay(tmp_path):
    mp.start_processes(_rank_main, args=(WORLD, _free_port(), str(tmp_path)), nprocs=WORLD, join=True,
                       start_method="spawn")
    w0, w1 = np.load(tmp_path / "w0.npy"), np.load(tmp_path / "w1.npy")
    s0, s1 = np.load(tmp_path / "s0.npy"), np.load(tmp_path / "s1.npy")
    assert np.array_equal(w0, w1) and np.array_equal(s0, s1)       # replicas identical after a mix
    # single-process replay of the same schedule
    cfg = _oracle.config_from(hm.parse_options("train_classifier", OPTIONS))
    data = synth.criteo_like(ROWS, dims=4096, seed=11).numpy()
    shards = [mix.shard_rows(ROWS, r, WORLD, cfg.mini_batch) for r in range(WORLD)]
    W = [np.zeros(4096, np.float32) for _ in range(WORLD)]
    S = [np.zeros(4096, np.float32) for _ in range(WORLD)]
    T = [0] * WORLD
    step_rows = INTERVAL * cfg.mini_batch
    n_steps = (shards[0][1] - shards[0][0] + step_rows - 1) // step_rows
    for k in range(n_steps):
        for r, (lo, hi) in enumerate(shards):
            a = lo + k * step_rows
            W[r], S[r], T[r] = _train_segment(cfg, data, a, min(hi, a + step_rows), W[r], S[r], T[r])
        mw = ((W[0] + W[1]) * np.float32(0.5)).astype(np.float32)
        ms = ((S[0] + S[1]) * np.float32(0.5)).astype(np.float32)
        W = [mw.copy() for _ in range(WORLD)]
        S = [ms.copy() for _ in range(WORLD)]
    assert np.array_equal(w0, W[0]) and np.array_equal(s0, S[0])
    assert np.count_nonzero(w0) > 100
