"""Two GPUs (skipped on a one-GPU box): replicas on different devices of one process, and the
MIX step over NCCL -- two ranks that train their own row shard on their own GPU and average
every INTERVAL mini-batches end with identical replicas, equal to a CPU-oracle replay of the
same schedule within the weight tolerance."""
import numpy as np
import pytest

import hivemall_b200 as hm
from hivemall_b200 import _lib, synth

import _mix_schedule

pytestmark = pytest.mark.gpu


def _need_two():
    if _lib.load().hm_device_count() < 2:
        pytest.skip("needs two GPUs")


def test_two_devices_in_one_process_agree():
    _need_two()
    data = synth.criteo_like(6000, dims=1 << 16, seed=31)
    out = []
    for dev in (0, 1):
        for fn, opts in (("train_classifier", f"-opt ftrl -mini_batch 1000 -dims {1 << 16}"),
                         ("train_fm", f"-factors 8 -classification -opt adagrad -mini_batch 1000 -dims {1 << 16}")):
            u = hm.UDTF(opts, dev, function_name=fn)
            u.process_batch(*data.numpy())
            u.close()
            out.append((dev, fn, {name: u.array_to_host(name) for name in (("w", "state") if fn != "train_fm" else ("w", "V"))}))
            u.destroy()
    for (d0, f0, a0), (d1, f1, a1) in zip(out[:2], out[2:]):
        assert f0 == f1 and d0 != d1
        for name in a0:
            assert np.array_equal(a0[name], a1[name]), (f0, name)


def test_two_ranks_mix_over_nccl(tmp_path):
    _need_two()
    _mix_schedule.run_and_check(str(tmp_path))
