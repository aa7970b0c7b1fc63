"""hm_model_table_noalloc: the model table for callers that are not allowed to allocate
device memory (a task confined to a fixed share of a shared GPU, or code that cannot take the
device-wide stall of cudaMalloc/cudaFree next to other running work).

The entry point keeps its only scratch in an array compiled into the library, so what has to
hold is: same table as hm_udtf_forward (which allocates), call after call, although every call
reuses that one array and finds the previous call's leftovers in it.  This file is named so
that pytest collects it last.
"""
import numpy as np
import pytest

import hivemall_b200 as hm
from hivemall_b200 import synth, udtf

import _oracle

pytestmark = pytest.mark.gpu


def _trained(seed, dims, rows):
    data = synth.criteo_like(rows, dims=dims, seed=seed)
    rp, f, v, y = data.numpy()
    u = hm.train_classifier(f"-opt adagrad -mini_batch 2048 -dims {dims}")
    u.process_batch(rp, f, v, y)
    u.close()
    return u


def _noalloc_table(u, cap):
    import torch
    dims = u.config.dims
    w_dev = torch.from_numpy(u.array_to_host("w")).cuda()      # dense weights, one float per feature
    assert w_dev.numel() == dims
    feat = torch.full((cap,), -7, dtype=torch.int32, device="cuda")
    wt = torch.full((cap,), float("nan"), dtype=torch.float32, device="cuda")
    n_rows = torch.full((1,), -1, dtype=torch.int64, device="cuda")
    u.sync()
    torch.cuda.synchronize()
    udtf.model_table_noalloc(w_dev, dims, feat, wt, cap, n_rows, stream=0)
    torch.cuda.synchronize()
    return int(n_rows.item()), feat.cpu().numpy(), wt.cpu().numpy()


def test_two_calls_in_a_row_match_the_allocating_entry_point():
    # two different models of different sizes: the second call meets the first call's counters
    a = _trained(seed=21, dims=1 << 20, rows=20000)
    b = _trained(seed=22, dims=(1 << 18) + 123, rows=6000)
    for u in (a, b):
        want_f, want_w = u.forward()                       # existing entry point (allocates)
        assert len(want_f) > 100
        n, got_f, got_w = _noalloc_table(u, cap=len(want_f) + 50)
        assert n == len(want_f)
        assert np.array_equal(got_f[:n], want_f)           # bit-exact feature order
        assert np.array_equal(got_w[:n], want_w)           # weights are copies: bit-exact too
        assert (got_f[n:] == -7).all()                     # nothing written past the table
        # and against the CPU oracle's model table of the same dense weights (not only against the
        # library's own allocating path)
        ora_f, ora_w = _oracle.model_table(u.array_to_host("w"))
        assert np.array_equal(got_f[:n], ora_f) and np.array_equal(got_w[:n], ora_w)
    # and once more on the first model: leftovers of a smaller call do not leak into a larger one
    want_f, want_w = a.forward()
    n, got_f, got_w = _noalloc_table(a, cap=len(want_f))
    assert n == len(want_f) and np.array_equal(got_f, want_f) and np.array_equal(got_w, want_w)
    a.destroy()
    b.destroy()


def test_capacity_is_respected_and_limit_is_enforced():
    import torch
    u = _trained(seed=23, dims=1 << 16, rows=3000)
    want_f, _ = u.forward()
    n, got_f, _ = _noalloc_table(u, cap=10)
    assert n == len(want_f) > 10                           # true count reported, only `cap` rows written
    assert np.array_equal(got_f, want_f[:10])
    dummy = torch.zeros(8, device="cuda")
    cnt = torch.zeros(1, dtype=torch.int64, device="cuda")
    with pytest.raises(hm.UDFArgumentError) as e:
        udtf.model_table_noalloc(dummy, udtf.NOALLOC_MAX_DIMS + 1, dummy, dummy, 8, cnt)
    assert "exceeds" in str(e.value) or "outside" in str(e.value)
    with pytest.raises(hm.UDFArgumentError):
        udtf.model_table_noalloc(0, 16, dummy, dummy, 8, cnt)
    u.destroy()
