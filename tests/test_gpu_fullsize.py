"""BASELINE.json-sized runs (2^24 hashed dims, -mini_batch 65536, Criteo-shaped rows) checked
through size-independent properties, since the CPU oracle would take too long at this size:
determinism, host path == device path bit for bit, model-table order, a conserved quantity of
the update rule, and that training actually reduces the loss.  A slice small enough for the
oracle is still compared directly."""
import numpy as np
import pytest

import hivemall_b200 as hm
from hivemall_b200 import synth

import _oracle
from _util import assert_weights_close

pytestmark = pytest.mark.gpu
DIMS = 1 << 24
MB = 65536


def _device_rows(n_rows, seed=1):
    import torch
    parts = [synth.criteo_like(min(1 << 20, n_rows - r), dims=DIMS, seed=seed, device="cuda", row0=r)
             for r in range(0, n_rows, 1 << 20)]
    return synth.Csr(torch.arange(n_rows + 1, device="cuda", dtype=torch.int64) * 39,
                     torch.cat([p.feature for p in parts]), torch.cat([p.value for p in parts]),
                     torch.cat([p.label for p in parts]))


def test_config1_shape_properties():
    import torch
    n = 32 * MB                                   # 2.1 M rows, 82 M elements, 32 mini-batches
    data = _device_rows(n)
    opts = f"-loss logloss -opt ftrl -mini_batch {MB} -dims {DIMS} -alpha 0.05 -lambda1 0.0001"
    a = hm.train_classifier(opts)
    a.process_device(n, data.nnz, data.rowptr, data.feature, data.value, data.label)
    rows_a = a.close()
    assert a.step == 32
    # (1) same rows through the host-buffer path, pushed in uneven pieces: bit-identical replica
    host = [t.cpu().numpy() for t in (data.rowptr, data.feature, data.value, data.label)]
    b = hm.train_classifier(opts)
    cuts = [0, 100001, 7 * MB + 5, 19 * MB, n]
    for lo, hi in zip(cuts, cuts[1:]):
        b.process_batch(host[0][lo:hi + 1], host[1], host[2], host[3][lo:hi])
    assert b.close() == rows_a
    wa, wb = a.array_to_host("w"), b.array_to_host("w")
    assert np.array_equal(wa, wb)
    assert np.array_equal(a.array_to_host("state"), b.array_to_host("state"))
    # (2) model table: ascending, exactly the non-zero weights
    feat, wt = a.forward()
    assert len(feat) == rows_a == np.count_nonzero(wa)
    assert np.all(np.diff(feat.astype(np.int64)) > 0) and np.array_equal(wt, wa[feat])
    # (3) FTRL invariant: n_j = sum of squared mean gradients >= 0, and z, n untouched where the
    # feature never occurred
    st = a.array_to_host("state").reshape(-1, 2)
    seen = np.zeros(DIMS, bool)
    seen[np.unique(host[1])] = True
    assert (st[:, 1] >= 0).all() and not st[~seen].any() and not wa[~seen].any()
    assert (st[seen, 1] > 0).mean() > 0.99
    # (4) it learns: held-out logloss below the untrained 0.693 and AUC well above chance
    held = synth.criteo_like(200000, dims=DIMS, seed=1, row0=1 << 30)
    rp, f, v, y = held.numpy()
    m = torch.from_numpy(a.predict(rp, f, v))
    assert synth.logloss(m, torch.from_numpy(y)) < 0.68
    assert synth.auc(m, torch.from_numpy(y)) > 0.6
    # (5) the first 3 mini-batches against the oracle directly
    k = 3 * MB
    c = hm.train_classifier(opts)
    c.process_batch(host[0][:k + 1], host[1], host[2], host[3][:k])
    c.close()
    ref = _oracle.train_linear_mt(_oracle.config_from(hm.parse_options("train_classifier", opts)),
                                  host[0][:k + 1], host[1], host[2], host[3][:k], 8)
    assert_weights_close(c.array_to_host("w"), ref.w)
    for u in (a, b, c):
        u.destroy()


def test_config3_shape_adam_2pow26_dims():
    """configs[3] shape on one GPU: squared loss + Adam, 2^26 dims (1 GiB replica), rerun identical."""
    n = 8 * MB
    import torch
    data = synth.criteo_like(n, dims=1 << 26, seed=3, device="cuda", regression=True)
    opts = f"-loss squared -opt adam -mini_batch {MB} -dims {1 << 26} -eta0 0.001"
    ws = []
    for _ in range(2):
        u = hm.train_regressor(opts)
        u.process_device(n, data.nnz, data.rowptr, data.feature, data.value, data.label)
        u.close()
        feat, wt = u.forward()
        ws.append((feat, wt))
        assert np.all(np.diff(feat.astype(np.int64)) > 0) and len(feat) > 100000
        u.destroy()
    assert np.array_equal(ws[0][0], ws[1][0]) and np.array_equal(ws[0][1], ws[1][1])
    torch.cuda.synchronize()
