"""BASELINE.json-sized runs (2^24 / 2^26 hashed dims, -mini_batch 65536, Criteo- and KDD12-shaped rows):
size-independent properties (determinism, host path == device path bit for bit, model-table order, a
conserved quantity of the update rule, training reduces the loss) AND the oracle itself on the same rows --
all 32 mini-batches of the configs[1] run, the configs[3] run (Adam, 2^26 dims), and two mini-batches of
train_fm at the configs[2] (k = 16, Criteo-shaped) and configs[4] (k = 64, KDD12-shaped) sizes."""
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
    # (5) the whole run -- all 32 mini-batches, hot features and all -- against the oracle
    import os
    ref = _oracle.train_linear_mt(_oracle.config_from(hm.parse_options("train_classifier", opts)),
                                  host[0], host[1], host[2], host[3], min(16, os.cpu_count() or 1))
    assert ref.t == 32
    assert_weights_close(wa, ref.w)
    assert_weights_close(a.array_to_host("state"), ref.state(2), "state")
    for u in (a, b):
        u.destroy()


def test_config3_shape_adam_2pow26_dims():
    """configs[3] shape on one GPU: squared loss + Adam, 2^26 dims (1 GiB replica): rerun identical, and the
    oracle on the same 8 mini-batches."""
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
    import os
    rp, f, v, y = data.numpy()
    ref = _oracle.train_linear_mt(_oracle.config_from(hm.parse_options("train_regressor", opts)), rp, f, v, y,
                                  min(16, os.cpu_count() or 1))
    ofeat, owt = _oracle.model_table(ref.w)
    assert np.array_equal(ws[0][0], ofeat)                 # the same features carry a weight, in the same order
    assert_weights_close(ws[0][1], owt, "model table weights")


@pytest.mark.parametrize("gen,k,nnz", [("criteo", 16, 39), ("kdd12", 64, 11)])
def test_train_fm_at_baseline_dims(gen, k, nnz):
    """configs[2] / configs[4] shapes: train_fm AdaGrad at 2^24 dims, -mini_batch 65536, two mini-batches against
    the (scalar) oracle: w0, w, V and their accumulators on every feature the rows touch."""
    n = 2 * MB
    data = (synth.criteo_like(n, dims=DIMS, seed=2) if gen == "criteo" else synth.kdd12_like(n, dims=DIMS, seed=4))
    rp, f, v, y = data.numpy()
    opts = f"-factors {k} -classification -opt adagrad -mini_batch {MB} -dims {DIMS} -eta0 0.05"
    u = hm.train_fm(opts)
    u.process_batch(rp, f, v, y)
    u.close()
    ref = _oracle.train_fm(_oracle.config_from(hm.parse_options("train_fm", opts)), rp, f, v, y)
    assert u.step == ref.t == 2
    touched = np.unique(f)
    assert_weights_close(u.array_to_host("w0"), ref.w0, "w0")
    assert_weights_close(u.array_to_host("w")[touched], ref.w[touched], "w")
    V = u.array_to_host("V").reshape(-1, k)
    assert_weights_close(V[touched].reshape(-1), ref.V.reshape(-1, k)[touched].reshape(-1), "V")
    gV = u.array_to_host("gV").reshape(-1, k)
    assert_weights_close(gV[touched].reshape(-1), ref.gv.reshape(-1, k)[touched].reshape(-1), "gV")
    # untouched features keep their initial factors bit for bit
    rest = np.setdiff1d(np.arange(0, DIMS, 4099), touched)
    assert np.array_equal(V[rest], ref.V.reshape(-1, k)[rest])
    u.destroy()


@pytest.mark.parametrize("fn,extra", [("train_pa1", "-c 0.05"), ("train_pa2_regr", "-c 0.05 -epsilon 0.02")])
def test_online_learners_at_the_configs1_size(fn, extra):
    """The online learners of DESIGN.md 3.10 on 65,536-row mini-batches of Criteo-shaped rows at 2^24 dims: the
    launch shape of the headline (rows worked on by half a warp, hot features reduced in the forward kernel,
    |x|^2 from the per-block kernel) against the oracle, eight mini-batches, and a bit-identical rerun."""
    import os
    n = 8 * MB
    data = _device_rows(n, seed=3)
    host = [t.cpu().numpy() for t in (data.rowptr, data.feature, data.value, data.label)]
    opts = f"-mini_batch {MB} -dims {DIMS} {extra}"
    ws = []
    for _ in range(2):
        u = hm.UDTF(opts, 0, function_name=fn)
        u.process_device(n, data.nnz, data.rowptr, data.feature, data.value, data.label)
        u.close()
        assert u.step == 8 and len(u.hot_features()) >= 20
        ws.append(u.array_to_host("w"))
        u.destroy()
    assert np.array_equal(ws[0], ws[1])
    ref = _oracle.train_linear_mt(_oracle.config_from(hm.parse_options(fn, opts)), host[0], host[1], host[2], host[3],
                                  min(16, os.cpu_count() or 1))
    assert ref.t == 8 and np.count_nonzero(ref.w) > 100000
    assert_weights_close(ws[0], ref.w)
