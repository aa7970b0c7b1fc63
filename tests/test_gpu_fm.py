"""Parity of the CUDA factorization machine (train_fm) with the CPU oracle, through the C-ABI.
PARITY UNPINNED w.r.t. Hivemall itself (no reference source is mounted); both sides state
DESIGN.md section 3.8."""
import numpy as np
import pytest

import hivemall_b200 as hm
from hivemall_b200 import synth

import _oracle
from _util import assert_weights_close

pytestmark = pytest.mark.gpu


def run_both(options, data, pushes=None):
    rp, f, v, y = data.numpy()
    u = hm.train_fm(options)
    n = len(y)
    r0 = 0
    for cnt in (pushes or [n]):
        r1 = min(n, r0 + cnt)
        u.process_batch(rp[r0:r1 + 1], f, v, y[r0:r1])
        r0 = r1
    u.close()
    cfg = _oracle.config_from(hm.parse_options("train_fm", options))
    ref = _oracle.train_fm(cfg, rp, f, v, y)
    return u, ref, cfg


def check(u, ref, adagrad):
    assert u.step == ref.t
    assert_weights_close(u.array_to_host("w0"), ref.w0, "w0")
    assert_weights_close(u.array_to_host("w"), ref.w, "w")
    assert_weights_close(u.array_to_host("V"), ref.V, "V")
    if adagrad:
        assert_weights_close(u.array_to_host("gw"), ref.gw, "gw")
        assert_weights_close(u.array_to_host("gV"), ref.gv, "gV")
        assert_weights_close(u.array_to_host("gw0"), ref.gw0, "gw0")


def test_initial_factors_are_bit_identical():
    u = hm.train_fm("-factors 7 -dims 1000 -sigma 0.3 -seed 5")
    cfg = _oracle.config_from(hm.parse_options("train_fm", "-factors 7 -dims 1000 -sigma 0.3 -seed 5"))
    assert np.array_equal(u.array_to_host("V"), _oracle.fm_init_v(cfg))
    assert not u.array_to_host("w").any()
    u.destroy()


@pytest.mark.parametrize("opt", ["sgd", "adagrad"])
@pytest.mark.parametrize("k", [4, 16])
def test_criteo_shaped_classification(opt, k):
    """configs[2] shape at oracle-sized scale: 39 feats/row, k factors, hot features that span
    many update chunks (mini-batch 4096 rows -> the 13 dense features have 4096 occurrences)."""
    data = synth.criteo_like(12000, dims=1 << 16, seed=3)
    options = f"-factors {k} -classification -opt {opt} -mini_batch 4096 -dims {1 << 16} -eta0 0.05"
    u, ref, _ = run_both(options, data, pushes=[5000, 7000])
    check(u, ref, opt == "adagrad")
    u.destroy()


def test_k64_regression_ragged():
    """configs[4] factor count; uneven rows incl. empty ones and duplicate features in a row."""
    data = synth.ragged(1500, dims=3000, max_nnz=30, seed=4, regression=True)
    options = "-factors 64 -opt adagrad -mini_batch 200 -dims 3000 -eta0 0.02 -iters 2"
    u, ref, _ = run_both(options, data)
    check(u, ref, True)
    u.destroy()


def test_model_table_and_prediction():
    data = synth.kdd12_like(6000, dims=1 << 14, seed=4)
    options = f"-factors 8 -classification -opt adagrad -mini_batch 500 -dims {1 << 14} -eta0 0.1 -iters 2"
    u, ref, cfg = run_both(options, data)
    feat, wt, vif = u.forward()
    seen = np.flatnonzero(ref.touched).astype(np.int32)
    assert feat[0] == -1 and np.array_equal(feat[1:], seen)          # w0 row, then ascending features seen
    assert_weights_close(wt[:1], ref.w0, "w0 row")
    assert_weights_close(wt[1:], ref.w[seen], "Wi")
    assert_weights_close(vif[1:].reshape(-1), ref.V.reshape(-1, 8)[seen].reshape(-1), "Vif")
    held = synth.kdd12_like(3000, dims=1 << 14, seed=4, row0=1 << 20)
    rp, f, v, y = held.numpy()
    import torch
    p_gpu = torch.from_numpy(u.predict(rp, f, v))
    p_ref = torch.from_numpy(_oracle.predict_fm(cfg, rp, f, v, ref.w0[0], ref.w, ref.V))
    yt = torch.from_numpy(y)
    assert abs(synth.auc(p_gpu, yt) - synth.auc(p_ref, yt)) < 1e-4
    assert abs(synth.logloss(p_gpu, yt) - synth.logloss(p_ref, yt)) < 1e-4
    u.destroy()


def test_fm_rerun_is_bit_identical():
    data = synth.criteo_like(9000, dims=1 << 15, seed=8)
    options = f"-factors 16 -classification -opt adagrad -mini_batch 3000 -dims {1 << 15}"
    a, _, _ = run_both(options, data)
    b, _, _ = run_both(options, data)
    assert np.array_equal(a.array_to_host("V"), b.array_to_host("V"))
    assert np.array_equal(a.array_to_host("w"), b.array_to_host("w"))


def test_too_many_factors_is_reported():
    with pytest.raises(hm.HivemallError) as e:
        hm.train_fm("-factors 128 -dims 100")
    assert e.value.status == 5
