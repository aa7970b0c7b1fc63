"""The online learners on the CUDA path against the oracle: perceptron and the passive-aggressive family run
through the same forward / sort / update kernels as train_classifier, with the step size in the per-row
coefficient (hm_linear.cu: online_coef, row_sqnorm_kernel).  Every launch shape is covered: the persistent
small-batch kernel (-mini_batch 1, 128), the two-kernel path with and without hot features, rows worked on
by half a warp, several pushes, replays.

PARITY UNPINNED w.r.t. Hivemall (no reference source is mounted); checked against oracle/.
"""
import numpy as np
import pytest

import hivemall_b200 as hm
from hivemall_b200 import synth

from _util import assert_weights_close, run_gpu_linear, run_oracle_linear

pytestmark = pytest.mark.gpu

CLASSIFIERS = ["train_perceptron", "train_pa", "train_pa1", "train_pa2"]
REGRESSORS = ["train_pa1_regr", "train_pa2_regr"]


def check(fn, options, data, pushes=None):
    u, w, _ = run_gpu_linear(fn, options, data, pushes=pushes)
    ref, _ = run_oracle_linear(fn, options, data)
    assert u.step == ref.t
    assert_weights_close(w, ref.w)
    assert np.count_nonzero(ref.w) > 0
    u.destroy()
    return w


@pytest.mark.parametrize("fn", CLASSIFIERS)
@pytest.mark.parametrize("mb", [1, 128])
def test_small_batches_a9a_shaped(fn, mb):
    data = synth.a9a_like(2048 + 37, seed=2)
    extra = " -c 0.1" if fn in ("train_pa1", "train_pa2") else ""
    check(fn, f"-dims 124 -mini_batch {mb} -iters 2{extra}", data)


@pytest.mark.parametrize("fn", REGRESSORS)
@pytest.mark.parametrize("mb", [1, 100])
def test_small_batches_regression(fn, mb):
    data = synth.ragged(3000, dims=256, max_nnz=12, seed=5, regression=True)
    check(fn, f"-dims 256 -mini_batch {mb} -c 0.5 -epsilon 0.05", data)


@pytest.mark.parametrize("fn", CLASSIFIERS + REGRESSORS)
@pytest.mark.parametrize("hot", ["1", "0"])
def test_two_kernel_path_criteo_shaped(monkeypatch, fn, hot):
    """Mini-batches of 4096 Criteo-shaped rows: quantitative columns with real values (|x|^2 is not the row
    length), hot features reduced in the forward kernel."""
    monkeypatch.setenv("HM_HOT_FEATURES", hot)
    dims = 1 << 18
    data = synth.criteo_like(3 * 4096 + 555, dims=dims, seed=7)
    if fn in REGRESSORS:
        rng = np.random.default_rng(1)
        import torch
        data.label = data.label * 2.0 + torch.from_numpy(rng.normal(0, 0.3, data.rows).astype(np.float32))
    norm = " -grad_norm count" if fn in ("train_pa", "train_pa2_regr") else ""
    check(fn, f"-dims {dims} -mini_batch 4096{norm}", data, pushes=[5000, 4096 * 2, 10000])


@pytest.mark.parametrize("lpr", ["16", "8"])
def test_rows_side_by_side(monkeypatch, lpr):
    monkeypatch.setenv("HM_FWD_LPR", lpr)
    dims = 1 << 16
    data = synth.criteo_like(2 * 2048 + 100, dims=dims, seed=9)
    check("train_pa1", f"-dims {dims} -mini_batch 2048 -c 0.2", data)
    monkeypatch.setenv("HM_HOT_FEATURES", "0")
    check("train_pa2", f"-dims {dims} -mini_batch 2048 -c 0.2", data)


def test_out_of_range_elements_do_not_count_in_the_norm():
    """An index >= dims is ignored everywhere, also in |x|^2: the model is what the oracle gets for the rows
    without that element (and close() reports it)."""
    data = synth.ragged(4096, dims=512, max_nnz=10, seed=11)
    rowptr, feature, value, label = data.numpy()
    bad_f = feature.copy()
    r = 1234 + int(np.argmax(np.diff(rowptr[1234:]) > 1))       # a row with at least two elements
    assert rowptr[r + 1] - rowptr[r] > 1
    p = int(rowptr[r])
    bad_f[p] = 512 + 9
    options = "-dims 512 -mini_batch 2048"
    u = hm.UDTF(options, 0, function_name="train_pa")
    u.process_batch(rowptr, bad_f, value, label)
    with pytest.raises(hm.UDFArgumentError):
        u.close()
    w = u.array_to_host("w")
    # the same rows with that element removed
    keep = np.ones(len(feature), bool)
    keep[p] = False
    rp2 = rowptr.copy()
    rp2[r + 1:] -= 1
    clean = synth.Csr(*(__import__("torch").from_numpy(a) for a in (rp2, feature[keep], value[keep], label)))
    ref, _ = run_oracle_linear("train_pa", options, clean)
    assert_weights_close(w, ref.w)
    u.destroy()


def test_replays_and_determinism():
    dims = 1 << 16
    data = synth.criteo_like(3 * 2048, dims=dims, seed=13)
    options = f"-dims {dims} -mini_batch 2048 -iters 3 -c 0.05"
    w1 = check("train_pa1", options, data)
    u, w2, _ = run_gpu_linear("train_pa1", options, data)
    assert np.array_equal(w1, w2)
    u.destroy()


def test_model_table_and_prediction():
    data = synth.a9a_like(4096, seed=3)
    rowptr, feature, value, label = data.numpy()
    with hm.train_pa1("-dims 124 -mini_batch 64 -c 0.05") as u:
        u.process_batch(rowptr, feature, value, label)
        u.close()
        feat, wt = u.forward()[:2]
        w = u.array_to_host("w")
        assert np.all(np.diff(feat) > 0) and np.array_equal(wt, w[feat]) and set(np.flatnonzero(w)) <= set(feat.tolist())
        import torch
        assert synth.auc(torch.from_numpy(u.predict(rowptr, feature, value)), torch.from_numpy(label)) > 0.7


# ---- learners that keep a variance per feature (AROW): forward with margin + variance, two update launches ---

COV = ["train_arow", "train_arow_regr", "train_arowe_regr", "train_cw", "train_scw", "train_scw2"]
COV_REGR = ["train_arow_regr", "train_arowe_regr"]
COV_OPTS = {"train_arow": "-r 1", "train_arow_regr": "-r 2", "train_arowe_regr": "-r 2 -epsilon 0.05",
            "train_cw": "-phi 0.5", "train_scw": "-phi 0.8 -c 0.1", "train_scw2": "-c 0.1"}


def check_cov(fn, options, data, pushes=None):
    u, w, state = run_gpu_linear(fn, options, data, pushes=pushes)
    ref, ref_state = run_oracle_linear(fn, options, data)
    assert u.step == ref.t and u.state_width == 1
    assert_weights_close(w, ref.w)
    assert_weights_close(state, ref_state, "variance")
    assert np.count_nonzero(ref.w) > 0 and (ref_state <= 1.0).all() and (ref_state < 1.0).any()
    u.destroy()
    return w, state


@pytest.mark.parametrize("fn", COV)
@pytest.mark.parametrize("mb", [1, 64])
def test_cov_small_batches(fn, mb):
    if fn not in COV_REGR:
        data = synth.a9a_like(1024 + 11, seed=4)
        check_cov(fn, f"-dims 124 -mini_batch {mb} {COV_OPTS[fn]}", data)
    else:
        data = synth.ragged(1500, dims=256, max_nnz=12, seed=6, regression=True)
        check_cov(fn, f"-dims 256 -mini_batch {mb} {COV_OPTS[fn]}", data)


@pytest.mark.parametrize("fn", COV)
def test_cov_criteo_shaped_every_tile_size(fn):
    """Criteo-shaped rows with real-valued quantitative columns (x^2 matters), mini-batches of 1024, 8192 and
    20,000 rows: the three tile sizes of the update kernel, several pushes, a ragged tail."""
    dims = 1 << 18
    for mb, n in ((1024, 5 * 1024 + 77), (8192, 2 * 8192 + 1000), (20000, 45000)):
        data = synth.criteo_like(n, dims=dims, seed=8)
        if fn in COV_REGR:
            import torch
            rng = np.random.default_rng(2)
            data.label = data.label * 2.0 + torch.from_numpy(rng.normal(0, 0.3, data.rows).astype(np.float32))
        norm = " -grad_norm count" if mb == 8192 else ""
        check_cov(fn, f"-dims {dims} -mini_batch {mb} {COV_OPTS[fn]}{norm}", data, pushes=[n // 3, n])


def test_cov_determinism_replays_and_table():
    dims = 1 << 16
    data = synth.criteo_like(3 * 4096, dims=dims, seed=10)
    options = f"-dims {dims} -mini_batch 4096 -iters 2 -r 0.5"
    w1, s1 = check_cov("train_arow", options, data)
    u, w2, s2 = run_gpu_linear("train_arow", options, data)
    assert np.array_equal(w1, w2) and np.array_equal(s1, s2)
    feat, wt = u.forward()[:2]
    assert np.all(np.diff(feat) > 0) and np.array_equal(wt, w2[feat])
    rowptr, feature, value, label = data.numpy()
    import torch
    assert synth.auc(torch.from_numpy(u.predict(rowptr, feature, value)), torch.from_numpy(label)) > 0.52    # (six averaged steps: a weak model, better than chance)
    u.destroy()
