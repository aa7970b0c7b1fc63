"""The online learners that share the mini-batch skeleton: Rosenblatt's
perceptron and the passive-aggressive family (Crammer, Dekel, Keshet, Shalev-Shwartz, Singer, JMLR 2006) --
the oracle's statement of them checked against what the papers promise.  CPU only.

PARITY UNPINNED w.r.t. Hivemall (no reference source is mounted): which learners its base class carries,
their option names and how they behave under -mini_batch are not known here; DESIGN.md 3.10 states
this library's contract.
"""
import numpy as np
import pytest

import hivemall_b200 as hm
from hivemall_b200 import synth

import _oracle


def one_row(fn, options, f, v, y, w0=None, dims=16):
    cfg = _oracle.config_from(hm.parse_options(fn, f"-dims {dims} " + options))
    rp = np.array([0, len(f)], dtype=np.int64)
    w = np.zeros(dims, np.float32) if w0 is None else np.asarray(w0, np.float32).copy()
    res = _oracle.train_linear(cfg, rp, np.asarray(f), np.asarray(v, np.float32), np.asarray([y], np.float32), w=w)
    return res.w


def test_pa_step_lands_on_the_margin():
    """PA: after the step the row's margin is exactly 1 (the smallest change of w that removes the loss)."""
    f, v = [1, 5, 9], [0.5, -2.0, 1.5]
    w = one_row("train_pa", "", f, v, 1.0)
    assert abs(float(np.dot(w[f], v)) - 1.0) < 1e-6
    w = one_row("train_pa", "", f, v, 0.0)          # label <= 0 is the negative class
    assert abs(float(np.dot(w[f], v)) + 1.0) < 1e-6


def test_pa1_caps_the_step_and_pa2_softens_it():
    f, v = [2, 3], [1.0, 1.0]                       # |x|^2 = 2, loss at w = 0 is 1: PA's tau = 0.5
    assert np.allclose(one_row("train_pa", "", f, v, 1.0)[f], 0.5)
    assert np.allclose(one_row("train_pa1", "-c 0.1", f, v, 1.0)[f], 0.1)          # min(C, 0.5)
    assert np.allclose(one_row("train_pa1", "-c 10", f, v, 1.0)[f], 0.5)
    assert np.allclose(one_row("train_pa2", "-c 1", f, v, 1.0)[f], 1.0 / (2 + 0.5))  # l / (|x|^2 + 1/(2C))


def test_no_loss_no_step():
    f, v = [0, 7], [1.0, 2.0]
    w0 = np.zeros(16, np.float32)
    w0[0], w0[7] = 1.0, 1.0                          # margin 3
    for fn in ("train_perceptron", "train_pa", "train_pa1", "train_pa2"):
        assert np.array_equal(one_row(fn, "", f, v, 1.0, w0), w0), fn
    # regression: inside the insensitive zone nothing moves, outside the step stops at its edge
    assert np.array_equal(one_row("train_pa1_regr", "-epsilon 0.5", f, v, 3.3, w0), w0)
    w = one_row("train_pa1_regr", "-epsilon 0.5 -c 100", f, v, 5.0, w0)
    assert abs(float(np.dot(w[f], v)) - 4.5) < 1e-6
    w = one_row("train_pa1_regr", "-epsilon 0.5 -c 100", f, v, 1.0, w0)
    assert abs(float(np.dot(w[f], v)) - 1.5) < 1e-6


def test_perceptron_moves_by_the_row_on_a_mistake():
    f, v = [4, 6], [2.0, -1.0]
    assert np.allclose(one_row("train_perceptron", "", f, v, 1.0)[f], v)           # margin 0 counts as a mistake
    assert np.allclose(one_row("train_perceptron", "", f, v, -1.0)[f], np.negative(v))


def test_empty_row_and_zero_row_are_harmless():
    for fn in ("train_pa", "train_pa1", "train_pa2", "train_pa2_regr"):
        assert not one_row(fn, "", [], [], 1.0).any()
        assert np.isfinite(one_row(fn, "", [3], [0.0], 1.0)).all()


@pytest.mark.parametrize("fn", ["train_perceptron", "train_pa", "train_pa1", "train_pa2"])
def test_classifiers_learn_a_separable_problem(fn):
    data = synth.a9a_like(4096, seed=1)
    rp, f, v, y = data.numpy()
    cfg = _oracle.config_from(hm.parse_options(fn, "-dims 124 -mini_batch 1 -iters 2"))
    res = _oracle.train_linear(cfg, rp, f, v, y)
    import torch
    assert synth.auc(torch.from_numpy(_oracle.predict_linear(rp, f, v, res.w)), torch.from_numpy(y)) > 0.7     # (noisy labels: the last iterate of an online learner wanders)


@pytest.mark.parametrize("fn", ["train_pa1_regr", "train_pa2_regr"])
def test_regressors_reduce_the_error(fn):
    data = synth.ragged(3000, dims=64, max_nnz=8, seed=4, regression=True)
    rp, f, v, y = data.numpy()
    cfg = _oracle.config_from(hm.parse_options(fn, "-dims 64 -mini_batch 1 -iters 3 -epsilon 0.01"))
    res = _oracle.train_linear(cfg, rp, f, v, y)
    err = np.abs(_oracle.predict_linear(rp, f, v, res.w) - y).mean()
    assert err < 0.5 * np.abs(y).mean()


@pytest.mark.parametrize("fn", ["train_pa1", "train_pa2_regr"])
def test_openmp_variant_is_bit_identical(fn):
    data = synth.criteo_like(2000, dims=1 << 12, seed=3)
    rp, f, v, y = data.numpy()
    cfg = _oracle.config_from(hm.parse_options(fn, f"-mini_batch 128 -dims {1 << 12} -iters 2 -c 0.5"))
    a = _oracle.train_linear(cfg, rp, f, v, y)
    b = _oracle.train_linear_mt(cfg, rp, f, v, y, 4)
    assert np.array_equal(a.w, b.w) and a.t == b.t


# ---- learners that keep a variance per feature: AROW (Crammer, Kulesza, Dredze 2009), diagonal --------------

def one_row_cov(fn, options, f, v, y, dims=16):
    cfg = _oracle.config_from(hm.parse_options(fn, f"-dims {dims} " + options))
    rp = np.array([0, len(f)], dtype=np.int64)
    res = _oracle.train_linear(cfg, rp, np.asarray(f), np.asarray(v, np.float32), np.asarray([y], np.float32))
    return res.w, res.s1


def test_arow_first_step_is_the_closed_form():
    """From w = 0, sigma = 1: v = |x|^2, beta = 1 / (v + r), alpha = beta; w_j = alpha y x_j,
    sigma_j = 1 - beta x_j^2; features the row does not have keep sigma = 1."""
    f, x, r = [1, 4, 9], np.array([0.5, -2.0, 1.0], np.float32), 0.5
    w, sg = one_row_cov("train_arow", f"-r {r}", f, x, 1.0)
    beta = 1.0 / (float((x * x).sum()) + r)
    assert np.allclose(w[f], beta * x, rtol=1e-6)
    assert np.allclose(sg[f], 1.0 - beta * x * x, rtol=1e-6)
    untouched = np.setdiff1d(np.arange(16), f)
    assert (sg[untouched] == 1.0).all() and not w[untouched].any()
    # the variances never grow and stay positive
    assert (sg <= 1.0).all() and (sg > 0.0).all()


def test_arow_no_loss_no_step_and_regression_always_steps():
    f, x = [2, 3], [1.0, 1.0]
    cfg = _oracle.config_from(hm.parse_options("train_arow", "-dims 8"))
    w0 = np.zeros(8, np.float32); w0[2] = 2.0
    res = _oracle.train_linear(cfg, np.array([0, 2]), np.asarray(f), np.asarray(x, np.float32), np.array([1.0], np.float32), w=w0,
                               s1=np.ones(8, np.float32))
    assert np.array_equal(res.w, w0) and (res.s1 == 1.0).all()          # margin 2 >= 1: untouched
    w, sg = one_row_cov("train_arow_regr", "-r 1", f, x, 0.3, dims=8)
    assert abs(float(w[2] + w[3]) - 0.3 * 2 / 3) < 1e-6 and (sg[[2, 3]] < 1.0).all()   # e beta |x|^2 = 0.3 * 2 / (2 + 1)
    w, sg = one_row_cov("train_arowe_regr", "-r 1 -epsilon 0.5", f, x, 0.3, dims=8)
    assert not w.any() and (sg == 1.0).all()                            # inside the insensitive zone


def test_arow_learns_and_confidence_grows_with_evidence():
    data = synth.a9a_like(4096, seed=1)
    rp, f, v, y = data.numpy()
    cfg = _oracle.config_from(hm.parse_options("train_arow", "-dims 124 -mini_batch 1 -r 1"))
    res = _oracle.train_linear(cfg, rp, f, v, y)
    import torch
    assert synth.auc(torch.from_numpy(_oracle.predict_linear(rp, f, v, res.w)), torch.from_numpy(y)) > 0.8
    counts = np.bincount(f, minlength=124)
    seen = counts > 0
    assert (res.s1[~seen] == 1.0).all() and (res.s1[seen] < 1.0).all() and (res.s1 > 0).all()
    # frequent features end up with smaller variances than rare ones
    order = np.argsort(counts[seen])
    sg = res.s1[seen][order]
    assert sg[: len(sg) // 4].mean() > sg[-len(sg) // 4:].mean()


# ---- confidence-weighted / soft confidence-weighted (diagonal) ---------------------------------------------

@pytest.mark.parametrize("fn", ["train_cw", "train_scw", "train_scw2"])
def test_confidence_weighted_first_step(fn):
    """From w = 0, sigma = 1 the row is uncertain (margin 0 < phi sqrt(v)): a step along y x, variances of the
    row's features shrink, nothing else moves; SCW-I's step is capped by C."""
    f, x = [1, 4, 9], np.array([0.5, -2.0, 1.0], np.float32)
    w, sg = one_row_cov(fn, "", f, x, 1.0)
    assert (np.sign(w[f]) == np.sign(x)).all()
    assert (sg[f] < 1.0).all() and (sg[f] > 0.0).all()
    untouched = np.setdiff1d(np.arange(16), f)
    assert (sg[untouched] == 1.0).all() and not w[untouched].any()
    wn, _ = one_row_cov(fn, "", f, x, 0.0)                 # the negative class mirrors it
    assert np.allclose(wn, -w)
    if fn == "train_scw":
        wc, _ = one_row_cov(fn, "-c 0.01", f, x, 1.0)
        assert np.allclose(wc[f], 0.01 * x, rtol=1e-6)    # alpha = C, sigma = 1


@pytest.mark.parametrize("fn", ["train_cw", "train_scw", "train_scw2"])
def test_confident_and_correct_rows_do_not_step(fn):
    cfg = _oracle.config_from(hm.parse_options(fn, "-dims 8 -phi 0.5"))
    w0 = np.zeros(8, np.float32); w0[2] = 5.0
    sg0 = np.full(8, 0.01, np.float32)
    res = _oracle.train_linear(cfg, np.array([0, 2]), np.array([2, 3]), np.array([1.0, 1.0], np.float32),
                               np.array([1.0], np.float32), w=w0, s1=sg0)
    assert np.array_equal(res.w, w0) and np.array_equal(res.s1, sg0)


@pytest.mark.parametrize("fn,opts", [("train_cw", "-phi 0.5"), ("train_scw", "-c 0.1"), ("train_scw2", "-c 0.1")])
def test_confidence_weighted_learns(fn, opts):
    data = synth.a9a_like(4096, seed=1)
    rp, f, v, y = data.numpy()
    cfg = _oracle.config_from(hm.parse_options(fn, f"-dims 124 -mini_batch 1 {opts}"))
    res = _oracle.train_linear(cfg, rp, f, v, y)
    import torch
    # (hard-margin CW is known to suffer from label noise -- the reason SCW exists; the soft variants do better)
    assert synth.auc(torch.from_numpy(_oracle.predict_linear(rp, f, v, res.w)), torch.from_numpy(y)) > (0.7 if fn == "train_cw" else 0.78)
    assert (res.s1 > 0).all() and (res.s1 <= 1.0).all() and np.isfinite(res.w).all()
