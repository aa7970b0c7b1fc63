"""Hot features: the few features that carry a large share of a mini-batch's elements (the
quantitative features every row has, the head of a categorical field) are picked per ingest block
from a sample of its rows, reduced by the forward kernel and never enter the sort (hm_sort.cu:
hot_select_kernel, hm_linear.cu: forward_rows_hot / hot_features_step).  The result must be what
the oracle computes for the plain rows -- the oracle knows nothing about hot features -- whatever
the rows look like: a clean prefix, the same feature again later in a row or twice in one row,
rows without it, empty rows, long rows.

PARITY UNPINNED w.r.t. Hivemall (no reference source is mounted); checked against oracle/.
"""
import os

import numpy as np
import pytest
import torch

import hivemall_b200 as hm
from hivemall_b200 import synth

from _util import assert_weights_close, run_gpu_linear, run_oracle_linear

pytestmark = pytest.mark.gpu

DIMS = 1 << 16
PREFIX = [7, 40000, 123, 65535, 9, 31000, 2]


def prefix_rows(n_rows, prefix=PREFIX, seed=3, tail_max=24, regression=False):
    """Rows = `prefix` features with real values, then 0..tail_max other features (duplicates inside
    a row allowed, none of them equal to a prefix feature)."""
    rng = np.random.default_rng(seed)
    tails = rng.integers(0, tail_max + 1, n_rows)
    rows_f, rows_v = [], []
    banned = set(prefix)
    for r in range(n_rows):
        t = rng.integers(0, DIMS, tails[r])
        t = np.array([x if int(x) not in banned else (int(x) + 1) % DIMS for x in t], dtype=np.int64)
        t = np.array([x if int(x) not in banned else 1000 for x in t], dtype=np.int64)
        rows_f.append(np.concatenate([np.asarray(prefix, dtype=np.int64), t]))
        rows_v.append(np.concatenate([rng.normal(0, 1, len(prefix)), np.ones(len(t))]))
    return rows_f, rows_v, rng


def to_csr(rows_f, rows_v, rng, regression=False):
    lens = np.array([len(f) for f in rows_f])
    rowptr = np.zeros(len(rows_f) + 1, dtype=np.int64)
    rowptr[1:] = np.cumsum(lens)
    feature = np.concatenate(rows_f).astype(np.int32)
    value = np.concatenate(rows_v).astype(np.float32)
    w_true = rng.normal(0, 0.3, DIMS)
    m = np.array([float(np.dot(w_true[f[f < DIMS]], v[f < DIMS])) for f, v in zip(rows_f, rows_v)])
    if regression:
        label = m.astype(np.float32)
    else:
        label = (rng.random(len(rows_f)) < 1 / (1 + np.exp(-m))).astype(np.float32)
    return synth.Csr(torch.from_numpy(rowptr), torch.from_numpy(feature), torch.from_numpy(value),
                     torch.from_numpy(label))


def check(fn, options, data, want_hot, pushes=None):
    """want_hot: the hot features of the LAST ingest block (that is what hot_features() reports)."""
    u, w, state = run_gpu_linear(fn, options, data, pushes=pushes)
    got = sorted(int(x) for x in u.hot_features())
    assert got == sorted(want_hot), f"hot features {got}, expected {sorted(want_hot)}"
    ref, ref_state = run_oracle_linear(fn, options, data)
    assert u.step == ref.t
    assert_weights_close(w, ref.w)
    assert_weights_close(state, ref_state, "state")
    u.destroy()


@pytest.mark.parametrize("opt", ["sgd", "adagrad", "adadelta", "adam", "ftrl"])
def test_clean_prefix_every_optimizer(opt):
    rows_f, rows_v, rng = prefix_rows(6000 + 77)
    data = to_csr(rows_f, rows_v, rng)
    check("train_classifier", f"-loss logloss -opt {opt} -mini_batch 2048 -dims {DIMS} -eta0 0.05", data, PREFIX)


@pytest.mark.parametrize("norm", ["count", "sum"])
def test_gradient_normalisations(norm):
    rows_f, rows_v, rng = prefix_rows(5000, seed=5, regression=True)
    data = to_csr(rows_f, rows_v, rng, regression=True)
    check("train_regressor", f"-loss squared -opt adagrad -mini_batch 2048 -dims {DIMS} -eta0 0.01 -grad_norm {norm}",
          data, PREFIX)


def test_feature_again_later_in_a_row():
    """Feature PREFIX[2] also turns up in the tail of one row: every occurrence of a hot feature, wherever
    it sits, is reduced in the forward kernel, so that they all meet in ONE optimizer step."""
    rows_f, rows_v, rng = prefix_rows(4096, seed=7)      # two whole mini-batches: one ingest block
    rows_f[3321] = np.concatenate([rows_f[3321], [PREFIX[2]]])
    rows_v[3321] = np.concatenate([rows_v[3321], [0.75]])
    data = to_csr(rows_f, rows_v, rng)
    check("train_classifier", f"-loss logloss -opt adagrad -mini_batch 2048 -dims {DIMS} -eta0 0.05", data, PREFIX)


def test_row_that_breaks_the_pattern_and_short_rows():
    rows_f, rows_v, rng = prefix_rows(4096, seed=9)
    rows_f[17][1] = 555                      # position 1 holds another feature in one row
    rows_f[2500] = rows_f[2500][:5]          # a row with only positions 0..4
    rows_v[2500] = rows_v[2500][:5]
    data = to_csr(rows_f, rows_v, rng)
    # (a feature is hot by its frequency, not by sitting at the same place in every row)
    check("train_classifier", f"-loss logloss -opt ftrl -mini_batch 2048 -dims {DIMS} -alpha 0.05", data, PREFIX)


def test_same_feature_twice_in_a_row():
    """Two lanes of a warp hit the same accumulator: the adds are serialised in lane order."""
    prefix = [5, 9, 5, 11]
    rows_f, rows_v, rng = prefix_rows(4096, prefix=prefix, seed=11)
    data = to_csr(rows_f, rows_v, rng)
    check("train_classifier", f"-loss logloss -opt adam -mini_batch 2048 -dims {DIMS} -eta0 0.01", data, [5, 9, 11])


def test_empty_rows_and_long_rows():
    """Empty rows, and rows of more than 64 elements whose tail holds hot features again (the forward kernel
    keeps a lane's first two elements in registers and re-reads longer tails)."""
    rows_f, rows_v, rng = prefix_rows(4096, seed=13)
    rows_f[100] = rows_f[100][:0]
    rows_v[100] = rows_v[100][:0]
    for r in (7, 1000, 3000, 4095):
        extra = rng.integers(0, DIMS, 90)
        extra[[40, 70, 88]] = [PREFIX[0], PREFIX[3], PREFIX[0]]
        rows_f[r] = np.concatenate([rows_f[r], extra])
        rows_v[r] = np.concatenate([rows_v[r], rng.normal(0, 1, 90)])
    data = to_csr(rows_f, rows_v, rng)
    check("train_classifier", f"-loss hinge -opt sgd -mini_batch 2048 -dims {DIMS} -eta0 0.05", data, PREFIX)


def test_blocks_with_different_columns():
    """Two pushes whose rows start differently: hot features are picked per ingest block."""
    a_f, a_v, rng = prefix_rows(4096, seed=15)
    b_f, b_v, _ = prefix_rows(4096, prefix=[3, 8], seed=16)
    data = to_csr(a_f + b_f, a_v + b_v, rng)
    check("train_classifier", f"-loss logloss -opt adagrad -mini_batch 2048 -dims {DIMS} -eta0 0.05", data, [3, 8],
          pushes=[4096, 4096])


def test_switched_off_by_environment():
    rows_f, rows_v, rng = prefix_rows(5000, seed=17)
    data = to_csr(rows_f, rows_v, rng)
    os.environ["HM_HOT_FEATURES"] = "0"
    try:
        check("train_classifier", f"-loss logloss -opt adagrad -mini_batch 2048 -dims {DIMS} -eta0 0.05", data, [])
    finally:
        del os.environ["HM_HOT_FEATURES"]


def test_criteo_shaped_rows():
    """The 13 quantitative columns and the head of the small categorical fields are hot."""
    data = synth.criteo_like(20000, dims=1 << 20, seed=1)
    options = f"-loss logloss -opt ftrl -mini_batch 4096 -dims {1 << 20} -alpha 0.05"
    u, w, state = run_gpu_linear("train_classifier", options, data)
    hot = set(int(x) for x in u.hot_features())
    assert set(int(x) for x in data.feature[:13]) <= hot and 20 <= len(hot) <= 64
    ref, ref_state = run_oracle_linear("train_classifier", options, data)
    assert_weights_close(w, ref.w)
    assert_weights_close(state, ref_state, "state")
    # re-runs are bit-identical (fixed summation order, no float atomics)
    u2, w2, state2 = run_gpu_linear("train_classifier", options, data)
    assert np.array_equal(w, w2) and np.array_equal(state, state2)
    u.destroy()
    u2.destroy()


def test_out_of_range_feature_is_reported_and_ignored():
    """An index >= dims is dropped from the sort (it must not split a valid feature's run) and reported
    at close(); the rest of the model is what the oracle gets for the rows without that element."""
    rows_f, rows_v, rng = prefix_rows(5000, seed=19)
    clean = to_csr([f.copy() for f in rows_f], [v.copy() for v in rows_v], np.random.default_rng(1))
    rows_f[1234] = np.concatenate([rows_f[1234], [DIMS + 5]])
    rows_v[1234] = np.concatenate([rows_v[1234], [1.0]])
    bad = to_csr(rows_f, rows_v, np.random.default_rng(1))
    bad.label = clean.label
    options = f"-loss logloss -opt adagrad -mini_batch 2048 -dims {DIMS} -eta0 0.05"
    rowptr, feature, value, label = bad.numpy()
    u = hm.UDTF(options, 0, function_name="train_classifier")
    u.process_batch(rowptr, feature, value, label)
    with pytest.raises(hm.UDFArgumentError):
        u.close()
    w = u.array_to_host("w")
    ref, _ = run_oracle_linear("train_classifier", options, clean)
    assert_weights_close(w, ref.w)
    u.destroy()


@pytest.mark.parametrize("hot", ["1", "0"])
@pytest.mark.parametrize("lpr", ["16", "8"])
def test_rows_side_by_side_in_a_warp(monkeypatch, lpr, hot):
    """Short rows are worked on by half (or a quarter) of a warp each, two (four) rows side by side
    (hm_linear.cu: forward_rows_lpr / forward_rows_hot<LD, 16>; the library picks this for mini-batches that
    fill the GPU, HM_FWD_LPR forces it here).  Same data as the hard cases above: a hot feature twice in a row
    (the row group's serial fallback), empty rows, rows longer than the 48 elements a group keeps in registers
    with hot features in the re-read tail, a ragged last mini-batch."""
    monkeypatch.setenv("HM_FWD_LPR", lpr)
    monkeypatch.setenv("HM_HOT_FEATURES", hot)
    prefix = [5, 9, 5, 11, 40000]
    rows_f, rows_v, rng = prefix_rows(4096 + 333, prefix=prefix, seed=23)
    rows_f[100] = rows_f[100][:0]
    rows_v[100] = rows_v[100][:0]
    rows_f[101] = rows_f[101][:2]
    rows_v[101] = rows_v[101][:2]
    for r in (6, 7, 1000, 3001, 4096, 4428):
        extra = rng.integers(0, DIMS, 90)
        extra[[10, 41, 70, 88]] = [9, 5, 11, 5]
        rows_f[r] = np.concatenate([rows_f[r], extra])
        rows_v[r] = np.concatenate([rows_v[r], rng.normal(0, 1, 90)])
    data = to_csr(rows_f, rows_v, rng)
    want = [5, 9, 11, 40000] if hot == "1" else []
    check("train_classifier", f"-loss logloss -opt adagrad -mini_batch 2048 -dims {DIMS} -eta0 0.05", data, want)
    check("train_classifier", f"-loss logloss -opt ftrl -mini_batch 2048 -dims {DIMS} -alpha 0.05", data, want)
