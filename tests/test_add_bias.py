"""add_bias ("bias / add_bias-style preprocessing"): the host mirrors, and the
bias clause "0:1.0" travelling through the device-side string ingest into training.

PARITY UNPINNED w.r.t. Hivemall (no reference source is mounted): that the bias clause is the string "0:1.0"
is this library's statement, a constant in hivemall_b200/udtf.py.
"""
import numpy as np
import pytest

import hivemall_b200 as hm
from hivemall_b200 import synth


def test_add_bias_strings():
    assert hm.add_bias(["a:1", "b"]) == ["a:1", "b", "0:1.0"]
    assert hm.add_bias([]) == ["0:1.0"]
    assert hm.add_bias([b"a:1"]) == [b"a:1", b"0:1.0"]


def test_add_bias_csr_ragged_rows():
    rowptr = np.array([0, 2, 2, 5], dtype=np.int64)
    feature = np.array([7, 9, 1, 2, 3], dtype=np.int32)
    value = np.array([0.5, 1.5, 2.0, 3.0, 4.0], dtype=np.float32)
    rp, f, v = hm.add_bias_csr(rowptr, feature, value)
    assert rp.tolist() == [0, 3, 4, 8]
    assert f.tolist() == [7, 9, 0, 0, 1, 2, 3, 0]
    assert v.tolist() == [0.5, 1.5, 1.0, 1.0, 2.0, 3.0, 4.0, 1.0]
    rp, f, v = hm.add_bias_csr(rowptr, feature, None)
    assert v is None and f.tolist() == [7, 9, 0, 0, 1, 2, 3, 0]
    rp, f, v = hm.add_bias_csr(rowptr, feature, None, bias_feature=100, bias_value=0.5)
    assert f.tolist() == [7, 9, 100, 100, 1, 2, 3, 100] and v.tolist() == [1, 1, 0.5, 0.5, 1, 1, 1, 0.5]
    # a slice of a larger CSR (rowptr not starting at 0)
    rp, f, v = hm.add_bias_csr(rowptr[1:], feature, value)
    assert rp.tolist() == [0, 1, 5] and f.tolist() == [0, 1, 2, 3, 0] and v.tolist() == [1.0, 2.0, 3.0, 4.0, 1.0]


@pytest.mark.gpu
def test_bias_clause_through_string_ingest_matches_the_oracle():
    """Rows pushed as strings with add_bias applied: the device parses "0:1.0" as feature 0 with value 1, it is
    hot (every row has it), and the model -- bias weight included -- is what the oracle computes for the CSR
    rows with the element appended."""
    import _oracle
    from _util import assert_weights_close
    dims = 1 << 12
    rng = np.random.default_rng(3)
    n = 6000
    rows, labels = [], []
    for r in range(n):
        k = int(rng.integers(0, 6))
        feats = [f"{int(j)}:{float(np.round(x, 3))}" for j, x in zip(rng.integers(1, dims, k), rng.normal(0, 1, k))]
        rows.append(hm.add_bias(feats))
        labels.append(float(rng.random() < 0.7))       # unbalanced: the bias has something to learn
    options = f"-loss logloss -opt adagrad -mini_batch 2048 -dims {dims} -eta0 0.1"
    u = hm.train_classifier(options)
    for row, y in zip(rows, labels):
        u.process(row, y)
    u.close()
    w = u.array_to_host("w")
    assert 0 in set(int(x) for x in u.hot_features())
    # the same rows as CSR for the oracle
    idx, val = hm.feature_hashing([s for row in rows for s in row], num_features=dims)
    rowptr = np.zeros(n + 1, dtype=np.int64)
    rowptr[1:] = np.cumsum([len(r) for r in rows])
    cfg = _oracle.config_from(hm.parse_options("train_classifier", options))
    ref = _oracle.train_linear(cfg, rowptr, idx, val, np.asarray(labels, np.float32))
    assert_weights_close(w, ref.w)
    assert w[0] > 0.1            # the bias learned the positive prior
    # and add_bias_csr over the bias-free CSR gives the very same rows
    plain_rows = [row[:-1] for row in rows]
    idx0, val0 = hm.feature_hashing([s for row in plain_rows for s in row], num_features=dims)
    rp0 = np.zeros(n + 1, dtype=np.int64)
    rp0[1:] = np.cumsum([len(r) for r in plain_rows])
    rp1, f1, v1 = hm.add_bias_csr(rp0, idx0, val0)
    assert np.array_equal(rp1, rowptr) and np.array_equal(f1, idx) and np.array_equal(v1, val)
    u.destroy()
