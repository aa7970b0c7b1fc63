"""Bit-exact parity of GPU feature hashing with the oracle and the public MurmurHash3 vectors."""
import json
import os
import random

import numpy as np
import pytest

import hivemall_b200 as hm

import _oracle

pytestmark = pytest.mark.gpu
GOLDEN = os.path.join(os.path.dirname(__file__), "golden", "murmur3_kat.json")


def test_public_known_answers_on_gpu():
    kat = json.load(open(GOLDEN))
    for v in kat["vectors"]:
        data = bytes.fromhex(v["hex"])
        got = hm.mhash([data], num_features=0x7FFFFFFF, seed=v["seed"])
        assert int(got[0]) == _oracle.lib().hmo_fold(v["hash"], 0x7FFFFFFF)


def test_random_strings_bit_exact():
    rng = random.Random(5)
    strings = []
    for i in range(20000):
        n = rng.choice([0, 1, 2, 3, 4, 5, 7, 8, 9, 15, 16, 17, 31, 33, 64, 100])
        strings.append(bytes(rng.getrandbits(8) for _ in range(n)))
    strings += ["", "a", "feature", "ユーザー:属性", "x" * 5000, "age:39", "workclass#State-gov"]
    for nf, seed in [(1 << 24, 0x9747B28C), (16777213, 0), (1, 7), (123, 0xFFFFFFFF)]:
        got = hm.mhash(strings, num_features=nf, seed=seed)
        want = _oracle.mhash(strings, nf, seed)
        assert np.array_equal(got, want)
        assert got.min() >= 0 and got.max() < nf


def test_empty_batch():
    assert len(hm.mhash([])) == 0


def test_feature_hashing_parses_name_value():
    feats = ["age:39", "height:1.5e2", "country=jp", "price:-0.25", "7:3.0", "123456", "a:b:2", "weird:", ":5",
             "n:1e-3", "big:123456789.125", "zero:0", "ratio:.5", "url:http://x", "t:+4"]
    idx, val = hm.feature_hashing(feats, num_features=1 << 24)
    names = ["age", "height", "country=jp", "price", None, None, "a:b", "weird:", "", "n", "big", "zero", "ratio",
             "url:http://x", "t"]
    vals = [39, 150, 1, -0.25, 3, 1, 2, 1, 5, 1e-3, 123456789.125, 0, 0.5, 1, 4]
    for i, (nm, v) in enumerate(zip(names, vals)):
        assert val[i] == np.float32(v), feats[i]
        if nm is not None:
            assert idx[i] == _oracle.mhash([nm], 1 << 24, hm.DEFAULT_HASH_SEED)[0], feats[i]
    assert idx[4] == 7 and idx[5] == 123456


def test_process_strings_equals_process_indices():
    rows = [["age:3", "job=clerk", "edu:1.5"], ["job=clerk"], [], ["age:1", "age:2", "zip=94110"]] * 50
    labels = [1, 0, 1, 0] * 50
    opts = "-dims 4096 -mini_batch 16 -opt adagrad"
    a = hm.train_classifier(opts)
    for r, y in zip(rows, labels):
        a.process(r, y)
    a.close()
    b = hm.train_classifier(opts)
    for r, y in zip(rows, labels):
        i, v = hm.feature_hashing(r, num_features=4096) if r else (np.zeros(0, np.int32), np.zeros(0, np.float32))
        b.process(list(zip(i.tolist(), v.tolist())), y)
    b.close()
    assert np.array_equal(a.array_to_host("w"), b.array_to_host("w"))
    assert np.count_nonzero(a.array_to_host("w")) > 0


def test_process_strings_many_slices_bit_identical():
    """Rows pushed as feature strings are hashed on the device straight into the learner's CSR.  Several
    H2D slices, two pushes, a tail that carries over: the model is bit-identical to pushing the
    indices / values that the feature_hashing UDF returns for the same strings."""
    n_rows, per_row, dims = 300_000, 7, 1 << 18
    rng = np.random.default_rng(5)
    cat = (rng.random((n_rows, per_row)) ** 3 * 50_000).astype(np.int64)
    val = np.round(rng.random((n_rows, per_row)) * 4, 3)
    cols = []
    for j in range(per_row):
        if j < 2:
            cols.append(np.char.add(f"q{j}:", val[:, j].astype(str)))
        else:
            cols.append(np.char.add(f"c{j}=", cat[:, j].astype(str)))
    enc = np.char.encode(np.stack(cols, axis=1).reshape(-1), "ascii")
    lens = np.char.str_len(enc).astype(np.int64)
    offsets = np.zeros(len(enc) + 1, dtype=np.int64)
    offsets[1:] = np.cumsum(lens)
    data = np.frombuffer(b"".join(enc.tolist()), dtype=np.uint8)
    rowptr = np.arange(n_rows + 1, dtype=np.int64) * per_row
    label = (rng.random(n_rows) < 0.3).astype(np.float32)
    opts = f"-loss logloss -opt ftrl -mini_batch 4096 -dims {dims} -alpha 0.05"
    cut = 170_001
    a = hm.train_classifier(opts)
    a.process_strings(rowptr[: cut + 1], data, offsets, label[:cut])
    a.process_strings(rowptr[cut:], data, offsets, label[cut:])
    a.close()
    idx, v = hm.feature_hashing([bytes(e) for e in enc.tolist()], num_features=dims)
    assert np.array_equal(v.reshape(n_rows, per_row)[:, :2], val[:, :2].astype(np.float32))
    b = hm.train_classifier(opts)
    b.process_batch(rowptr, idx, v, label)
    b.close()
    assert np.array_equal(a.array_to_host("state"), b.array_to_host("state"))
    assert np.count_nonzero(a.array_to_host("w")) > 1000
    a.destroy()
    b.destroy()
