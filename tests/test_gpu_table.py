"""The model table as a file (hm_udtf_save_model_table) and warm start from it
(hm_udtf_load_model_table): text (Hive text-table layout, delimiter a parameter) and binary.
The reference's own table layout is unknown (no source is mounted)."""
import numpy as np
import pytest

import hivemall_b200 as hm
from hivemall_b200 import synth

from _util import assert_weights_close

pytestmark = pytest.mark.gpu


@pytest.mark.parametrize("binary,delim", [(False, "\t"), (False, "\x01"), (True, "\t")])
def test_sgd_continues_exactly_from_a_saved_table(tmp_path, binary, delim):
    """train -> save -> load into a fresh learner -> continue == the uninterrupted run, bit for bit (-opt sgd
    has no optimizer slots, so the table holds the whole model)."""
    dims, mb = 1 << 14, 512
    data = synth.criteo_like(8 * mb, dims=dims, seed=7)
    rp, f, v, y = data.numpy()
    opts = f"-loss logloss -opt sgd -mini_batch {mb} -dims {dims} -eta0 0.05 -eta inverse -power_t 0.3"
    cut = 5 * mb
    whole = hm.train_classifier(opts)
    whole.process_batch(rp, f, v, y)
    whole.close()
    first = hm.train_classifier(opts)
    first.process_batch(rp[: cut + 1], f, v, y[:cut])
    path = str(tmp_path / ("model.bin" if binary else "model.tsv"))
    n = first.save_model_table(path, binary=binary, delimiter=delim)
    feat, wt = first.forward()
    assert n == len(feat) > 100
    if not binary:
        lines = open(path).read().splitlines()
        assert len(lines) == n and lines[0].split(delim)[0] == str(feat[0])
        assert np.float32(lines[0].split(delim)[1]) == wt[0]
    second = hm.train_classifier(opts)
    assert second.load_model_table(path, binary=binary, delimiter=delim) == n
    assert np.array_equal(second.array_to_host("w"), first.array_to_host("w"))
    second.step = first.step
    second.process_batch(rp[cut:], f, v, y[cut:])
    second.close()
    assert np.array_equal(second.array_to_host("w"), whole.array_to_host("w"))
    for u in (whole, first, second):
        u.destroy()


@pytest.mark.parametrize("binary", [False, True])
def test_train_fm_table_round_trip(tmp_path, binary):
    dims, k = 1 << 12, 8
    data = synth.criteo_like(3000, dims=dims, seed=9)
    rp, f, v, y = data.numpy()
    opts = f"-factors {k} -classification -opt adagrad -mini_batch 256 -dims {dims} -eta0 0.05"
    a = hm.train_fm(opts)
    a.process_batch(rp, f, v, y)
    path = str(tmp_path / "fm.table")
    n = a.save_model_table(path, binary=binary)
    feat, wt, vif = a.forward()
    assert n == len(feat) and feat[0] == -1 and vif.shape == (n, k)
    b = hm.train_fm(opts)
    assert b.load_model_table(path, binary=binary) == n
    feat2, wt2, vif2 = b.forward()                 # the loaded features count as seen: the same table comes back
    assert np.array_equal(feat2, feat) and np.array_equal(wt2, wt) and np.array_equal(vif2, vif)
    held = synth.criteo_like(500, dims=dims, seed=9, row0=1 << 20)
    hr, hf, hv, _ = held.numpy()
    assert np.array_equal(a.predict(hr, hf, hv), b.predict(hr, hf, hv))
    a.destroy()
    b.destroy()


def test_ftrl_weights_are_rebuilt_and_stateful_optimizers_are_lossy(tmp_path):
    dims = 1 << 14
    data = synth.criteo_like(4096, dims=dims, seed=11)
    rp, f, v, y = data.numpy()
    opts = f"-loss logloss -opt ftrl -mini_batch 512 -dims {dims} -alpha 0.1 -lambda1 0.001 -lambda2 0.01"
    a = hm.train_classifier(opts)
    a.process_batch(rp, f, v, y)
    path = str(tmp_path / "ftrl.tsv")
    a.save_model_table(path)
    b = hm.train_classifier(opts)
    b.load_model_table(path)
    # FTRL stores (z, n): z is rebuilt from the weight with n = 0, the weights agree to rounding
    assert_weights_close(b.array_to_host("w"), a.array_to_host("w"), rtol=1e-6, atol=1e-9)
    assert np.count_nonzero(b.array_to_host("state").reshape(-1, 2)[:, 1]) == 0      # n is not in a model table
    a.destroy()
    b.destroy()


def test_malformed_files_are_argument_errors(tmp_path):
    u = hm.train_classifier("-opt sgd -dims 64 -mini_batch 4")
    bad = tmp_path / "bad.tsv"
    bad.write_text("3\t0.5\nseven\t1.0\n")
    with pytest.raises(hm.UDFArgumentError) as e:
        u.load_model_table(str(bad))
    assert "line 2" in str(e.value)
    bad.write_text("64\t0.5\n")
    with pytest.raises(hm.UDFArgumentError):
        u.load_model_table(str(bad))                         # feature >= dims
    with pytest.raises(hm.UDFArgumentError):
        u.load_model_table(str(tmp_path / "missing.tsv"))
    with pytest.raises(hm.UDFArgumentError):
        u.load_model_table(str(bad), binary=True)            # not a binary table
    good = tmp_path / "good.tsv"
    good.write_text("3\t0.5\n\n10\t-1.25e-3\n")
    assert u.load_model_table(str(good)) == 2
    w = u.array_to_host("w")
    assert w[3] == np.float32(0.5) and w[10] == np.float32(-1.25e-3) and np.count_nonzero(w) == 2
    u.destroy()


def test_online_learners_tables(tmp_path):
    """The passive-aggressive learners have no optimizer slots: train -> save -> load -> continue is exact.  The
    variance-keeping learners save their weights; a fresh learner that loads them starts from variance 1 again,
    and the variances of a run can be carried over as array `state`, after which continuing is exact too."""
    dims, mb = 1 << 14, 512
    data = synth.criteo_like(6 * mb, dims=dims, seed=9)
    rp, f, v, y = data.numpy()
    cut = 4 * mb
    for fn, opts in (("train_pa1", f"-mini_batch {mb} -dims {dims} -c 0.1"), ("train_arow", f"-mini_batch {mb} -dims {dims} -r 1")):
        whole = hm.UDTF(opts, 0, function_name=fn)
        whole.process_batch(rp, f, v, y)
        whole.close()
        first = hm.UDTF(opts, 0, function_name=fn)
        first.process_batch(rp[: cut + 1], f, v, y[:cut])
        path = str(tmp_path / f"{fn}.tsv")
        n = first.save_model_table(path)
        second = hm.UDTF(opts, 0, function_name=fn)
        assert second.load_model_table(path) == n > 100
        assert np.array_equal(second.array_to_host("w"), first.array_to_host("w"))
        if fn == "train_arow":
            assert (second.array_to_host("state") == 1.0).all()
            second.array_from_host("state", first.array_to_host("state"))
        second.step = first.step
        second.process_batch(rp[cut:] - rp[cut], f[rp[cut]:], v[rp[cut]:], y[cut:])
        second.close()
        assert np.array_equal(second.array_to_host("w"), whole.array_to_host("w")), fn
        if fn == "train_arow":
            assert np.array_equal(second.array_to_host("state"), whole.array_to_host("state"))
        for u in (whole, first, second):
            u.destroy()
