"""Entry points around the training loop: warm start, replica scaling (the 1/P step of MIX),
device-side prediction and hashing, last-mini-batch margins, counters."""
import ctypes as C

import numpy as np
import pytest

import hivemall_b200 as hm
from hivemall_b200 import _lib, synth

import _oracle

pytestmark = pytest.mark.gpu


@pytest.mark.parametrize("opt", ["adagrad", "adam", "ftrl"])
def test_warm_start_round_trip(opt):
    """Save (arrays, step) after half of the rows, load them into a fresh learner, continue:
    bit-identical to the uninterrupted run."""
    data = synth.criteo_like(8192, dims=1 << 16, seed=21)
    rp, f, v, y = data.numpy()
    opts = f"-opt {opt} -mini_batch 1024 -dims {1 << 16} -eta inverse"
    full = hm.train_classifier(opts)
    full.process_batch(rp, f, v, y)
    full.close()
    a = hm.train_classifier(opts)
    a.process_batch(rp[:4097], f, v, y[:4096])
    a.sync()
    saved = {"state": a.array_to_host("state"), "step": a.step}
    if opt != "ftrl":
        saved["w"] = a.array_to_host("w")
    b = hm.train_classifier(opts)
    if opt != "ftrl":
        b.array_from_host("w", saved["w"])
    else:
        with pytest.raises(hm.HivemallError):        # FTRL weights are derived: only the state can be written
            b.array_from_host("w", np.zeros(1 << 16, np.float32))
    b.array_from_host("state", saved["state"])
    b.step = saved["step"]
    b.process_batch(rp[4096:], f, v, y[4096:])
    b.close()
    assert b.step == full.step == 8
    assert np.array_equal(b.array_to_host("w"), full.array_to_host("w"))
    assert np.array_equal(b.array_to_host("state"), full.array_to_host("state"))
    for u in (full, a, b):
        u.destroy()


def test_scale_replica_and_device_arrays():
    import torch
    data = synth.criteo_like(4096, dims=1 << 14, seed=22)
    u = hm.train_classifier(f"-opt adagrad -mini_batch 512 -dims {1 << 14}")
    u.process_batch(*data.numpy())
    u.sync()
    w0, s0 = u.array_to_host("w"), u.array_to_host("state")
    arrays = u.arrays()
    assert list(arrays) == ["replica"] and arrays["replica"][1:] == ((1 << 14) * 2, 2)
    t = torch.as_tensor(u.array_view("replica"), device="cuda")          # zero-copy view of the library's memory
    assert np.array_equal(t.cpu().numpy().reshape(-1, 2)[:, 0], w0)
    u.scale_replica(0.5)
    u.sync()
    assert np.array_equal(u.array_to_host("w"), w0 * np.float32(0.5))
    assert np.array_equal(u.array_to_host("state"), s0 * np.float32(0.5))
    u.destroy()


def test_predict_device_and_last_margins():
    import torch
    data = synth.criteo_like(3000, dims=1 << 14, seed=23, device="cuda")
    u = hm.train_classifier(f"-opt sgd -mini_batch 1000 -dims {1 << 14} -eta0 0.05")
    u.process_device(data.rows, data.nnz, data.rowptr, data.feature, data.value, data.label)
    u.sync()
    # margins of the LAST mini-batch as they were when it was scored = prediction with the model
    # of two mini-batches, which the oracle can reproduce
    host = synth.Csr(data.rowptr.cpu(), data.feature.cpu(), data.value.cpu(), data.label.cpu())
    rp, f, v, y = host.numpy()
    cfg = _oracle.config_from(hm.parse_options("train_classifier", f"-opt sgd -mini_batch 1000 -dims {1 << 14} -eta0 0.05"))
    two = _oracle.train_linear(cfg, rp[:2001], f, v, y[:2000])
    want = _oracle.predict_linear(rp[2000:] - rp[2000], f[rp[2000]:], v[rp[2000]:], two.w)
    got = u.last_margins()
    assert len(got) == 1000
    np.testing.assert_allclose(got, want, rtol=1e-5, atol=1e-6)
    # device-side prediction equals host-side prediction
    out = torch.empty(data.rows, dtype=torch.float32, device="cuda")
    u.predict_device(data.rows, data.rowptr, data.feature, data.value, out)
    u.sync()
    assert np.array_equal(out.cpu().numpy(), u.predict(rp, f, v))
    st = u.stats()
    assert (st.rows, st.nnz, st.mini_batches) == (3000, 3000 * 39, 3) and st.kernel_launches > 0
    u.destroy()


def test_mhash_device_matches_host_entry_point():
    import torch
    names = [f"field{i % 37}=v{i * 7919 % 100003}" for i in range(5000)]
    data, off = hm.udtf._pack_strings(names)
    d_data, d_off = torch.from_numpy(data).cuda(), torch.from_numpy(off).cuda()
    d_out = torch.empty(len(names), dtype=torch.int32, device="cuda")
    torch.cuda.synchronize()
    _lib.check(_lib.load().hm_mhash_device(C.c_void_p(d_data.data_ptr()), C.c_void_p(d_off.data_ptr()), len(names),
                                           hm.DEFAULT_HASH_SEED, 1 << 24, C.c_void_p(d_out.data_ptr()), C.c_void_p(0)))
    torch.cuda.synchronize()
    assert np.array_equal(d_out.cpu().numpy(), hm.mhash(names))
    assert np.array_equal(d_out.cpu().numpy(), _oracle.mhash(names, 1 << 24, hm.DEFAULT_HASH_SEED))


def test_profiling_counters_cover_every_launch():
    data = synth.criteo_like(4096, dims=1 << 14, seed=24)
    u = hm.train_classifier(f"-opt adagrad -mini_batch 2048 -dims {1 << 14}")
    u.set_profiling(True)
    u.process_batch(*data.numpy())
    u.sync()
    u.set_profiling(False)
    kt = {k: v for k, v in u.kernel_times().items() if v["launches"]}
    assert kt["linear_forward"]["launches"] == 2 and kt["linear_update"]["launches"] == 2
    assert kt["sort_scatter"]["launches"] == 2           # 14-bit keys: two digit passes, one block
    assert all(v["ms"] > 0 for v in kt.values())
    assert sum(v["launches"] for v in kt.values()) == u.stats().kernel_launches
    u.destroy()


def test_null_values_mean_one():
    """value == NULL on the host path: index-only features, every value 1.0 (nothing but the indices crosses the bus)."""
    data = synth.kdd12_like(5000, dims=1 << 16, seed=25)
    rp, f, v, y = data.numpy()
    assert np.all(v == 1.0)
    opts = f"-opt adagrad -mini_batch 512 -dims {1 << 16}"
    a = hm.train_classifier(opts)
    a.process_batch(rp, f, v, y)
    a.close()
    b = hm.train_classifier(opts)
    b.process_batch(rp[:1234], f, None, y[:1233])        # ragged pushes: carried tails keep their ones
    b.process_batch(rp[1233:], f, None, y[1233:])
    b.close()
    assert np.array_equal(a.array_to_host("w"), b.array_to_host("w"))
    assert b.stats().h2d_bytes < a.stats().h2d_bytes
    a.destroy(); b.destroy()


def _split(rp, f, v, rng):
    """Cut every row of a CSR into a quantitative group (a random-length prefix) and an index-only group (the
    rest, whose values are forced to 1.0); returns the split arrays and the merged CSR they stand for."""
    n = len(rp) - 1
    q_rp, c_rp = np.zeros(n + 1, np.int64), np.zeros(n + 1, np.int64)
    qf, qv, cf, mv = [], [], [], v.copy()
    for i in range(n):
        a, b = int(rp[i]), int(rp[i + 1])
        k = int(rng.integers(0, b - a + 1))           # 0 = no quantitative element, b - a = no index-only element
        qf.append(f[a:a + k]); qv.append(v[a:a + k]); cf.append(f[a + k:b])
        mv[a + k:b] = 1.0
        q_rp[i + 1] = q_rp[i] + k
        c_rp[i + 1] = c_rp[i] + (b - a - k)
    cat = lambda xs, dt: np.concatenate(xs).astype(dt) if xs else np.zeros(0, dt)
    return q_rp, cat(qf, np.int32), cat(qv, np.float32), c_rp, cat(cf, np.int32), mv


@pytest.mark.parametrize("fn,opts", [
    ("train_classifier", f"-opt ftrl -mini_batch 700 -dims {1 << 16}"),
    ("train_regressor", f"-loss squared -opt adam -mini_batch 64 -dims 5000 -iters 2"),
    ("train_fm", f"-factors 8 -classification -opt adagrad -mini_batch 512 -dims 5000"),
])
def test_split_rows_equal_merged_rows(fn, opts):
    """hm_udtf_process_split: quantitative + index-only groups merged on the device == the merged CSR pushed
    through hm_udtf_process, bit for bit, whatever the push boundaries (rows with an empty group included)."""
    dims = 5000 if "5000" in opts else 1 << 16
    rng = np.random.default_rng(7)
    data = synth.ragged(4000, dims, 30, seed=41, regression=(fn == "train_regressor"))
    rp, f, v, y = data.numpy()
    q_rp, qf, qv, c_rp, cf, mv = _split(rp, f, v, rng)
    a = hm.UDTF(opts, 0, function_name=fn)
    a.process_batch(rp, f, mv, y)
    a.close()
    b = hm.UDTF(opts, 0, function_name=fn)
    cuts = [0, 1, 1234, 1235, 3000, 4000]
    for lo, hi in zip(cuts, cuts[1:]):
        b.process_split(q_rp[lo:hi + 1], qf, qv, c_rp[lo:hi + 1], cf, y[lo:hi])
    b.close()
    assert a.step == b.step
    for name in (("w", "state") if fn != "train_fm" else ("w0", "w", "V", "gV")):
        assert np.array_equal(a.array_to_host(name), b.array_to_host(name)), name
    assert b.stats().h2d_bytes < a.stats().h2d_bytes
    a.destroy(); b.destroy()


def test_predict_reports_out_of_range_indices_itself():
    """Scoring rows with an index >= dims is the predict call's own error; it must not make a later
    close() complain about training data that was fine, and a clean predict afterwards succeeds."""
    data = synth.a9a_like(1024, seed=0)
    rp, f, v, y = data.numpy()
    u = hm.train_classifier("-opt adagrad -mini_batch 128 -dims 124")
    u.process_batch(rp, f, v, y)
    u.sync()
    bad_f = f.copy()
    bad_f[5] = 124
    with pytest.raises(hm.UDFArgumentError) as e:
        u.predict(rp[:9], bad_f, v)
    assert "scored row" in str(e.value)
    m = u.predict(rp[:9], f, v)
    assert np.isfinite(m).all()
    assert u.close() > 0
    u.destroy()


@pytest.mark.parametrize("fn,opts", [
    ("train_classifier", f"-opt ftrl -mini_batch 4096 -dims {1 << 18}"),
    ("train_classifier", f"-opt adagrad -mini_batch 300 -dims {1 << 18} -iters 2"),
    ("train_fm", f"-factors 8 -classification -opt adagrad -mini_batch 1000 -dims {1 << 18}"),
])
def test_column_rows_equal_expanded_rows(fn, opts):
    """hm_udtf_process_columns: a quantitative block whose indices are given once + index-only columns, expanded on
    the device == the same rows as plain CSR through hm_udtf_process, bit for bit; pushes of column rows and of
    CSR rows may alternate (the carried tail keeps its own row pointer), several H2D slices included."""
    n = 300_000 if "4096" in opts else 20_000
    data = synth.criteo_like(n, dims=1 << 18, seed=13)
    rp, f, v, y = data.numpy()
    f2, v2 = f.reshape(n, 39), v.reshape(n, 39)
    assert (f2[:, :13] == f2[0, :13]).all() and (v2[:, 13:] == 1.0).all()
    a = hm.UDTF(opts, 0, function_name=fn)
    a.process_batch(rp, f, v, y)
    a.close()
    b = hm.UDTF(opts, 0, function_name=fn)
    cuts = [0, 777, 778, n // 2 + 5, n]
    for i, (lo, hi) in enumerate(zip(cuts, cuts[1:])):
        if i == 1:      # one push as plain CSR in between
            b.process_batch(rp[lo:hi + 1], f, v, y[lo:hi])
        else:
            b.process_columns(f2[0, :13], v2[lo:hi, :13], f2[lo:hi, 13:], y[lo:hi])
    b.close()
    assert a.step == b.step
    for name in (("w", "state") if fn != "train_fm" else ("w0", "w", "V", "gV")):
        assert np.array_equal(a.array_to_host(name), b.array_to_host(name)), name
    assert b.stats().h2d_bytes < 0.6 * a.stats().h2d_bytes
    a.destroy(); b.destroy()
