"""Parity of the CUDA linear learner with the CPU oracle, through the C-ABI.

PARITY UNPINNED w.r.t. Hivemall itself (no reference source is mounted): what is checked
is the product against oracle/hm_oracle.c, two independent statements of DESIGN.md
section 3, on identical rows, order, seed and mini-batch size.
"""
import numpy as np
import pytest

import hivemall_b200 as hm
from hivemall_b200 import synth

import _oracle
from _util import assert_weights_close, run_gpu_linear, run_oracle_linear

pytestmark = pytest.mark.gpu

OPTS = ["sgd", "adagrad", "adadelta", "adam", "ftrl"]


@pytest.mark.parametrize("opt", OPTS)
def test_a9a_shaped_all_optimizers(opt):
    """configs[0] shape: 123 binary features, -mini_batch 128, every optimizer."""
    data = synth.a9a_like(4096 + 37, seed=0)      # ragged tail: last mini-batch is short
    options = f"-loss logloss -opt {opt} -mini_batch 128 -dims 124 -eta0 0.1"
    u, w, state = run_gpu_linear("train_classifier", options, data)
    ref, ref_state = run_oracle_linear("train_classifier", options, data)
    assert u.step == ref.t == (4096 + 37 + 127) // 128
    assert_weights_close(w, ref.w, f"w[{opt}]")
    assert_weights_close(state, ref_state, f"state[{opt}]")
    u.destroy()


def test_config0_full_model_table():
    """configs[0] in full: 32K rows, logloss + AdaGrad, -mini_batch 128; the model table
    (feature order and weights) and held-out AUC / logloss agree with the oracle."""
    data = synth.a9a_like(32768, seed=0)
    options = "-loss logloss -opt AdaGrad -mini_batch 128 -dims 124"
    u, w, _ = run_gpu_linear("train_classifier", options, data, pushes=[10000, 10000, 12768])
    ref, _ = run_oracle_linear("train_classifier", options, data)
    assert_weights_close(w, ref.w)
    feat, wt = u.forward()
    ofeat, owt = _oracle.model_table(ref.w)
    assert np.array_equal(feat, ofeat)            # bit-exact feature ordering of the model table
    assert_weights_close(wt, owt, "model table weights")
    held = synth.a9a_like(8192, seed=0, row0=1 << 20)
    rp, f, v, y = held.numpy()
    import torch
    m_gpu = torch.from_numpy(u.predict(rp, f, v))
    m_ref = torch.from_numpy(_oracle.predict_linear(rp, f, v, ref.w))
    yt = torch.from_numpy(y)
    assert abs(synth.auc(m_gpu, yt) - synth.auc(m_ref, yt)) < 1e-4
    assert abs(synth.logloss(m_gpu, yt) - synth.logloss(m_ref, yt)) < 1e-4
    assert synth.auc(m_gpu, yt) > 0.7             # it learned something
    u.destroy()


@pytest.mark.parametrize("loss,fn", [("logloss", "train_classifier"), ("hinge", "train_classifier"),
                                     ("squared", "train_regressor")])
def test_losses_on_ragged_rows(loss, fn):
    """Uneven rows incl. empty ones, duplicate features inside a row, signed values."""
    data = synth.ragged(3000, dims=5000, max_nnz=60, seed=3, regression=(fn == "train_regressor"))
    options = f"-loss {loss} -opt adagrad -mini_batch 256 -dims 5000 -eta0 0.05"
    u, w, state = run_gpu_linear(fn, options, data)
    ref, ref_state = run_oracle_linear(fn, options, data)
    assert_weights_close(w, ref.w)
    assert_weights_close(state, ref_state, "state")
    u.destroy()


@pytest.mark.parametrize("opt", ["ftrl", "adam"])
def test_criteo_shaped_hashed_dims(opt):
    """configs[1]/[3] shape at oracle-sized scale: 39 feats/row, 2^20 hashed dims, mini-batches
    large enough that hot features span many CTAs of the update kernel."""
    data = synth.criteo_like(40000, dims=1 << 20, seed=1, regression=(opt == "adam"))
    fn = "train_regressor" if opt == "adam" else "train_classifier"
    loss = "squared" if opt == "adam" else "logloss"
    options = f"-loss {loss} -opt {opt} -mini_batch 8192 -dims {1 << 20} -eta0 0.01"
    u, w, state = run_gpu_linear(fn, options, data)
    ref, ref_state = run_oracle_linear(fn, options, data)
    assert_weights_close(w, ref.w)
    assert_weights_close(state, ref_state, "state")
    u.destroy()


@pytest.mark.parametrize("dims", [300, 511, (1 << 17) - 3, 1 << 18, (1 << 24) + 1, 1 << 25, (1 << 27) - 1])
def test_key_widths_that_use_nine_bit_digits(dims):
    """The sort takes 9-bit digits where that saves a pass (9, 17-18, 25-27 key bits): 1, 2 or 3 passes
    of mixed widths must group features exactly like 8-bit digits do."""
    data = synth.criteo_like(12000, dims=dims, seed=5)
    options = f"-opt adagrad -mini_batch 4096 -dims {dims} -eta0 0.05"
    u, w, state = run_gpu_linear("train_classifier", options, data)
    ref, ref_state = run_oracle_linear("train_classifier", options, data)
    assert_weights_close(w, ref.w)
    assert_weights_close(state, ref_state, "state")
    u.destroy()


def test_push_boundaries_do_not_matter():
    """Mini-batch boundaries belong to the row stream, not to the process() calls."""
    data = synth.criteo_like(5000, dims=1 << 16, seed=2)
    options = f"-opt adagrad -mini_batch 300 -dims {1 << 16}"
    _, w1, s1 = run_gpu_linear("train_classifier", options, data, pushes=[5000])
    _, w2, s2 = run_gpu_linear("train_classifier", options, data, pushes=[1, 299, 17, 1000, 2, 3681])
    assert np.array_equal(w1, w2) and np.array_equal(s1, s2)   # same launches on the same data: bit-identical


def test_iters_and_schedules():
    data = synth.ragged(2000, dims=800, max_nnz=20, seed=5)
    for extra in ["-iters 3 -eta inverse -power_t 0.2", "-iters 2 -eta simple -total_steps 10 -reg l2 -lambda 0.01",
                  "-iters 2 -reg l1 -lambda 0.001 -opt sgd", "-grad_norm count", "-grad_norm sum -eta0 0.001"]:
        options = f"-opt adagrad -mini_batch 128 -dims 800 {extra}"
        u, w, state = run_gpu_linear("train_classifier", options, data, pushes=[700, 1300])
        ref, ref_state = run_oracle_linear("train_classifier", options, data)
        assert u.step == ref.t
        assert_weights_close(w, ref.w, f"w[{extra}]")
        assert_weights_close(state, ref_state, f"state[{extra}]")
        u.destroy()


def test_replays_do_not_sort_again():
    """-iters 3: the later passes train from the sort results the first pass kept: sort kernels are
    launched in pass 1 only, results match the oracle, and host pushes in several pieces (a carried tail,
    a short last mini-batch) replay in the same order."""
    data = synth.criteo_like(20000 + 333, dims=1 << 16, seed=4)
    rowptr, feature, value, label = data.numpy()
    base = f"-opt adagrad -mini_batch 4096 -dims {1 << 16} -eta0 0.05"
    counts = {}
    for iters in (1, 3):
        u = hm.UDTF(f"{base} -iters {iters}", 0, function_name="train_classifier")
        u.set_profiling(True)
        r0 = 0
        for cnt in (9000, 11333):
            u.process_batch(rowptr[r0:r0 + cnt + 1], feature, value, label[r0:r0 + cnt])
            r0 += cnt
        u.close()
        kt = u.kernel_times()
        counts[iters] = (kt["sort_scatter"]["launches"], kt["sort_hist"]["launches"], kt["linear_update"]["launches"])
        if iters == 3:
            w, state = u.array_to_host("w"), u.array_to_host("state")
        u.destroy()
    assert counts[3][0] == counts[1][0] > 0 and counts[3][1] == counts[1][1]
    assert counts[3][2] == 3 * counts[1][2]
    ref, ref_state = run_oracle_linear("train_classifier", f"{base} -iters 3", data)
    assert_weights_close(w, ref.w)
    assert_weights_close(state, ref_state, "state")


def test_deterministic_rerun():
    """No float atomics anywhere: two runs are bit-identical."""
    data = synth.criteo_like(30000, dims=1 << 18, seed=7)
    options = f"-opt ftrl -mini_batch 4096 -dims {1 << 18}"
    _, w1, s1 = run_gpu_linear("train_classifier", options, data)
    _, w2, s2 = run_gpu_linear("train_classifier", options, data)
    assert np.array_equal(w1, w2) and np.array_equal(s1, s2)


def test_edge_cases():
    # empty input: close() on a learner that saw nothing gives an empty model table
    u = hm.train_classifier("-dims 100 -mini_batch 8")
    assert u.close() == 0
    f, w = u.forward()
    assert len(f) == 0 and len(w) == 0
    u.destroy()
    # rows with no features at all
    u = hm.train_classifier("-dims 100 -mini_batch 4")
    u.process_batch(np.zeros(11, np.int64), np.zeros(0, np.int32), np.zeros(0, np.float32), np.ones(10, np.float32))
    assert u.close() == 0 and u.step == 3
    u.destroy()
    # a single row, mini-batch larger than the data
    u = hm.train_classifier("-dims 10 -mini_batch 1000 -opt sgd -eta0 1.0")
    u.process([(3, 2.0), (7, -1.0)], 1)
    assert u.close() == 2
    f, w = u.forward()
    assert list(f) == [3, 7]
    np.testing.assert_allclose(w, [1.0, -0.5], rtol=1e-6)   # -eta * dloss * x, dloss = -1/(1+e^0) = -0.5
    u.destroy()
    # out-of-range feature index is an error at close(), not a crash
    u = hm.train_classifier("-dims 10 -mini_batch 2")
    u.process([(10, 1.0)], 1)
    with pytest.raises(hm.UDFArgumentError):
        u.close()
    u.destroy()
    # lifecycle errors
    u = hm.train_classifier("-dims 10")
    with pytest.raises(hm.HivemallError):   # forward() before close()
        hm.udtf.check(u._lib.hm_udtf_forward(u._h, None, None, None, 0, None))
    u.close()
    with pytest.raises(hm.HivemallError):
        u.process_batch(np.zeros(2, np.int64), np.zeros(0, np.int32), np.zeros(0, np.float32), np.ones(1, np.float32))
    u.destroy()


def test_process_device_matches_host_path():
    import torch
    data = synth.criteo_like(16384, dims=1 << 18, seed=11, device="cuda")
    options = f"-opt ftrl -mini_batch 4096 -dims {1 << 18}"
    u = hm.train_classifier(options)
    u.process_device(data.rows, data.nnz, data.rowptr, data.feature, data.value, data.label)
    u.close()
    w_dev = u.array_to_host("w")
    host = synth.Csr(data.rowptr.cpu(), data.feature.cpu(), data.value.cpu(), data.label.cpu())
    _, w_host, _ = run_gpu_linear("train_classifier", options, host)
    assert np.array_equal(w_dev, w_host)
    ref, _ = run_oracle_linear("train_classifier", options, host)
    assert_weights_close(w_dev, ref.w)
    torch.cuda.synchronize()
    u.destroy()


def _csr_from_lengths(lens, dims, seed):
    import torch
    lens = np.asarray(lens, dtype=np.int64)
    rowptr = np.zeros(len(lens) + 1, np.int64)
    rowptr[1:] = np.cumsum(lens)
    rng = np.random.default_rng(seed)
    nnz = int(rowptr[-1])
    feat = (rng.random(nnz) ** 2 * dims).astype(np.int32)
    val = ((rng.random(nnz) - 0.5) * 2).astype(np.float32)
    label = (rng.random(len(lens)) < 0.5).astype(np.float32)
    return synth.Csr(torch.from_numpy(rowptr), torch.from_numpy(feat), torch.from_numpy(val), torch.from_numpy(label))


def test_rows_longer_than_a_sort_tile_and_runs_of_empty_rows():
    """Rows of 5,000-20,000 elements (longer than the 4096-element sort tile), next to thousands of
    empty and 1-2 element rows (more than 1024 rows inside one tile: the row lookup's slow path)."""
    rng = np.random.default_rng(7)
    lens = []
    for _ in range(6):
        lens += [int(rng.integers(5000, 20000))]
        lens += list(rng.integers(0, 3, size=3000))
        lens += [0] * 2500
    data = _csr_from_lengths(lens, dims=50000, seed=8)
    options = "-opt adagrad -mini_batch 4096 -dims 50000"
    u, w, state = run_gpu_linear("train_classifier", options, data, pushes=[7000, 9000, 20000])
    ref, ref_state = run_oracle_linear("train_classifier", options, data)
    assert u.step == ref.t
    assert_weights_close(w, ref.w)
    assert_weights_close(state, ref_state, "state")
    u.destroy()


def test_mini_batch_of_one_row():
    data = synth.ragged(300, dims=200, max_nnz=12, seed=12)
    options = "-opt sgd -mini_batch 1 -dims 200 -eta0 0.05"
    u, w, _ = run_gpu_linear("train_classifier", options, data)
    ref, _ = run_oracle_linear("train_classifier", options, data)
    assert u.step == ref.t == 300
    assert_weights_close(w, ref.w)
    u.destroy()


def test_one_mini_batch_larger_than_the_block_cap():
    """A single mini-batch of 2^21 rows x 39 = 82 M elements exceeds the 2^26-element block cap:
    it becomes a block of its own and every update CTA walks many tiles."""
    import torch
    n = (1 << 21) + 12345
    parts = [synth.criteo_like(min(1 << 20, n - r), dims=1 << 22, seed=9, device="cuda", row0=r) for r in range(0, n, 1 << 20)]
    data = synth.Csr(torch.arange(n + 1, device="cuda", dtype=torch.int64) * 39, torch.cat([p.feature for p in parts]),
                     torch.cat([p.value for p in parts]), torch.cat([p.label for p in parts]))
    options = f"-opt adagrad -mini_batch {1 << 21} -dims {1 << 22}"
    u = hm.train_classifier(options)
    u.process_device(n, data.nnz, data.rowptr, data.feature, data.value, data.label)
    u.close()
    assert u.step == 2
    host = synth.Csr(data.rowptr.cpu(), data.feature.cpu(), data.value.cpu(), data.label.cpu())
    ref = _oracle.train_linear_mt(_oracle.config_from(hm.parse_options("train_classifier", options)), *host.numpy(), 8)
    assert_weights_close(u.array_to_host("w"), ref.w)
    assert_weights_close(u.array_to_host("state"), ref.s1, "state")
    u.destroy()


@pytest.mark.parametrize("fn,loss", [("train_classifier", "squared_hinge"), ("train_classifier", "modified_huber"),
                                     ("train_regressor", "huber -huber_c 0.5"), ("train_regressor", "quantile -quantile 0.8"),
                                     ("train_regressor", "epsilon_insensitive -epsilon 0.1"),
                                     ("train_regressor", "squared_epsilon_insensitive -epsilon 0.1")])
def test_further_losses_on_the_hot_path(fn, loss):
    """The further plain losses through the launch shape of the headline in small: Criteo-shaped rows, 4096-row
    mini-batches, hot features reduced in the forward kernel, several pushes."""
    import torch
    dims = 1 << 18
    data = synth.criteo_like(3 * 4096 + 99, dims=dims, seed=21)
    if fn == "train_regressor":
        rng = np.random.default_rng(4)
        data.label = data.label * 2.0 + torch.from_numpy(rng.normal(0, 0.5, data.rows).astype(np.float32))
    options = f"-loss {loss} -opt adagrad -mini_batch 4096 -dims {dims} -eta0 0.05"
    u, w, state = run_gpu_linear(fn, options, data, pushes=[5000, 8000])
    assert len(u.hot_features()) >= 13
    ref, ref_state = run_oracle_linear(fn, options, data)
    assert_weights_close(w, ref.w)
    assert_weights_close(state, ref_state, "state")
    assert np.count_nonzero(ref.w) > 1000
    u.destroy()
