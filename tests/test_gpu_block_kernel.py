"""Mid-size mini-batches are trained by one persistent grid per block (linear_block_kernel: forward,
grid barrier, reduce + step, grid barrier, next mini-batch) instead of two launches per mini-batch.
Both paths obey the same arithmetic contract: each is compared with the oracle, with HM_BLOCK_KERNEL=0
selecting the two-kernel path for the same rows."""
import os

import numpy as np
import pytest

import hivemall_b200 as hm
from hivemall_b200 import synth

from _util import assert_weights_close, run_gpu_linear, run_oracle_linear

pytestmark = pytest.mark.gpu


def both_paths(fn, options, data, pushes=None):
    out = {}
    for flag in ("1", "0"):
        os.environ["HM_BLOCK_KERNEL"] = flag
        try:
            u, w, state = run_gpu_linear(fn, options, data, pushes=pushes)
        finally:
            del os.environ["HM_BLOCK_KERNEL"]
        out[flag] = (u.step, w, state, u.stats().kernel_launches)
        u.destroy()
    ref, ref_state = run_oracle_linear(fn, options, data)
    for flag, (step, w, state, _) in out.items():
        assert step == ref.t
        assert_weights_close(w, ref.w, f"w[block_kernel={flag}]")
        assert_weights_close(state, ref_state, f"state[block_kernel={flag}]")
    return out


@pytest.mark.parametrize("opt", ["sgd", "adagrad", "adadelta", "adam", "ftrl"])
def test_criteo_shaped_every_optimizer(opt):
    """Hot features, hot categorical heads that straddle CTAs, a short last mini-batch; far fewer launches."""
    data = synth.criteo_like(6 * 4096 + 1234, dims=1 << 18, seed=2)
    out = both_paths("train_classifier", f"-loss logloss -opt {opt} -mini_batch 4096 -dims {1 << 18} -eta0 0.05", data)
    assert out["1"][3] < out["0"][3]


def test_ragged_rows_no_hot_features():
    data = synth.ragged(20000, dims=30000, max_nnz=40, seed=8, regression=True)
    both_paths("train_regressor", "-loss squared -opt adam -mini_batch 2500 -dims 30000 -eta0 0.01", data,
               pushes=[7000, 1, 12999])


def test_more_than_64_mini_batches_in_a_block_and_iters():
    """70 mini-batches: two launches of the grid per pass; -iters 2 replays them from the kept sort results."""
    data = synth.kdd12_like(70 * 1100 + 17, dims=1 << 16, seed=4)
    both_paths("train_classifier", f"-loss hinge -opt adagrad -mini_batch 1100 -dims {1 << 16} -iters 2 -grad_norm count", data)


def test_block_path_is_deterministic():
    data = synth.criteo_like(5 * 8192, dims=1 << 20, seed=6)
    options = f"-loss logloss -opt ftrl -mini_batch 8192 -dims {1 << 20} -alpha 0.05"
    _, w1, s1 = run_gpu_linear("train_classifier", options, data)
    _, w2, s2 = run_gpu_linear("train_classifier", options, data)
    assert np.array_equal(w1, w2) and np.array_equal(s1, s2)
