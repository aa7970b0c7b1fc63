"""Mid-size mini-batches launch forward and update as programmatic dependents (a kernel starts while its
predecessor is still running and synchronises at griddepcontrol.wait).  That changes when things run, not
what is computed: the result must be bit-identical to plain stream-ordered launches (HM_PDL=0), run after
run -- a stale read across the overlap would show up here."""
import os

import numpy as np
import pytest

import hivemall_b200 as hm
from hivemall_b200 import synth

from _util import run_gpu_linear

pytestmark = pytest.mark.gpu


def run(fn, options, data, pdl, pushes=None):
    os.environ["HM_PDL"] = "1" if pdl else "0"
    try:
        u, w, state = run_gpu_linear(fn, options, data, pushes=pushes)
    finally:
        del os.environ["HM_PDL"]
    u.destroy()
    return w, state


@pytest.mark.parametrize("opt", ["sgd", "adagrad", "adadelta", "adam", "ftrl"])
def test_same_bits_with_and_without_overlap(opt):
    data = synth.criteo_like(40 * 1024 + 99, dims=1 << 16, seed=12)     # small model: every mini-batch rewrites hot records
    options = f"-loss logloss -opt {opt} -mini_batch 1024 -dims {1 << 16} -eta0 0.05"
    w0, s0 = run("train_classifier", options, data, pdl=False)
    for rep in range(3):
        w1, s1 = run("train_classifier", options, data, pdl=True)
        assert np.array_equal(w0, w1) and np.array_equal(s0, s1), f"{opt}: run {rep} differs from the stream-ordered result"


def test_no_hot_features_and_several_pushes():
    data = synth.ragged(30000, dims=2000, max_nnz=30, seed=14, regression=True)
    options = "-loss squared -opt adagrad -mini_batch 700 -dims 2000 -eta0 0.01 -iters 2"
    w0, s0 = run("train_regressor", options, data, pdl=False, pushes=[12345, 17655])
    w1, s1 = run("train_regressor", options, data, pdl=True, pushes=[12345, 17655])
    assert np.array_equal(w0, w1) and np.array_equal(s0, s1)
