"""Randomised differential test of the linear learner: option strings, shapes and push patterns drawn
from a seeded generator, CUDA path vs. CPU oracle on identical rows (ragged rows with empty ones,
mini-batches from 1 row to larger than the input, dims from 17 to 2^20, every loss / optimizer /
regulariser / learning-rate schedule / gradient normalisation, one or two passes)."""
import numpy as np
import pytest

from hivemall_b200 import synth

from _util import assert_weights_close, run_gpu_linear, run_oracle_linear

pytestmark = pytest.mark.gpu


def draw_case(seed: int):
    r = np.random.default_rng(seed)
    fn = r.choice(["train_classifier", "train_regressor"])
    loss = r.choice(["logloss", "hinge", "squared"]) if fn == "train_classifier" else "squared"
    opt = r.choice(["sgd", "adagrad", "adadelta", "adam", "ftrl"])
    dims = int(r.choice([17, 300, 5000, (1 << 17) + 1, 1 << 20]))
    rows = int(r.integers(1, 6000))
    mb = int(r.choice([1, 7, 64, 333, 1024, 2500, 8192]))
    if mb == 1:
        rows = min(rows, 300)
    max_nnz = int(r.choice([1, 3, 20, 60]))
    opts = [f"-loss {loss}", f"-opt {opt}", f"-mini_batch {mb}", f"-dims {dims}", f"-iters {int(r.choice([1, 1, 2]))}"]
    opts.append(f"-eta {r.choice(['fixed', 'inverse', 'simple'])} -eta0 {r.choice([0.01, 0.05, 0.2])} -total_steps 100")
    if opt != "ftrl":
        reg = r.choice(["no", "l1", "l2"])
        opts.append(f"-reg {reg} -lambda {r.choice([1e-4, 1e-2])}")
    else:
        opts.append(f"-alpha {r.choice([0.05, 0.5])} -lambda1 {r.choice([0.0, 1e-3])} -lambda2 {r.choice([0.0, 1e-2])}")
    opts.append(f"-grad_norm {r.choice(['batch', 'count', 'sum'])}")
    pushes = None
    if r.random() < 0.5:       # ragged pushes: mini-batch boundaries must not depend on them
        cuts = np.sort(r.integers(0, rows + 1, size=int(r.integers(1, 5))))
        pushes = [int(b - a) for a, b in zip(np.concatenate([[0], cuts]), np.concatenate([cuts, [rows]])) if b > a]
    return fn, " ".join(opts), dims, rows, max_nnz, pushes


@pytest.mark.parametrize("seed", range(40))
def test_random_configuration_matches_the_oracle(seed):
    fn, options, dims, rows, max_nnz, pushes = draw_case(seed)
    data = synth.ragged(rows, dims, max_nnz, seed=100 + seed, regression=(fn == "train_regressor"))
    u, w, state = run_gpu_linear(fn, options, data, pushes=pushes)
    ref, ref_state = run_oracle_linear(fn, options, data)
    assert u.step == ref.t, options
    assert_weights_close(w, ref.w, f"w [{fn} {options}]")
    assert_weights_close(state, ref_state, f"state [{fn} {options}]")
    u.destroy()


def draw_fm_case(seed: int):
    r = np.random.default_rng(1000 + seed)
    k = int(r.choice([1, 3, 4, 5, 8, 16, 33, 64]))
    opt = r.choice(["sgd", "adagrad"])
    cls = bool(r.random() < 0.5)
    dims = int(r.choice([50, 3000, (1 << 16) + 7]))
    rows = int(r.integers(1, 2500))
    mb = int(r.choice([1, 16, 200, 1024, 4096]))
    if mb == 1:
        rows = min(rows, 150)
    max_nnz = int(r.choice([1, 4, 30]))
    opts = [f"-factors {k}", f"-opt {opt}", f"-mini_batch {mb}", f"-dims {dims}", f"-iters {int(r.choice([1, 1, 2]))}",
            f"-eta0 {r.choice([0.01, 0.05])}", f"-sigma {r.choice([0.05, 0.3])}", f"-seed {int(r.integers(0, 99))}",
            f"-lambda0 {r.choice([0.0, 0.01])} -lambdaW {r.choice([0.0, 0.01])} -lambdaV {r.choice([0.0, 0.01])}",
            f"-grad_norm {r.choice(['batch', 'count', 'sum'])}"]
    if cls:
        opts.append("-classification")
    pushes = None
    if r.random() < 0.5:
        cuts = np.sort(r.integers(0, rows + 1, size=int(r.integers(1, 4))))
        pushes = [int(b - a) for a, b in zip(np.concatenate([[0], cuts]), np.concatenate([cuts, [rows]])) if b > a]
    return " ".join(opts), dims, rows, max_nnz, pushes, cls, opt == "adagrad"


@pytest.mark.parametrize("seed", range(24))
def test_random_fm_configuration_matches_the_oracle(seed):
    from test_gpu_fm import check, run_both
    options, dims, rows, max_nnz, pushes, cls, adagrad = draw_fm_case(seed)
    data = synth.ragged(rows, dims, max_nnz, seed=200 + seed, regression=not cls)
    u, ref, _ = run_both(options, data, pushes=pushes)
    check(u, ref, adagrad)
    u.destroy()


ONLINE = {  # function -> (regression?, option generator)
    "train_perceptron": (False, lambda r: ""),
    "train_pa": (False, lambda r: ""),
    "train_pa1": (False, lambda r: f"-c {r.choice([0.01, 0.3, 5.0])}"),
    "train_pa2": (False, lambda r: f"-c {r.choice([0.01, 0.3, 5.0])}"),
    "train_pa1_regr": (True, lambda r: f"-c {r.choice([0.05, 1.0])} -epsilon {r.choice([0.0, 0.05, 0.5])}"),
    "train_pa2_regr": (True, lambda r: f"-c {r.choice([0.05, 1.0])} -epsilon {r.choice([0.0, 0.05, 0.5])}"),
    "train_arow": (False, lambda r: f"-r {r.choice([0.1, 1.0, 10.0])}"),
    "train_arow_regr": (True, lambda r: f"-r {r.choice([0.1, 1.0, 10.0])}"),
    "train_arowe_regr": (True, lambda r: f"-r {r.choice([0.1, 1.0])} -epsilon {r.choice([0.01, 0.3])}"),
    "train_cw": (False, lambda r: f"-phi {r.choice([0.2, 1.0, 2.0])}"),
    "train_scw": (False, lambda r: f"-phi {r.choice([0.2, 1.0, 2.0])} -c {r.choice([0.05, 1.0])}"),
    "train_scw2": (False, lambda r: f"-phi {r.choice([0.2, 1.0, 2.0])} -c {r.choice([0.05, 1.0])}"),
}


def draw_online_case(seed: int):
    r = np.random.default_rng(5000 + seed)
    fn = list(ONLINE)[seed % len(ONLINE)]
    regression, gen = ONLINE[fn]
    dims = int(r.choice([17, 300, 5000, (1 << 17) + 1]))
    rows = int(r.integers(1, 5000))
    mb = int(r.choice([1, 7, 64, 333, 1024, 2500]))
    if mb == 1:
        rows = min(rows, 300)
    max_nnz = int(r.choice([1, 3, 20, 60]))
    opts = [gen(r), f"-mini_batch {mb}", f"-dims {dims}", f"-iters {int(r.choice([1, 1, 2]))}",
            f"-grad_norm {r.choice(['batch', 'count', 'sum'])}"]
    pushes = None
    if r.random() < 0.5:
        cuts = np.sort(r.integers(0, rows + 1, size=int(r.integers(1, 5))))
        pushes = [int(b - a) for a, b in zip(np.concatenate([[0], cuts]), np.concatenate([cuts, [rows]])) if b > a]
    return fn, " ".join(o for o in opts if o), dims, rows, max_nnz, pushes, regression


def draw_loss_case(seed: int):
    r = np.random.default_rng(9000 + seed)
    losses = [("train_classifier", "squared_hinge", ""), ("train_classifier", "modified_huber", ""),
              ("train_regressor", "huber", f"-huber_c {r.choice([0.1, 1.0, 3.0])}"),
              ("train_regressor", "quantile", f"-quantile {r.choice([0.1, 0.5, 0.9])}"),
              ("train_regressor", "epsilon_insensitive", f"-epsilon {r.choice([0.0, 0.1, 0.5])}"),
              ("train_regressor", "squared_epsilon_insensitive", f"-epsilon {r.choice([0.0, 0.1, 0.5])}")]
    fn, loss, extra = losses[seed % len(losses)]
    opt = r.choice(["sgd", "adagrad", "adadelta", "adam", "ftrl"])
    dims = int(r.choice([17, 300, 5000, (1 << 17) + 1]))
    rows = int(r.integers(1, 5000))
    mb = int(r.choice([1, 7, 64, 333, 1024, 2500]))
    if mb == 1:
        rows = min(rows, 300)
    opts = [f"-loss {loss}", extra, f"-opt {opt}", f"-mini_batch {mb}", f"-dims {dims}", f"-eta0 {r.choice([0.01, 0.05])}",
            f"-grad_norm {r.choice(['batch', 'count', 'sum'])}"]
    return fn, " ".join(o for o in opts if o), dims, rows, int(r.choice([1, 3, 20, 60]))


@pytest.mark.parametrize("seed", range(24))
def test_random_further_losses_match_the_oracle(seed):
    """squared hinge, modified Huber, Huber, quantile, epsilon-insensitive and its square with every optimizer."""
    fn, options, dims, rows, max_nnz = draw_loss_case(seed)
    data = synth.ragged(rows, dims, max_nnz, seed=900 + seed, regression=(fn == "train_regressor"))
    u, w, state = run_gpu_linear(fn, options, data)
    ref, ref_state = run_oracle_linear(fn, options, data)
    assert u.step == ref.t, options
    assert_weights_close(w, ref.w, f"w [{fn} {options}]")
    assert_weights_close(state, ref_state, f"state [{fn} {options}]")
    u.destroy()


@pytest.mark.parametrize("seed", range(36))
def test_random_online_learner_matches_the_oracle(seed):
    """The online learners (perceptron, passive-aggressive, AROW, CW, SCW) under the same randomised shapes:
    weights and, where the learner keeps one, the per-feature variance."""
    fn, options, dims, rows, max_nnz, pushes, regression = draw_online_case(seed)
    data = synth.ragged(rows, dims, max_nnz, seed=700 + seed, regression=regression)
    u, w, state = run_gpu_linear(fn, options, data, pushes=pushes)
    ref, ref_state = run_oracle_linear(fn, options, data)
    assert u.step == ref.t, options
    assert_weights_close(w, ref.w, f"w [{fn} {options}]")
    assert_weights_close(state, ref_state, f"state [{fn} {options}]")
    u.destroy()
