"""UDTF option-string parsing (host logic, no GPU)."""
import pytest

import hivemall_b200 as hm


def test_baseline_config0_string():
    c = hm.parse_options("train_classifier", "-loss logloss -opt AdaGrad -mini_batch 128")
    assert (c.task, c.loss, c.opt, c.mini_batch, c.dims, c.iters) == (0, 0, 1, 128, 1 << 24, 1)


def test_defaults_per_function():
    assert hm.parse_options("train_regressor", "").loss == 2
    fm = hm.parse_options("train_fm", "-factors 16 -classification")
    assert (fm.task, fm.factors, fm.fm_classification, fm.loss) == (2, 16, 1, 0)
    assert hm.parse_options("train_fm", "").loss == 2


@pytest.mark.parametrize("opt,code", [("sgd", 0), ("AdaGrad", 1), ("ADADELTA", 2), ("adam", 3), ("ftrl", 4)])
def test_optimizers(opt, code):
    assert hm.parse_options("train_classifier", f"-opt {opt}").opt == code


def test_numeric_options_round_trip():
    c = hm.parse_options("train_classifier",
                         "-opt ftrl -alpha 0.05 -beta 2 -lambda1 0.5 -lambda2 0.25 -dims 1000 -iters 3 "
                         "-eta inverse -eta0 0.3 -power_t 0.2 -reg l2 -lambda 0.01 -grad_norm count "
                         "-mix_interval 64 -hash_seed 7 -seed 9 -eps 1e-8 -rho 0.9 -beta1 0.8 -beta2 0.99")
    assert c.dims == 1000 and c.iters == 3 and c.eta_sched == 1 and c.reg == 2 and c.grad_norm == 1
    assert c.mix_interval == 64 and c.hash_seed == 7 and c.seed == 9
    assert abs(c.alpha - 0.05) < 1e-9 and c.beta == 2 and c.lambda1 == 0.5 and c.lambda2 == 0.25
    assert abs(c.eta0 - 0.3) < 1e-7 and abs(c.power_t - 0.2) < 1e-7 and abs(c.lambda_ - 0.01) < 1e-9
    assert abs(c.beta1 - 0.8) < 1e-7 and abs(c.beta2 - 0.99) < 1e-7


@pytest.mark.parametrize("fn,opts", [
    ("train_classifier", "-bogus 1"),
    ("train_classifier", "-mini_batch"),
    ("train_classifier", "-mini_batch 0"),
    ("train_classifier", "-mini_batch abc"),
    ("train_classifier", "-opt lbfgs"),
    ("train_classifier", "-loss huber"),
    ("train_classifier", "loss logloss"),
    ("train_classifier", "-factors 4"),
    ("train_regressor", "-loss logloss"),
    ("train_fm", "-opt adam"),
    ("train_fm", "-loss logloss"),
    ("train_everything", ""),
    ("train_classifier", "-eta0 nan"),
    ("train_classifier", "-dims -5"),
])
def test_bad_options_are_argument_errors(fn, opts):
    with pytest.raises(hm.UDFArgumentError):
        hm.parse_options(fn, opts)


def test_error_names_the_offending_token():
    with pytest.raises(hm.UDFArgumentError) as e:
        hm.parse_options("train_classifier", "-opt adagrad -frobnicate 3")
    assert "-frobnicate" in str(e.value)


@pytest.mark.parametrize("fn,opts", [
    ("train_classifier", "-loss logloss -opt AdaGrad -mini_batch 128"),
    ("train_classifier", "-loss logloss -opt ftrl -mini_batch 65536 -dims 16777216 -alpha 0.05 -lambda1 0.0001"),
    ("train_classifier", "-loss hinge -opt adadelta -rho 0.9 -eps 1e-7 -reg l1 -lambda 0.001 -grad_norm sum"),
    ("train_regressor", "-loss squared -opt adam -mini_batch 65536 -dims 67108864 -eta0 0.001 -beta1 0.8"),
    ("train_regressor", "-eta inverse -eta0 0.3 -power_t 0.2 -total_steps 500 -iters 3 -mix_interval 64 -hash_seed 7"),
    ("train_fm", "-factors 16 -classification -opt adagrad -mini_batch 65536 -dims 16777216 -eta0 0.05"),
    ("train_fm", "-factors 8 -sigma 0.2 -lambda0 0.1 -lambdaW 0.2 -lambdaV 0.3 -seed 5"),
    ("train_fm", ""),
])
def test_oracle_side_parser_agrees_with_the_library(fn, opts):
    """bench.py's CPU arm builds its oracle config with tests/_oracle.config_from_options so that its
    process never loads the product library; the two parsers must agree field by field."""
    import _oracle
    a = _oracle.config_from(hm.parse_options(fn, opts))
    b = _oracle.config_from_options(fn, opts)
    for name, _ in _oracle.HmoConfig._fields_:
        assert getattr(a, name) == getattr(b, name), name
