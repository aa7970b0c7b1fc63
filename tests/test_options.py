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
    ("train_regressor", "-loss squared_hinge"),
    ("train_classifier", "-loss huber"),
    ("train_classifier", "-loss quantile"),
    ("train_classifier", "-epsilon 0.1"),
    ("train_regressor", "-quantile 1.5"),
    ("train_pa1", "-huber_c 1"),
    ("train_fm", "-opt adam"),
    ("train_fm", "-loss logloss"),
    ("train_everything", ""),
    ("train_pa", "-c 1"),                       # plain PA has no aggressiveness
    ("train_perceptron", "-eta0 0.5"),          # fixed by the function
    ("train_pa1", "-loss hinge"),
    ("train_pa1", "-opt adagrad"),
    ("train_pa1", "-epsilon 0.1"),              # regression only
    ("train_pa2", "-c 0"),
    ("train_pa2", "-c"),
    ("train_classifier", "-aggressiveness 1"),
    ("train_arow", "-c 1"),
    ("train_arow", "-epsilon 0.1"),             # only the regression with an insensitive zone has it
    ("train_arow", "-r 0"),
    ("train_pa1", "-r 1"),
    ("train_arow_regr", "-opt sgd"),
    ("train_cw", "-c 1"),                       # plain CW has no aggressiveness
    ("train_cw", "-r 1"),
    ("train_scw", "-phi 0"),
    ("train_cw", "-eta 0.5"),                   # a probability strictly above one half
    ("train_cw", "-eta fixed"),
    ("train_arow", "-phi 1"),
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
    ("train_perceptron", "-mini_batch 16 -dims 1000 -iters 2"),
    ("train_pa", "-mini_batch 1 -grad_norm count"),
    ("train_pa1", "-c 0.25 -mini_batch 128 -dims 4096 -mix_interval 4"),
    ("train_pa2", "-aggressiveness 3"),
    ("train_pa1_regr", "-c 0.5 -epsilon 0.01 -mini_batch 64"),
    ("train_pa2_regr", "-e 0.2"),
    ("train_arow", "-r 0.5 -mini_batch 32"),
    ("train_arow_regr", ""),
    ("train_arowe_regr", "-r 2 -epsilon 0.3 -dims 1024"),
    ("train_cw", "-phi 0.7 -mini_batch 16"),
    ("train_scw", "-phi 1.2 -c 0.5"),
    ("train_scw2", ""),
    ("train_cw", "-eta 0.85"),
    ("train_scw", "-eta 0.99 -c 0.5"),
    ("train_scw2", "-eta 0.51"),
    ("train_classifier", "-loss squared_hinge -opt sgd"),
    ("train_classifier", "-loss ModifiedHuber"),
    ("train_regressor", "-loss huber -huber_c 0.7"),
    ("train_regressor", "-loss quantile -tau 0.9 -mini_batch 32"),
    ("train_regressor", "-loss epsilon_insensitive -epsilon 0.3"),
    ("train_regressor", "-loss squared_epsilon_insensitive"),
])
def test_oracle_side_parser_agrees_with_the_library(fn, opts):
    """bench.py's CPU arm builds its oracle config with tests/_oracle.config_from_options so that its
    process never loads the product library; the two parsers must agree field by field."""
    import _oracle
    a = _oracle.config_from(hm.parse_options(fn, opts))
    b = _oracle.config_from_options(fn, opts)
    for name, _ in _oracle.HmoConfig._fields_:
        assert getattr(a, name) == getattr(b, name), name


def test_online_learner_functions():
    """The functions fix loss, optimizer and step size; C and epsilon have the stated defaults."""
    for fn, task, loss in [("train_perceptron", 0, 3), ("train_pa", 0, 4), ("train_pa1", 0, 5), ("train_pa2", 0, 6),
                           ("train_pa1_regr", 1, 7), ("train_pa2_regr", 1, 8)]:
        c = hm.parse_options(fn, "")
        assert (c.task, c.loss, c.opt, c.eta_sched, c.reg, c.mini_batch) == (task, loss, 0, 0, 0, 1), fn
        assert c.eta0 == 1.0 and c.pa_c == 1.0 and abs(c.pa_epsilon - 0.1) < 1e-8
    c = hm.parse_options("train_pa1_regr", "-c 0.25 -e 0.5")
    assert c.pa_c == 0.25 and c.pa_epsilon == 0.5


def test_variance_keeping_learner_functions():
    for fn, task, loss in [("train_arow", 0, 9), ("train_arow_regr", 1, 10), ("train_arowe_regr", 1, 11)]:
        c = hm.parse_options(fn, "")
        assert (c.task, c.loss, c.opt, c.mini_batch) == (task, loss, 5, 1), fn
        assert abs(c.cov_r - 0.1) < 1e-8 and c.eta0 == 1.0
    assert hm.parse_options("train_arow", "-r 3").cov_r == 3.0
    for fn, loss in [("train_cw", 12), ("train_scw", 13), ("train_scw2", 14)]:
        c = hm.parse_options(fn, "")
        assert (c.task, c.loss, c.opt, c.cov_phi, c.pa_c) == (0, loss, 5, 1.0, 1.0), fn
    assert hm.parse_options("train_scw", "-phi 0.5 -c 0.25").cov_phi == 0.5


def test_further_plain_losses():
    c = hm.parse_options("train_regressor", "-loss huber")
    assert (c.loss, c.loss_param) == (17, 1.0)
    c = hm.parse_options("train_regressor", "-loss quantile")
    assert (c.loss, c.loss_param) == (18, 0.5)
    c = hm.parse_options("train_regressor", "-loss epsilon_insensitive -epsilon 0.25")
    assert (c.loss, c.pa_epsilon, c.loss_param) == (19, 0.25, 0.0)
    assert hm.parse_options("train_classifier", "-loss squaredhinge").loss == 15
    assert hm.parse_options("train_classifier", "-loss modified_huber").loss == 16


def test_confidence_as_a_probability():
    """-eta of the confidence-weighted learners is a probability: phi = probit(eta)."""
    from scipy.stats import norm
    for eta in (0.51, 0.7, 0.85, 0.975, 0.999):
        phi = hm.parse_options("train_cw", f"-eta {eta}").cov_phi
        assert abs(phi - norm.ppf(eta)) < 1e-6 * max(1.0, norm.ppf(eta))
    assert abs(hm.parse_options("train_scw", "-eta 0.975").cov_phi - 1.959964) < 1e-5
