"""hivemall_b200 -- B200-native (sm_100a CUDA) implementation of Hivemall's data-parallel
training hot path (train_classifier / train_regressor / train_fm, mhash), behind a C-ABI.

Only the path is here: csrc/ (kernels + C-ABI), udtf.py (host mirror of the UDTF
interface), mix.py (model averaging across GPUs), synth.py (synthetic inputs for
bench/tests).  See DESIGN.md.  PARITY WITH THE REFERENCE IS UNPINNED: the mounted
reference holds no code.
"""
from ._lib import HivemallError, UDFArgumentError, LIB_PATH  # noqa: F401
from .udtf import (  # noqa: F401
    UDTF, TrainClassifierUDTF, TrainRegressorUDTF, FactorizationMachineUDTF,
    train_classifier, train_regressor, train_fm, mhash, feature_hashing, parse_options,
    host_alloc, host_free, bind_host_to_gpu, DEFAULT_NUM_FEATURES, DEFAULT_HASH_SEED,
    PerceptronUDTF, PassiveAggressiveUDTF, PassiveAggressive1UDTF, PassiveAggressive2UDTF,
    PassiveAggressive1RegressionUDTF, PassiveAggressive2RegressionUDTF,
    train_perceptron, train_pa, train_pa1, train_pa2, train_pa1_regr, train_pa2_regr,
    add_bias, add_bias_csr, BIAS_FEATURE, BIAS_CLAUSE,
    AROWClassifierUDTF, AROWRegressionUDTF, AROWeRegressionUDTF, train_arow, train_arow_regr, train_arowe_regr,
    ConfidenceWeightedUDTF, SoftConfidenceWeighted1UDTF, SoftConfidenceWeighted2UDTF, train_cw, train_scw, train_scw2,
)

__all__ = [
    "UDTF", "TrainClassifierUDTF", "TrainRegressorUDTF", "FactorizationMachineUDTF",
    "train_classifier", "train_regressor", "train_fm", "mhash", "feature_hashing",
    "train_perceptron", "train_pa", "train_pa1", "train_pa2", "train_pa1_regr", "train_pa2_regr",
    "train_arow", "train_arow_regr", "train_arowe_regr", "train_cw", "train_scw", "train_scw2", "add_bias", "add_bias_csr", "parse_options", "host_alloc", "host_free", "bind_host_to_gpu", "HivemallError", "UDFArgumentError",
]
