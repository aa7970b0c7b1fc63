"""ctypes binding of the CPU oracle (oracle/hm_oracle.c).  Test infrastructure only:
nothing under hivemall_b200/ imports this."""
from __future__ import annotations

import ctypes as C
import os

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
ORACLE_LIB = os.path.join(ROOT, "oracle", "_build", "libhm_oracle.so")


class HmoConfig(C.Structure):
    _fields_ = [
        ("task", C.c_int32), ("loss", C.c_int32), ("opt", C.c_int32), ("eta_sched", C.c_int32),
        ("reg", C.c_int32), ("grad_norm", C.c_int32), ("mini_batch", C.c_int32), ("iters", C.c_int32),
        ("dims", C.c_int64), ("total_steps", C.c_double),
        ("eta0", C.c_float), ("power_t", C.c_float), ("lambda_", C.c_float), ("eps", C.c_float),
        ("rho", C.c_float), ("beta1", C.c_float), ("beta2", C.c_float),
        ("alpha", C.c_float), ("beta", C.c_float), ("lambda1", C.c_float), ("lambda2", C.c_float),
        ("factors", C.c_int32), ("fm_classification", C.c_int32), ("sigma", C.c_float),
        ("lambda_w0", C.c_float), ("lambda_w", C.c_float), ("lambda_v", C.c_float),
        ("seed", C.c_uint32), ("pa_c", C.c_float), ("pa_epsilon", C.c_float), ("cov_r", C.c_float), ("cov_phi", C.c_float), ("loss_param", C.c_float),
    ]


_P = C.c_void_p
_lib = None


def lib() -> C.CDLL:
    global _lib
    if _lib is None:
        if not os.path.exists(ORACLE_LIB):
            import subprocess
            subprocess.check_call(["make", "-C", os.path.join(ROOT, "oracle")])
        L = C.CDLL(ORACLE_LIB)
        L.hmo_murmur3_x86_32.restype = C.c_uint32
        L.hmo_murmur3_x86_32.argtypes = [C.c_char_p, C.c_int64, C.c_uint32]
        L.hmo_fold.restype = C.c_int32
        L.hmo_fold.argtypes = [C.c_uint32, C.c_int32]
        L.hmo_mhash_batch.restype = None
        L.hmo_mhash_batch.argtypes = [_P, _P, C.c_int64, C.c_uint32, C.c_int32, _P]
        L.hmo_exp.restype = C.c_float
        L.hmo_exp.argtypes = [C.c_float]
        L.hmo_dloss.restype = C.c_float
        L.hmo_dloss.argtypes = [C.c_int32, C.c_float, C.c_float]
        L.hmo_loss.restype = C.c_double
        L.hmo_loss.argtypes = [C.c_int32, C.c_float, C.c_float]
        L.hmo_eta.restype = C.c_float
        L.hmo_eta.argtypes = [C.POINTER(HmoConfig), C.c_int64]
        L.hmo_train_linear.restype = C.c_int
        L.hmo_train_linear.argtypes = [C.POINTER(HmoConfig), C.c_int64, _P, _P, _P, _P, _P, _P, _P,
                                       C.POINTER(C.c_int64), _P, C.POINTER(C.c_double)]
        L.hmo_train_linear_mt.restype = C.c_int
        L.hmo_train_linear_mt.argtypes = [C.POINTER(HmoConfig), C.c_int64, _P, _P, _P, _P, _P, _P, _P,
                                          C.POINTER(C.c_int64), C.c_int]
        L.hmo_predict_linear.restype = None
        L.hmo_predict_linear.argtypes = [C.c_int64, _P, _P, _P, _P, _P]
        L.hmo_fm_init_v.restype = None
        L.hmo_fm_init_v.argtypes = [C.POINTER(HmoConfig), _P]
        L.hmo_train_fm.restype = C.c_int
        L.hmo_train_fm.argtypes = [C.POINTER(HmoConfig), C.c_int64, _P, _P, _P, _P, _P, _P, _P, _P, _P, _P, _P,
                                   C.POINTER(C.c_int64), _P, C.POINTER(C.c_double)]
        L.hmo_predict_fm.restype = None
        L.hmo_predict_fm.argtypes = [C.POINTER(HmoConfig), C.c_int64, _P, _P, _P, C.c_float, _P, _P, _P]
        L.hmo_model_table.restype = C.c_int64
        L.hmo_model_table.argtypes = [_P, C.c_int64, _P, _P, C.c_int64]
        _lib = L
    return _lib


def _p(a):
    return C.c_void_p(a.ctypes.data) if a is not None else C.c_void_p(0)


def config_from(hm_cfg) -> HmoConfig:
    """Copy the numeric fields of the product's parsed hm_config into an oracle config."""
    c = HmoConfig()
    for name, _ in HmoConfig._fields_:
        setattr(c, name, getattr(hm_cfg, name))
    if c.task == 2:  # HM_TASK_FM has no oracle task id; the FM entry points do not read it
        c.task = 1
    return c


# Option strings without the product library: the same grammar and defaults as
# hivemall_b200/csrc/hm_options.cpp, restated so that the CPU arm of bench.py (and nothing else)
# can build an oracle config while its process maps only oracle/.  tests/test_options.py checks
# the two parsers against each other.
_ENUMS = {
    "loss": {"logloss": 0, "logistic": 0, "log": 0, "hinge": 1, "hingeloss": 1, "squared": 2, "squaredloss": 2,
             "squared_loss": 2,
             "squared_hinge": 15, "squaredhinge": 15, "squaredhingeloss": 15,
             "modified_huber": 16, "modifiedhuber": 16, "modifiedhuberloss": 16,
             "huber": 17, "huberloss": 17, "quantile": 18, "quantileloss": 18,
             "epsilon_insensitive": 19, "epsiloninsensitive": 19, "epsiloninsensitiveloss": 19,
             "squared_epsilon_insensitive": 20, "squaredepsiloninsensitive": 20, "squaredepsiloninsensitiveloss": 20},
    "opt": {"sgd": 0, "adagrad": 1, "adadelta": 2, "adam": 3, "ftrl": 4},
    "eta_sched": {"fixed": 0, "inverse": 1, "inv": 1, "invscaling": 1, "simple": 2},
    "reg": {"no": 0, "none": 0, "l1": 1, "l2": 2},
    "grad_norm": {"batch": 0, "count": 1, "sum": 2},
}
_OPTION_FIELDS = {
    "-loss": ("loss", "enum"), "-loss_function": ("loss", "enum"), "-opt": ("opt", "enum"), "-optimizer": ("opt", "enum"),
    "-mini_batch": ("mini_batch", int), "-mini_batch_size": ("mini_batch", int),
    "-iters": ("iters", int), "-iter": ("iters", int), "-iterations": ("iters", int),
    "-dims": ("dims", int), "-feature_dimensions": ("dims", int), "-p": ("dims", int), "-num_features": ("dims", int),
    "-eta0": ("eta0", float), "-eta_0": ("eta0", float), "-eta": ("eta_sched", "enum"), "-power_t": ("power_t", float),
    "-total_steps": ("total_steps", float), "-t": ("total_steps", float),
    "-reg": ("reg", "enum"), "-regularization": ("reg", "enum"), "-lambda": ("lambda_", float), "-eps": ("eps", float),
    "-rho": ("rho", float), "-decay": ("rho", float), "-beta1": ("beta1", float), "-beta2": ("beta2", float),
    "-alpha": ("alpha", float), "-beta": ("beta", float), "-lambda1": ("lambda1", float), "-l1": ("lambda1", float),
    "-lambda2": ("lambda2", float), "-l2": ("lambda2", float), "-grad_norm": ("grad_norm", "enum"),
    "-factors": ("factors", int), "-factor": ("factors", int), "-f": ("factors", int), "-k": ("factors", int),
    "-sigma": ("sigma", float), "-lambda0": ("lambda_w0", float), "-lambda_w0": ("lambda_w0", float),
    "-lambdaW": ("lambda_w", float), "-lambda_w": ("lambda_w", float), "-lambdaV": ("lambda_v", float),
    "-lambda_v": ("lambda_v", float), "-seed": ("seed", int),
    "-hash_seed": (None, int), "-mix_interval": (None, int),     # not oracle fields
    "-aggressiveness": ("pa_c", float), "-epsilon": ("pa_epsilon", float), "-e": ("pa_epsilon", float),
    "-r": ("cov_r", float), "-phi": ("cov_phi", float),
    "-huber_c": ("_huber_c", float), "-quantile": ("_quantile", float), "-tau": ("_quantile", float),
}
# online learners on the mini-batch skeleton: function -> (task, loss id)
_ONLINE = {"train_perceptron": (0, 3), "train_pa": (0, 4), "train_pa1": (0, 5), "train_pa2": (0, 6),
           "train_pa1_regr": (1, 7), "train_pa2_regr": (1, 8),
           "train_arow": (0, 9), "train_arow_regr": (1, 10), "train_arowe_regr": (1, 11),
           "train_cw": (0, 12), "train_scw": (0, 13), "train_scw2": (0, 14)}
_ONLINE_FIXED = {"-loss", "-loss_function", "-opt", "-optimizer", "-eta", "-eta0", "-eta_0", "-power_t", "-total_steps",
                 "-t", "-reg", "-regularization", "-lambda"}


def _probit(p: float) -> float:
    """Acklam's approximation of the inverse normal distribution function, the operations of hm_options.cpp."""
    import math
    a = (-3.969683028665376e+01, 2.209460984245205e+02, -2.759285104469687e+02, 1.383577518672690e+02,
         -3.066479806614716e+01, 2.506628277459239e+00)
    b = (-5.447609879822406e+01, 1.615858368580409e+02, -1.556989798598866e+02, 6.680131188771972e+01,
         -1.328068155288572e+01)
    c = (-7.784894002430293e-03, -3.223964580411365e-01, -2.400758277161838e+00, -2.549732539343734e+00,
         4.374664141464968e+00, 2.938163982698783e+00)
    d = (7.784695709041462e-03, 3.224671290700398e-01, 2.445134137142996e+00, 3.754408661907416e+00)
    plow = 0.02425
    phigh = 1.0 - plow
    if p < plow:
        q = math.sqrt(-2.0 * math.log(p))
        return (((((c[0] * q + c[1]) * q + c[2]) * q + c[3]) * q + c[4]) * q + c[5]) / \
               ((((d[0] * q + d[1]) * q + d[2]) * q + d[3]) * q + 1.0)
    if p > phigh:
        q = math.sqrt(-2.0 * math.log(1.0 - p))
        return -(((((c[0] * q + c[1]) * q + c[2]) * q + c[3]) * q + c[4]) * q + c[5]) / \
               ((((d[0] * q + d[1]) * q + d[2]) * q + d[3]) * q + 1.0)
    q = p - 0.5
    r = q * q
    return (((((a[0] * r + a[1]) * r + a[2]) * r + a[3]) * r + a[4]) * r + a[5]) * q / \
           (((((b[0] * r + b[1]) * r + b[2]) * r + b[3]) * r + b[4]) * r + 1.0)


def config_from_options(function_name: str, options: str = "") -> HmoConfig:
    """Oracle config of a UDTF option string (no product code involved)."""
    online = _ONLINE.get(function_name)
    task = online[0] if online else {"train_classifier": 0, "train_regressor": 1, "train_fm": 2}[function_name]
    takes_c = bool(online) and online[1] in (5, 6, 7, 8, 13, 14)
    takes_eps = bool(online) and online[1] in (7, 8, 11)
    takes_r = bool(online) and 9 <= online[1] <= 11
    takes_phi = bool(online) and online[1] >= 12
    c = HmoConfig()
    c.task = task
    c.pa_c, c.pa_epsilon, c.cov_r, c.cov_phi, c.loss_param = 1.0, 0.1, 0.1, 1.0, 0.0
    extra = {"_huber_c": 1.0, "_quantile": 0.5}
    c.loss = 2 if task == 1 else 0
    c.opt = 0 if task == 2 else 1
    c.eta_sched, c.reg, c.grad_norm, c.mini_batch, c.iters = 0, 0, 0, 1, 1
    c.dims, c.total_steps = 1 << 24, 10000.0
    c.eta0, c.power_t, c.lambda_, c.eps, c.rho = (0.05 if task == 2 else 0.1), 0.1, 1e-4, 1e-6, 0.95
    c.beta1, c.beta2, c.alpha, c.beta, c.lambda1, c.lambda2 = 0.9, 0.999, 0.1, 1.0, 0.001, 0.0
    c.factors, c.fm_classification, c.sigma = 5, 0, 0.1
    c.lambda_w0 = c.lambda_w = c.lambda_v = 0.01
    c.seed = 31
    if online:
        c.loss, c.opt, c.eta0 = online[1], (5 if online[1] >= 9 else 0), 1.0
    tok = options.split()
    i = 0
    while i < len(tok):
        name = tok[i]
        i += 1
        if name == "-classification" or (name == "-c" and not online):
            if task != 2:
                raise ValueError(f"{name} is a train_fm option")
            c.fm_classification = 1
            continue
        if name in ("-dense", "-densemodel"):
            continue
        if takes_phi and name == "-eta":
            if i >= len(tok):
                raise ValueError("option '-eta' needs a value")
            eta = float(tok[i])
            i += 1
            if not 0.5000001 <= eta <= 0.9999999:
                raise ValueError("option -eta: a probability in (0.5, 1)")
            c.cov_phi = _probit(eta)
            continue
        if online and name in _ONLINE_FIXED:
            raise ValueError(f"{name} is fixed by {function_name}")
        if name in ("-c", "-aggressiveness") and not takes_c:
            raise ValueError(f"{name} is an option of train_pa1, train_pa2, their _regr forms, train_scw and train_scw2")
        if name in ("-epsilon", "-e") and not takes_eps and not (not online and task == 1):
            raise ValueError(f"{name} is an option of train_pa1_regr, train_pa2_regr and train_arowe_regr")
        if name in ("-huber_c", "-quantile", "-tau") and (online or task == 2):
            raise ValueError(f"{name} goes with -loss huber / quantile")
        if name == "-phi" and not takes_phi:
            raise ValueError("-phi is an option of train_cw, train_scw and train_scw2")
        if name == "-r" and not takes_r:
            raise ValueError("-r is an option of train_arow, train_arow_regr and train_arowe_regr")
        if name == "-c":
            name = "-aggressiveness"
        if name not in _OPTION_FIELDS:
            raise ValueError(f"unknown option {name!r}")
        if i >= len(tok):
            raise ValueError(f"option {name!r} needs a value")
        field, kind = _OPTION_FIELDS[name]
        v = tok[i]
        i += 1
        if field is None:
            continue
        if field.startswith("_"):
            extra[field] = float(v)
            continue
        if kind == "enum":
            table = _ENUMS[field]
            if v.lower() not in table:
                raise ValueError(f"option {name}: unknown value {v!r}")
            setattr(c, field, table[v.lower()])
        else:
            setattr(c, field, kind(v))
    if task == 1 and not online and c.loss not in (2, 17, 18, 19, 20):
        raise ValueError("train_regressor takes -loss squared, huber, quantile, epsilon_insensitive or squared_epsilon_insensitive")
    if task == 0 and not online and c.loss >= 17:
        raise ValueError("train_classifier takes -loss logloss, hinge, squared_hinge, modified_huber or squared")
    if not online:
        c.loss_param = extra["_huber_c"] if c.loss == 17 else (extra["_quantile"] if c.loss == 18 else 0.0)
    if task == 2:
        c.loss = 0 if c.fm_classification else 2
        c.task = 1          # (as config_from: the FM entry points do not read it)
    return c


def murmur3(data: bytes, seed: int) -> int:
    return lib().hmo_murmur3_x86_32(data, len(data), seed & 0xFFFFFFFF)


def mhash(strings, num_features: int, seed: int) -> np.ndarray:
    enc = [s.encode("utf-8") if isinstance(s, str) else bytes(s) for s in strings]
    off = np.zeros(len(enc) + 1, dtype=np.int64)
    if enc:
        off[1:] = np.cumsum([len(b) for b in enc])
    data = np.frombuffer(b"".join(enc) + b"\0", dtype=np.uint8).copy()
    out = np.empty(len(enc), dtype=np.int32)
    lib().hmo_mhash_batch(_p(data), _p(off), len(enc), seed & 0xFFFFFFFF, num_features, _p(out))
    return out


class LinearResult:
    def __init__(self, w, s1, s2, t, margins, loss_sum):
        self.w, self.s1, self.s2, self.t, self.margins, self.loss_sum = w, s1, s2, t, margins, loss_sum

    def state(self, width: int) -> np.ndarray:
        """Optimizer slots interleaved the way the product stores them."""
        if width == 0:
            return np.zeros(0, dtype=np.float32)
        if width == 1:
            return self.s1.copy()
        return np.stack([self.s1, self.s2], axis=1).reshape(-1).copy()


def train_linear(cfg: HmoConfig, rowptr, feature, value, label, w=None, s1=None, s2=None, t=0) -> LinearResult:
    rowptr = np.ascontiguousarray(rowptr, dtype=np.int64)
    feature = np.ascontiguousarray(feature).astype(np.uint32)
    value = np.ascontiguousarray(value, dtype=np.float32)
    label = np.ascontiguousarray(label, dtype=np.float32)
    n = len(rowptr) - 1
    d = int(cfg.dims)
    w = np.zeros(d, np.float32) if w is None else np.array(w, np.float32)
    # (the AROW learners' slot is a variance: 1 at the start)
    s1 = np.full(d, 1.0 if cfg.opt == 5 else 0.0, np.float32) if s1 is None else np.array(s1, np.float32)
    s2 = np.zeros(d, np.float32) if s2 is None else np.array(s2, np.float32)
    margins = np.zeros(n, np.float32)
    tt = C.c_int64(t)
    ls = C.c_double(0.0)
    rc = lib().hmo_train_linear(C.byref(cfg), n, _p(rowptr), _p(feature), _p(value), _p(label), _p(w), _p(s1),
                                _p(s2), C.byref(tt), _p(margins), C.byref(ls))
    if rc != 0:
        raise RuntimeError(f"hmo_train_linear failed: {rc}")
    return LinearResult(w, s1, s2, tt.value, margins, ls.value)


def train_linear_mt(cfg: HmoConfig, rowptr, feature, value, label, threads: int) -> LinearResult:
    """OpenMP variant used as the timed CPU baseline; inputs must already be contiguous
    int64 / uint32-compatible / float32 arrays (no copies are made of the big ones)."""
    rowptr = np.ascontiguousarray(rowptr, dtype=np.int64)
    feature = np.ascontiguousarray(feature)
    assert feature.dtype in (np.int32, np.uint32)
    value = np.ascontiguousarray(value, dtype=np.float32)
    label = np.ascontiguousarray(label, dtype=np.float32)
    d = int(cfg.dims)
    w, s1, s2 = np.zeros(d, np.float32), np.full(d, 1.0 if cfg.opt == 5 else 0.0, np.float32), np.zeros(d, np.float32)
    tt = C.c_int64(0)
    rc = lib().hmo_train_linear_mt(C.byref(cfg), len(rowptr) - 1, _p(rowptr), _p(feature), _p(value), _p(label),
                                   _p(w), _p(s1), _p(s2), C.byref(tt), threads)
    if rc != 0:
        raise RuntimeError(f"hmo_train_linear_mt failed: {rc}")
    return LinearResult(w, s1, s2, tt.value, None, 0.0)


def predict_linear(rowptr, feature, value, w) -> np.ndarray:
    rowptr = np.ascontiguousarray(rowptr, dtype=np.int64)
    feature = np.ascontiguousarray(feature).astype(np.uint32)
    value = np.ascontiguousarray(value, dtype=np.float32)
    w = np.ascontiguousarray(w, dtype=np.float32)
    out = np.empty(len(rowptr) - 1, np.float32)
    lib().hmo_predict_linear(len(out), _p(rowptr), _p(feature), _p(value), _p(w), _p(out))
    return out


def model_table(w):
    w = np.ascontiguousarray(w, dtype=np.float32)
    n = lib().hmo_model_table(_p(w), len(w), None, None, 0)
    f = np.empty(n, np.int32)
    wt = np.empty(n, np.float32)
    lib().hmo_model_table(_p(w), len(w), _p(f), _p(wt), n)
    return f, wt


class FmResult:
    pass


def fm_init_v(cfg: HmoConfig) -> np.ndarray:
    V = np.empty(int(cfg.dims) * int(cfg.factors), np.float32)
    lib().hmo_fm_init_v(C.byref(cfg), _p(V))
    return V


def train_fm(cfg: HmoConfig, rowptr, feature, value, label) -> FmResult:
    rowptr = np.ascontiguousarray(rowptr, dtype=np.int64)
    feature = np.ascontiguousarray(feature).astype(np.uint32)
    value = np.ascontiguousarray(value, dtype=np.float32)
    label = np.ascontiguousarray(label, dtype=np.float32)
    n = len(rowptr) - 1
    d, k = int(cfg.dims), int(cfg.factors)
    r = FmResult()
    r.w0 = np.zeros(1, np.float32)
    r.gw0 = np.zeros(1, np.float32)
    r.w = np.zeros(d, np.float32)
    r.gw = np.zeros(d, np.float32)
    r.V = fm_init_v(cfg)
    r.gv = np.zeros(d * k, np.float32)
    r.touched = np.zeros(d, np.uint8)
    r.preds = np.zeros(n, np.float32)
    tt = C.c_int64(0)
    ls = C.c_double(0.0)
    rc = lib().hmo_train_fm(C.byref(cfg), n, _p(rowptr), _p(feature), _p(value), _p(label), _p(r.w0), _p(r.w),
                            _p(r.V), _p(r.gw0), _p(r.gw), _p(r.gv), _p(r.touched), C.byref(tt), _p(r.preds),
                            C.byref(ls))
    if rc != 0:
        raise RuntimeError(f"hmo_train_fm failed: {rc}")
    r.t, r.loss_sum = tt.value, ls.value
    return r


def predict_fm(cfg: HmoConfig, rowptr, feature, value, w0, w, V) -> np.ndarray:
    rowptr = np.ascontiguousarray(rowptr, dtype=np.int64)
    feature = np.ascontiguousarray(feature).astype(np.uint32)
    value = np.ascontiguousarray(value, dtype=np.float32)
    out = np.empty(len(rowptr) - 1, np.float32)
    lib().hmo_predict_fm(C.byref(cfg), len(out), _p(rowptr), _p(feature), _p(value), float(w0),
                         _p(np.ascontiguousarray(w, np.float32)), _p(np.ascontiguousarray(V, np.float32)), _p(out))
    return out
