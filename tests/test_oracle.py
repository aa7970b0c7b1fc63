"""The CPU oracle against (a) the public MurmurHash3 known answers -- the only part of the
path that CAN be pinned, since the reference mount holds no code -- and (b) an independent
numpy restatement of DESIGN.md section 3 on small inputs, (c) its own OpenMP variant."""
import json
import os
import struct

import numpy as np
import pytest

import hivemall_b200 as hm
from hivemall_b200 import synth

import _oracle

GOLDEN = os.path.join(os.path.dirname(__file__), "golden", "murmur3_kat.json")


def test_murmur3_public_vectors():
    kat = json.load(open(GOLDEN))
    assert len(kat["vectors"]) >= 19
    for v in kat["vectors"]:
        assert _oracle.murmur3(bytes.fromhex(v["hex"]), v["seed"]) == v["hash"], v


def test_murmur3_smhasher_verification_value():
    kat = json.load(open(GOLDEN))
    key = bytes(range(256))
    hashes = b"".join(struct.pack("<I", _oracle.murmur3(key[:i], 256 - i)) for i in range(256))
    assert _oracle.murmur3(hashes, 0) == kat["smhasher_verification"] == 0xB0F57EE3


def test_fold_is_java_remainder():
    L = _oracle.lib()
    assert L.hmo_fold(0xFFFFFFF9, 4) == 1          # (int)-7 % 4 = -3 -> +4
    assert L.hmo_fold(7, 4) == 3
    assert L.hmo_fold(0x80000000, 1 << 24) == 0
    for h in [0, 1, 0x7FFFFFFF, 0x80000001, 0xDEADBEEF]:
        assert L.hmo_fold(h, 1 << 24) == (h & 0xFFFFFF)  # power of two: same as masking
        assert 0 <= L.hmo_fold(h, 16777213) < 16777213


def test_exp_accuracy():
    L = _oracle.lib()
    xs = np.linspace(-87, 88, 20001).astype(np.float32)
    got = np.array([L.hmo_exp(float(x)) for x in xs], dtype=np.float64)
    want = np.exp(xs.astype(np.float64))
    assert np.max(np.abs(got - want) / want) < 4e-7
    assert L.hmo_exp(0.0) == 1.0 and L.hmo_exp(-1000.0) > 0 and np.isfinite(L.hmo_exp(1000.0))


# ---- independent numpy restatement of the linear learner (float32 ops via numpy scalars) ----

F = np.float32


def np_exp(x):
    x = F(min(max(x, F(-87.0)), F(88.0)))
    n = F(np.rint(F(x * F(1.44269504))))
    r = F(x - F(n * F(0.693359375)))
    r = F(r - F(n * F(-2.12194440e-4)))
    p = F(1.3888889e-3)
    for c in (8.3333333e-3, 4.1666667e-2, 1.6666667e-1, 0.5, 1.0, 1.0):
        p = F(F(p * r) + F(c))
    return F(p * F(2.0) ** int(n))


def np_dloss(loss, m, y):
    if loss == 0:
        return F(F(-y) / F(F(1.0) + np_exp(F(y * m))))
    if loss == 1:
        return F(-y) if F(y * m) < F(1.0) else F(0.0)
    return F(m - y)


def np_dloss_more(cfg, m, y):
    """squared hinge, modified Huber; Huber, quantile, epsilon-insensitive and its square."""
    loss, p, eps = cfg.loss, F(cfg.loss_param), F(cfg.pa_epsilon)
    if loss in (15, 16):
        z = F(y * m)
        if loss == 15:
            return F(F(F(-2.0) * y) * F(F(1.0) - z)) if z < 1 else F(0.0)
        if z >= 1:
            return F(0.0)
        return F(F(F(-2.0) * y) * F(F(1.0) - z)) if z >= -1 else F(F(-4.0) * y)
    e = F(m - y)
    if loss == 17:
        return e if abs(e) <= p else (F(-p) if e < 0 else p)
    if loss == 18:
        return F(-p) if e < 0 else F(F(1.0) - p)
    a = F(abs(e) - eps)
    if not a > 0:
        return F(0.0)
    if loss == 19:
        return F(-1.0) if e < 0 else F(1.0)
    return F((F(-2.0) if e < 0 else F(2.0)) * a)


def np_row_coef(cfg, m, y, xx):
    """The per-row coefficient for any loss id: the loss derivative, or the online learner's signed step
    (perceptron; passive-aggressive PA / PA-I / PA-II, Crammer et al. 2006, classification and regression)."""
    loss = cfg.loss
    if loss <= 2:
        return np_dloss(loss, m, y)
    if loss >= 15:
        return np_dloss_more(cfg, m, y)
    if loss == 3:
        return F(-y) if F(y * m) <= F(0.0) else F(0.0)
    if loss in (7, 8):
        e = F(y - m)
        l, d = F(abs(e) - F(cfg.pa_epsilon)), (F(-1.0) if e < 0 else F(1.0))
    else:
        l, d = F(F(1.0) - F(y * m)), y
    if not l > 0:
        return F(0.0)
    if loss in (6, 8):
        tau = F(l / F(xx + F(F(0.5) / F(cfg.pa_c))))
    else:
        if not xx > 0:
            return F(0.0)
        tau = F(l / xx)
        if loss != 4:
            tau = min(F(cfg.pa_c), tau)
    return F(-F(tau * d))


def np_cov_coefs(cfg, m, y, v):
    """AROW (Crammer, Kulesza, Dredze 2009), CW (Dredze, Crammer, Pereira 2008), SCW-I / SCW-II (Wang, Zhao, Hoi
    2012), diagonal: (c1, c2) for the weight and the variance sums."""
    Z = (F(0.0), F(0.0))
    if cfg.loss >= 12:
        phi, cc = F(cfg.cov_phi), F(cfg.pa_c)
        if not v > 0:
            return Z
        my, p2 = F(y * m), F(phi * phi)
        if cfg.loss == 12:
            q = F(F(1.0) + F(F(F(2.0) * phi) * my))
            rad = F(F(q * q) - F(F(F(8.0) * phi) * F(my - F(phi * v))))
            if not rad >= 0:
                return Z
            alpha = F(F(np.sqrt(rad) - q) / F(F(F(4.0) * phi) * v))
            if not alpha > 0:
                return Z
            return F(-F(alpha * y)), F(F(F(2.0) * alpha) * phi)
        if not F(F(phi * np.sqrt(v)) - my) > 0:
            return Z
        t1 = F(my * my)
        if cfg.loss == 13:
            psi, zeta = F(F(1.0) + F(p2 * F(0.5))), F(F(1.0) + p2)
            rad = F(F(F(t1 * F(p2 * p2)) * F(0.25)) + F(F(v * p2) * zeta))
            alpha = F(F(np.sqrt(rad) - F(my * psi)) / F(v * zeta))
            alpha = min(cc, alpha)
        else:
            n = F(v + F(F(0.5) / cc))
            vp2 = F(v * p2)
            g1 = F(F(F(p2 * t1) * v) * v)
            g2 = F(F(F(F(4.0) * n) * v) * F(n + vp2))
            gamma = F(phi * np.sqrt(F(g1 + g2)))
            num = F(gamma - F(F(F(F(2.0) * my) * n) + F(F(p2 * my) * v)))
            den = F(F(2.0) * F(F(n * n) + F(F(n * v) * p2)))
            alpha = F(num / den)
        if not alpha > 0:
            return Z
        avp = F(F(alpha * v) * phi)
        usq = F(F(0.5) * F(np.sqrt(F(F(avp * avp) + F(F(4.0) * v))) - avp))
        return F(-F(alpha * y)), F(F(alpha * phi) / F(usq + avp))
    if cfg.loss == 9:
        l, d = F(F(1.0) - F(y * m)), y
        if not l > 0:
            return F(0.0), F(0.0)
    else:
        e = F(y - m)
        d = F(-1.0) if e < 0 else F(1.0)
        l = F(abs(e)) if cfg.loss == 10 else F(abs(e) - F(cfg.pa_epsilon))
        if cfg.loss == 11 and not l > 0:
            return F(0.0), F(0.0)
    beta = F(F(1.0) / F(v + F(cfg.cov_r)))
    alpha = F(l * beta)
    return F(-F(alpha * d)), beta


def np_train_cov(cfg, rowptr, feat, val, label):
    D, B = int(cfg.dims), int(cfg.mini_batch)
    w = np.zeros(D, F); sg = np.ones(D, F)
    n = len(label); t = 0
    for _ in range(cfg.iters):
        for b0 in range(0, n, B):
            b1 = min(n, b0 + B)
            cs = []
            for i in range(b0, b1):
                am, av = 0.0, 0.0
                for p in range(rowptr[i], rowptr[i + 1]):
                    am += float(F(w[feat[p]] * val[p]))
                    av += float(F(sg[feat[p]] * F(val[p] * val[p])))
                y = (F(1.0) if label[i] > 0 else F(-1.0)) if cfg.task == 0 else F(label[i])
                cs.append(np_cov_coefs(cfg, F(am), y, F(av)))
            s1, s2, cnts = {}, {}, {}
            for i in range(b0, b1):
                for p in range(rowptr[i], rowptr[i + 1]):
                    j = int(feat[p])
                    s1[j] = s1.get(j, 0.0) + float(F(cs[i - b0][0] * val[p]))
                    s2[j] = s2.get(j, 0.0) + float(F(cs[i - b0][1] * F(val[p] * val[p])))
                    cnts[j] = cnts.get(j, 0) + 1
            t += 1
            for j in s1:
                def mean(sm):
                    return F(sm * (1.0 / float(b1 - b0))) if cfg.grad_norm == 0 else (F(sm / float(cnts[j])) if cfg.grad_norm == 1 else F(sm))
                g1, g2, s = mean(s1[j]), mean(s2[j]), sg[j]
                w[j] = F(w[j] - F(s * g1))
                sg[j] = F(F(1.0) / F(F(F(1.0) / s) + g2)) if cfg.loss == 12 else F(s - F(F(s * s) * g2))
    return w, sg, np.zeros(D, F), t


def np_train(cfg, rowptr, feat, val, label):
    if 9 <= cfg.loss <= 14:
        return np_train_cov(cfg, rowptr, feat, val, label)
    D, B = int(cfg.dims), int(cfg.mini_batch)
    w = np.zeros(D, F); s1 = np.zeros(D, F); s2 = np.zeros(D, F)
    n = len(label); t = 0
    for _ in range(cfg.iters):
        for b0 in range(0, n, B):
            b1 = min(n, b0 + B)
            dl = []
            for i in range(b0, b1):
                acc, xx = 0.0, 0.0
                for p in range(rowptr[i], rowptr[i + 1]):
                    acc += float(F(w[feat[p]] * val[p]))
                    xx += float(F(val[p] * val[p]))
                y = (F(1.0) if label[i] > 0 else F(-1.0)) if cfg.task == 0 else F(label[i])
                dl.append(np_row_coef(cfg, F(acc), y, F(xx)))
            sums, cnts = {}, {}
            for i in range(b0, b1):
                for p in range(rowptr[i], rowptr[i + 1]):
                    j = int(feat[p])
                    sums[j] = sums.get(j, 0.0) + float(F(dl[i - b0] * val[p]))
                    cnts[j] = cnts.get(j, 0) + 1
            t += 1
            eta = F(_oracle.lib().hmo_eta(cfg, t))
            # Adam's bias-corrected step size, in fp64 on the host side of both implementations
            alpha_t = F(float(eta) * np.sqrt(1.0 - float(F(cfg.beta2)) ** t) / (1.0 - float(F(cfg.beta1)) ** t)) \
                if cfg.opt == 3 else F(0.0)
            for j, sm in sums.items():
                g = F(sm * (1.0 / float(b1 - b0))) if cfg.grad_norm == 0 else (F(sm / float(cnts[j])) if cfg.grad_norm == 1 else F(sm))
                wv = w[j]
                if cfg.reg == 2 and cfg.opt != 4:
                    g = F(g + F(F(cfg.lambda_) * wv))
                if cfg.opt == 0:
                    wv = F(wv - F(eta * g))
                elif cfg.opt == 1:
                    s1[j] = F(s1[j] + F(g * g))
                    wv = F(wv - F(F(eta * g) / F(np.sqrt(s1[j]) + F(cfg.eps))))
                elif cfg.opt == 4:
                    n1 = F(s2[j] + F(g * g))
                    sig = F(F(np.sqrt(n1) - np.sqrt(s2[j])) / F(cfg.alpha))
                    z = F(F(s1[j] + g) - F(sig * wv))
                    s1[j], s2[j] = z, n1
                    if abs(z) <= F(cfg.lambda1):
                        wv = F(0.0)
                    else:
                        sg = F(-1.0) if z < 0 else F(1.0)
                        wv = F(-F(F(z - F(sg * F(cfg.lambda1))) /
                                  F(F(F(F(cfg.beta) + np.sqrt(n1)) / F(cfg.alpha)) + F(cfg.lambda2))))
                elif cfg.opt == 2:      # AdaDelta: (s1, s2) = (E[g^2], E[dx^2])
                    rho, eps = F(cfg.rho), F(cfg.eps)
                    omr = F(F(1.0) - rho)
                    E = F(F(rho * s1[j]) + F(F(omr * g) * g))
                    dx = F(F(np.sqrt(F(s2[j] + eps)) / np.sqrt(F(E + eps))) * g)
                    s2[j] = F(F(rho * s2[j]) + F(F(omr * dx) * dx))
                    s1[j] = E
                    wv = F(wv - dx)
                elif cfg.opt == 3:      # Adam: (s1, s2) = (m, v)
                    be1, be2 = F(cfg.beta1), F(cfg.beta2)
                    m = F(F(be1 * s1[j]) + F(F(F(1.0) - be1) * g))
                    v = F(F(be2 * s2[j]) + F(F(F(F(1.0) - be2) * g) * g))
                    s1[j], s2[j] = m, v
                    wv = F(wv - F(F(alpha_t * m) / F(np.sqrt(v) + F(cfg.eps))))
                else:
                    raise NotImplementedError
                if cfg.reg == 1 and cfg.opt != 4:      # L1: soft threshold by eta * lambda after the step
                    a = F(abs(wv) - F(eta * F(cfg.lambda_)))
                    wv = F(np.copysign(a, wv)) if a > 0 else F(0.0)
                w[j] = wv
    return w, s1, s2, t


@pytest.mark.parametrize("options,fn", [
    ("-loss logloss -opt sgd -mini_batch 7 -dims 50 -eta0 0.5", "train_classifier"),
    ("-loss logloss -opt adagrad -mini_batch 16 -dims 50 -iters 2", "train_classifier"),
    ("-loss hinge -opt adagrad -mini_batch 5 -dims 50 -grad_norm count", "train_classifier"),
    ("-loss logloss -opt ftrl -mini_batch 11 -dims 50 -alpha 0.5 -lambda1 0.01 -lambda2 0.1", "train_classifier"),
    ("-loss squared -opt adagrad -mini_batch 9 -dims 50 -reg l2 -lambda 0.1 -eta inverse", "train_regressor"),
    ("-loss logloss -opt adadelta -mini_batch 8 -dims 50 -rho 0.9 -eps 1e-5 -iters 2", "train_classifier"),
    ("-loss squared -opt adadelta -mini_batch 13 -dims 50 -grad_norm sum", "train_regressor"),
    ("-loss logloss -opt adam -mini_batch 6 -dims 50 -eta0 0.05 -iters 3", "train_classifier"),
    ("-loss squared -opt adam -mini_batch 10 -dims 50 -eta0 0.01 -beta1 0.8 -beta2 0.99 -eta simple -total_steps 30", "train_regressor"),
    ("-loss hinge -opt sgd -mini_batch 7 -dims 50 -reg l1 -lambda 0.01 -eta0 0.3", "train_classifier"),
    # the further plain losses
    ("-loss squared_hinge -opt adagrad -mini_batch 7 -dims 50 -iters 2", "train_classifier"),
    ("-loss modified_huber -opt sgd -mini_batch 5 -dims 50 -eta0 0.3 -iters 2", "train_classifier"),
    ("-loss huber -huber_c 0.3 -opt adagrad -mini_batch 9 -dims 50", "train_regressor"),
    ("-loss quantile -quantile 0.8 -opt sgd -mini_batch 6 -dims 50 -eta0 0.05", "train_regressor"),
    ("-loss epsilon_insensitive -epsilon 0.2 -opt adam -mini_batch 8 -dims 50 -eta0 0.01", "train_regressor"),
    ("-loss squared_epsilon_insensitive -epsilon 0.05 -opt ftrl -mini_batch 4 -dims 50", "train_regressor"),
    # online learners on the same skeleton
    ("-mini_batch 1 -dims 50 -iters 2", "train_perceptron"),
    ("-mini_batch 6 -dims 50", "train_perceptron"),
    ("-mini_batch 1 -dims 50", "train_pa"),
    ("-mini_batch 5 -dims 50 -grad_norm count -iters 2", "train_pa"),
    ("-mini_batch 1 -dims 50 -c 0.05", "train_pa1"),
    ("-mini_batch 9 -dims 50 -aggressiveness 0.5 -grad_norm sum", "train_pa1"),
    ("-mini_batch 1 -dims 50 -c 0.3", "train_pa2"),
    ("-mini_batch 4 -dims 50 -c 2", "train_pa2"),
    ("-mini_batch 1 -dims 50 -c 0.1 -epsilon 0.05", "train_pa1_regr"),
    ("-mini_batch 8 -dims 50 -e 0.2 -iters 2", "train_pa1_regr"),
    ("-mini_batch 1 -dims 50 -c 0.7", "train_pa2_regr"),
    ("-mini_batch 7 -dims 50 -c 5 -epsilon 0.01 -grad_norm count", "train_pa2_regr"),
    # learners that keep a variance per feature
    ("-mini_batch 1 -dims 50 -iters 2", "train_arow"),
    ("-mini_batch 6 -dims 50 -r 1.0 -grad_norm count", "train_arow"),
    ("-mini_batch 1 -dims 50 -r 0.5", "train_arow_regr"),
    ("-mini_batch 8 -dims 50 -grad_norm sum -r 10", "train_arow_regr"),
    ("-mini_batch 1 -dims 50 -epsilon 0.05", "train_arowe_regr"),
    ("-mini_batch 5 -dims 50 -epsilon 0.2 -r 2 -iters 2", "train_arowe_regr"),
    ("-mini_batch 1 -dims 50 -iters 2", "train_cw"),
    ("-mini_batch 6 -dims 50 -phi 0.5 -grad_norm count", "train_cw"),
    ("-mini_batch 1 -dims 50 -phi 0.8 -c 0.5", "train_scw"),
    ("-mini_batch 7 -dims 50 -c 0.05 -iters 2", "train_scw"),
    ("-mini_batch 1 -dims 50 -phi 1.5 -c 2", "train_scw2"),
    ("-mini_batch 4 -dims 50 -c 0.3 -grad_norm sum", "train_scw2"),
])
def test_oracle_matches_numpy_restatement(options, fn):
    data = synth.ragged(120, dims=50, max_nnz=9, seed=2, regression=fn in ("train_regressor", "train_pa1_regr", "train_pa2_regr", "train_arow_regr", "train_arowe_regr"))
    rp, f, v, y = data.numpy()
    cfg = _oracle.config_from(hm.parse_options(fn, options))
    res = _oracle.train_linear(cfg, rp, f, v, y)
    w, s1, s2, t = np_train(cfg, rp, f, v, y)
    assert res.t == t
    assert np.array_equal(res.w, w), np.abs(res.w - w).max()
    assert np.array_equal(res.s1, s1) and np.array_equal(res.s2, s2)


@pytest.mark.parametrize("opt", ["sgd", "adagrad", "adadelta", "adam", "ftrl"])
def test_openmp_variant_is_bit_identical(opt):
    data = synth.criteo_like(3000, dims=1 << 14, seed=3)
    rp, f, v, y = data.numpy()
    cfg = _oracle.config_from(hm.parse_options("train_classifier", f"-opt {opt} -mini_batch 256 -dims {1 << 14} -iters 2"))
    a = _oracle.train_linear(cfg, rp, f, v, y)
    for threads in (1, 3, 4):
        b = _oracle.train_linear_mt(cfg, rp, f, v, y, threads)
        assert np.array_equal(a.w, b.w) and np.array_equal(a.s1, b.s1) and np.array_equal(a.s2, b.s2)
        assert a.t == b.t


def test_oracle_learns_and_model_table_is_sorted():
    data = synth.a9a_like(4096, seed=0)
    rp, f, v, y = data.numpy()
    cfg = _oracle.config_from(hm.parse_options("train_classifier", "-opt adagrad -mini_batch 128 -dims 124"))
    res = _oracle.train_linear(cfg, rp, f, v, y)
    import torch
    assert synth.auc(torch.from_numpy(_oracle.predict_linear(rp, f, v, res.w)), torch.from_numpy(y)) > 0.7
    feat, wt = _oracle.model_table(res.w)
    assert np.all(np.diff(feat) > 0) and np.all(wt != 0) and len(feat) == np.count_nonzero(res.w)


def test_oracle_rejects_out_of_range_feature():
    cfg = _oracle.config_from(hm.parse_options("train_classifier", "-dims 4 -mini_batch 2"))
    with pytest.raises(RuntimeError):
        _oracle.train_linear(cfg, np.array([0, 1]), np.array([4]), np.array([1.0]), np.array([1.0]))


def test_fm_oracle_runs_and_learns():
    data = synth.criteo_like(2000, dims=1 << 10, seed=5)
    rp, f, v, y = data.numpy()
    cfg = _oracle.config_from(hm.parse_options("train_fm", f"-factors 4 -classification -mini_batch 100 -dims {1 << 10} -iters 3 -opt adagrad -eta0 0.1"))
    r = _oracle.train_fm(cfg, rp, f, v, y)
    import torch
    p = _oracle.predict_fm(cfg, rp, f, v, r.w0[0], r.w, r.V)
    assert np.isfinite(p).all() and synth.auc(torch.from_numpy(p), torch.from_numpy(y)) > 0.6
    assert r.touched.sum() == len(np.unique(f))


def test_synth_is_deterministic_and_shaped():
    a = synth.criteo_like(1000, dims=1 << 24, seed=1)
    b = synth.criteo_like(500, dims=1 << 24, seed=1, row0=500)
    assert a.nnz == 39000 and int(a.feature.max()) < (1 << 24) and int(a.feature.min()) >= 0
    assert np.array_equal(a.feature.numpy()[500 * 39:], b.feature.numpy())      # streams are position-addressed
    assert np.array_equal(a.label.numpy()[500:], b.label.numpy())
    k = synth.a9a_like(100)
    assert k.nnz == 1400 and int(k.feature.max()) < 123


# ---- independent numpy restatement of the factorization machine ----

def np_fmix32(h):
    h &= 0xFFFFFFFF
    h ^= h >> 16; h = (h * 0x85EBCA6B) & 0xFFFFFFFF
    h ^= h >> 13; h = (h * 0xC2B2AE35) & 0xFFFFFFFF
    h ^= h >> 16
    return h


def np_fm_init(cfg):
    n = int(cfg.dims) * int(cfg.factors)
    V = np.empty(n, F)
    for q in range(n):
        h = np_fmix32((q * 0x9E3779B1 + int(cfg.seed)) & 0xFFFFFFFF)
        u = F(F(F(F(h >> 8) * F(5.9604644775390625e-08)) - F(0.5)) * F(2.0))
        V[q] = F(F(cfg.sigma) * u)
    return V


def np_fm_step(cfg, eta, g, lam, th, acc):
    g = F(g + F(F(lam) * th))
    if cfg.opt == 1:
        G = F(acc + F(g * g))
        return F(th - F(F(eta * g) / F(np.sqrt(G) + F(cfg.eps)))), G
    return F(th - F(eta * g)), acc


def np_train_fm(cfg, rowptr, feat, val, label):
    D, B, k = int(cfg.dims), int(cfg.mini_batch), int(cfg.factors)
    w0, gw0 = F(0.0), F(0.0)
    w = np.zeros(D, F); gw = np.zeros(D, F)
    V = np_fm_init(cfg).reshape(D, k); gv = np.zeros((D, k), F)
    n = len(label); t = 0
    preds = np.zeros(n, F)
    for _ in range(cfg.iters):
        for b0 in range(0, n, B):
            b1 = min(n, b0 + B)
            dl, S = [], []
            g0 = 0.0
            for i in range(b0, b1):
                ps = range(rowptr[i], rowptr[i + 1])
                lin = 0.0
                for p in ps:
                    lin += float(F(w[feat[p]] * val[p]))
                inter = 0.0
                s_row = []
                for f in range(k):
                    Sf, Qf = 0.0, 0.0
                    for p in ps:
                        vx = F(V[feat[p], f] * val[p])
                        Sf += float(vx); Qf += float(F(vx * vx))
                    s32 = F(Sf)
                    s_row.append(s32)
                    inter += float(F(s32 * s32)) - Qf
                pred = F(float(w0) + lin + 0.5 * inter)
                preds[i] = pred
                if cfg.fm_classification:
                    y = F(1.0) if label[i] > 0 else F(-1.0)
                    d = np_dloss(0, pred, y)
                else:
                    d = F(pred - label[i])
                dl.append(d); S.append(s_row); g0 += float(d)
            A1, A2, GS, cnts, order = {}, {}, {}, {}, []
            for i in range(b0, b1):
                for p in range(rowptr[i], rowptr[i + 1]):
                    j = int(feat[p])
                    if j not in cnts:
                        cnts[j] = 0; A1[j] = 0.0; A2[j] = 0.0; GS[j] = [0.0] * k; order.append(j)
                    cnts[j] += 1
                    c1 = F(dl[i - b0] * val[p])
                    A1[j] += float(c1)
                    A2[j] += float(F(c1 * val[p]))
                    for f in range(k):
                        GS[j][f] += float(F(c1 * S[i - b0][f]))
            t += 1
            eta = F(_oracle.lib().hmo_eta(cfg, t))
            inv = 1.0 / float(b1 - b0)

            def mean(sm, j):
                return F(sm * inv) if cfg.grad_norm == 0 else (F(sm / float(cnts[j])) if cfg.grad_norm == 1 else F(sm))
            w0, gw0 = np_fm_step(cfg, eta, F(g0) if cfg.grad_norm == 2 else F(g0 * inv), cfg.lambda_w0, w0, gw0)
            for j in order:
                w[j], gw[j] = np_fm_step(cfg, eta, mean(A1[j], j), cfg.lambda_w, w[j], gw[j])
                for f in range(k):
                    G = GS[j][f] - float(V[j, f]) * A2[j]
                    V[j, f], gv[j, f] = np_fm_step(cfg, eta, mean(G, j), cfg.lambda_v, V[j, f], gv[j, f])
    return w0, w, V.reshape(-1), gw0, gw, gv.reshape(-1), preds, t


@pytest.mark.parametrize("options", [
    "-factors 3 -classification -opt sgd -mini_batch 7 -dims 40 -eta0 0.1",
    "-factors 4 -classification -opt adagrad -mini_batch 16 -dims 40 -eta0 0.1 -iters 2",
    "-factors 2 -opt sgd -mini_batch 5 -dims 40 -eta0 0.02 -lambda0 0.1 -lambdaW 0.05 -lambdaV 0.2",
    "-factors 5 -opt adagrad -mini_batch 9 -dims 40 -eta0 0.05 -grad_norm count -sigma 0.3 -seed 7",
    "-factors 1 -classification -opt adagrad -mini_batch 1 -dims 40 -grad_norm sum",
])
def test_fm_oracle_matches_numpy_restatement(options):
    regression = "-classification" not in options
    data = synth.ragged(90, dims=40, max_nnz=7, seed=4, regression=regression)
    rp, f, v, y = data.numpy()
    cfg = _oracle.config_from(hm.parse_options("train_fm", options))
    assert np.array_equal(_oracle.fm_init_v(cfg), np_fm_init(cfg))
    res = _oracle.train_fm(cfg, rp, f, v, y)
    w0, w, V, gw0, gw, gvv, preds, t = np_train_fm(cfg, rp, f, v, y)
    assert res.t == t
    assert np.array_equal(res.preds, preds)
    assert res.w0[0] == w0 and np.array_equal(res.w, w) and np.array_equal(res.V, V)
    if cfg.opt == 1:
        assert res.gw0[0] == gw0 and np.array_equal(res.gw, gw) and np.array_equal(res.gv, gvv)
