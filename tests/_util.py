"""Shared helpers of the parity tests."""
from __future__ import annotations

import numpy as np

import hivemall_b200 as hm
from hivemall_b200 import synth

import _oracle

# BASELINE.json north_star: "weights ... within a stated relative tolerance (1e-5 fp32)".
# The absolute floor covers weights that are (near) zero on one side.
W_RTOL = 1e-5
W_ATOL = 1e-7


def assert_weights_close(got: np.ndarray, want: np.ndarray, what: str = "w", rtol=W_RTOL, atol=W_ATOL):
    got = np.asarray(got, dtype=np.float64)
    want = np.asarray(want, dtype=np.float64)
    assert got.shape == want.shape, f"{what}: shape {got.shape} vs {want.shape}"
    # a diverged run (inf / nan) must diverge in the same places, with the same sign
    fin = np.isfinite(got) & np.isfinite(want)
    same_nonfinite = (np.isnan(got) & np.isnan(want)) | ((got == want) & ~fin)
    with np.errstate(invalid="ignore"):
        err = np.where(fin, np.abs(got - want), np.where(same_nonfinite, 0.0, np.inf))
        tol = np.where(fin, rtol * np.maximum(np.abs(got), np.abs(want)) + atol, 0.0)
    bad = err > tol
    if bad.any():
        i = int(np.argmax(err - tol))
        raise AssertionError(
            f"{what}: {int(bad.sum())} of {got.size} entries differ beyond rtol={rtol} atol={atol}; "
            f"worst at {i}: got {got[i]!r} want {want[i]!r}")


def run_gpu_linear(function_name: str, options: str, data: synth.Csr, pushes=None, device: int = 0):
    """Train through the C-ABI from host buffers; returns (udtf, w, state)."""
    rowptr, feature, value, label = data.numpy()
    u = hm.UDTF(options, device, function_name=function_name)
    n = len(label)
    if pushes is None:
        pushes = [n]
    r0 = 0
    for cnt in pushes:
        r1 = min(n, r0 + cnt)
        u.process_batch(rowptr[r0:r1 + 1], feature, value, label[r0:r1])
        r0 = r1
    assert r0 == n
    u.close()
    w = u.array_to_host("w")
    state = u.array_to_host("state")
    return u, w, state


def run_oracle_linear(function_name: str, options: str, data: synth.Csr):
    cfg = hm.parse_options(function_name, options)
    ocfg = _oracle.config_from(cfg)
    rowptr, feature, value, label = data.numpy()
    res = _oracle.train_linear(ocfg, rowptr, feature, value, label)
    width = 0 if cfg.opt == 0 else (1 if cfg.opt in (1, 5) else 2)
    return res, res.state(width)
