"""Synthetic CSR inputs of the shapes BASELINE.json's configs name (there is no network for
real datasets).  Pure torch integer/float ops so the same code fills host arrays for the
parity tests and device arrays for bench.py; deterministic for a given (seed, device type
is irrelevant: all randomness comes from a counter hash, not from torch's generators).

  a9a_like     123 one-hot features in 14 fields, 14 nnz/row, values 1.0        (configs[0])
  criteo_like  13 numeric (fixed indices, positive values) + 26 categorical
               (power-law categories hashed into [0, dims)), 39 nnz/row          (configs[1..2])
  kdd12_like   11 categorical fields, 11 nnz/row, values 1.0                    (configs[4])

Labels are Bernoulli draws from a planted sparse linear model so AUC / logloss are
meaningful.  Row order is the generation order; nothing is shuffled.
"""
from __future__ import annotations

import math
from dataclasses import dataclass

import torch

_M32 = 0xFFFFFFFF


def _mix(x: torch.Tensor) -> torch.Tensor:
    """murmur3 fmix32 on int64 tensors holding uint32 values."""
    x = x & _M32
    x = x ^ (x >> 16)
    x = (x * 0x85EBCA6B) & _M32
    x = x ^ (x >> 13)
    x = (x * 0xC2B2AE35) & _M32
    x = x ^ (x >> 16)
    return x


def _u01(counter: torch.Tensor, stream: int, seed: int) -> torch.Tensor:
    """Uniform [0,1) float64 from a counter hash (24 bits)."""
    h = _mix(counter * 0x9E3779B1 + (stream * 0x85EBCA77 + seed * 0xC2B2AE3D + 0x27D4EB2F))
    h = _mix(h + stream)
    return (h >> 8).to(torch.float64) * (1.0 / (1 << 24))


@dataclass
class Csr:
    rowptr: torch.Tensor   # int64 [rows+1]
    feature: torch.Tensor  # int32 [nnz]
    value: torch.Tensor    # float32 [nnz]
    label: torch.Tensor    # float32 [rows]  (0/1 for classification, real for regression)

    @property
    def rows(self) -> int:
        return self.label.numel()

    @property
    def nnz(self) -> int:
        return self.feature.numel()

    def numpy(self):
        return (self.rowptr.cpu().numpy(), self.feature.cpu().numpy(), self.value.cpu().numpy(),
                self.label.cpu().numpy())


def _planted_weight(idx: torch.Tensor, seed: int, scale: float) -> torch.Tensor:
    h = _mix(idx.to(torch.int64) * 0x2545F491 + seed * 7919 + 13)
    # one feature in four carries signal
    w = ((h & 0xFFFF).to(torch.float64) / 65536.0 - 0.5) * 2.0 * scale
    return torch.where(((h >> 16) & 3) == 0, w, torch.zeros_like(w))


def _finish(idx: torch.Tensor, val: torch.Tensor, n_rows: int, per_row: int, row0: int, seed: int,
            regression: bool, signal: float) -> Csr:
    dev = idx.device
    m = (_planted_weight(idx, seed, signal) * val.to(torch.float64)).view(n_rows, per_row).sum(dim=1)
    rows = torch.arange(row0, row0 + n_rows, device=dev, dtype=torch.int64)
    u = _u01(rows, 0xAB, seed)
    if regression:
        noise = (_u01(rows, 0xAC, seed) - 0.5) * 0.2
        label = (m + noise).to(torch.float32)
    else:
        label = (u < torch.sigmoid(m)).to(torch.float32)
    rowptr = torch.arange(0, n_rows + 1, device=dev, dtype=torch.int64) * per_row
    return Csr(rowptr, idx.to(torch.int32).reshape(-1), val.to(torch.float32).reshape(-1), label)


def criteo_like(n_rows: int, dims: int = 1 << 24, seed: int = 1, device="cpu", row0: int = 0,
                regression: bool = False) -> Csr:
    """39 nnz/row: 13 numeric features at fixed hashed indices with value log1p(x) of a
    heavy-tailed x, 26 categorical features whose category follows a power law over a
    per-field cardinality, hashed into [0, dims)."""
    dev = torch.device(device)
    rows = torch.arange(row0, row0 + n_rows, device=dev, dtype=torch.int64)
    fields = torch.arange(39, device=dev, dtype=torch.int64)
    ctr = rows[:, None] * 64 + fields[None, :]                  # unique counter per element
    u = _u01(ctr, 1, seed)
    # numeric: fixed index per field, value = log1p(pareto-ish)
    num_idx = _mix(fields[:13] * 0x632BE5AB + 0x1234567 + seed) % dims
    num_val = torch.log1p(1.0 / (1.0 - u[:, :13] * 0.999) - 1.0)
    # categorical: cardinality grows with the field id (10 .. ~10M), category ~ u^3 * card
    card = (10.0 * (10 ** (torch.arange(26, device=dev, dtype=torch.float64) * (6.0 / 25.0)))).to(torch.int64)
    cat = (u[:, 13:] ** 3 * card[None, :].to(torch.float64)).to(torch.int64)
    cat_idx = _mix(cat * 0x9E3779B1 + (fields[13:] * 0x7FEB352D)[None, :] + seed) % dims
    idx = torch.cat([num_idx[None, :].expand(n_rows, 13), cat_idx], dim=1)
    val = torch.cat([num_val, torch.ones(n_rows, 26, device=dev, dtype=torch.float64)], dim=1)
    return _finish(idx, val, n_rows, 39, row0, seed, regression, 0.5)


def a9a_like(n_rows: int = 32768, seed: int = 0, device="cpu", row0: int = 0) -> Csr:
    """123 binary features in 14 one-hot fields (cardinalities sum to 123), 14 nnz/row."""
    dev = torch.device(device)
    card = torch.tensor([5, 8, 16, 7, 14, 6, 5, 2, 41, 2, 3, 4, 5, 5], device=dev, dtype=torch.int64)
    assert int(card.sum()) == 123
    off = torch.cumsum(card, 0) - card
    rows = torch.arange(row0, row0 + n_rows, device=dev, dtype=torch.int64)
    fields = torch.arange(14, device=dev, dtype=torch.int64)
    u = _u01(rows[:, None] * 16 + fields[None, :], 2, seed)
    idx = off[None, :] + (u ** 2 * card[None, :].to(torch.float64)).to(torch.int64)
    val = torch.ones(n_rows, 14, device=dev, dtype=torch.float64)
    return _finish(idx, val, n_rows, 14, row0, seed, False, 1.5)


def kdd12_like(n_rows: int, dims: int = 1 << 24, seed: int = 4, device="cpu", row0: int = 0) -> Csr:
    """11 categorical fields, value 1.0, power-law categories hashed into [0, dims)."""
    dev = torch.device(device)
    rows = torch.arange(row0, row0 + n_rows, device=dev, dtype=torch.int64)
    fields = torch.arange(11, device=dev, dtype=torch.int64)
    u = _u01(rows[:, None] * 16 + fields[None, :], 3, seed)
    card = (100.0 * (10 ** (fields.to(torch.float64) * 0.5))).to(torch.int64)
    cat = (u ** 3 * card[None, :].to(torch.float64)).to(torch.int64)
    idx = _mix(cat * 0x9E3779B1 + (fields * 0x7FEB352D)[None, :] + seed) % dims
    val = torch.ones(n_rows, 11, device=dev, dtype=torch.float64)
    return _finish(idx, val, n_rows, 11, row0, seed, False, 0.8)


def ragged(n_rows: int, dims: int, max_nnz: int, seed: int = 9, device="cpu", regression: bool = False) -> Csr:
    """Rows of uneven length (0..max_nnz), duplicate features allowed inside a row,
    signed real values: the edge-case generator for the parity tests."""
    dev = torch.device(device)
    rows = torch.arange(n_rows, device=dev, dtype=torch.int64)
    lens = (_u01(rows, 5, seed) ** 2 * (max_nnz + 1)).to(torch.int64).clamp_(0, max_nnz)
    rowptr = torch.zeros(n_rows + 1, device=dev, dtype=torch.int64)
    rowptr[1:] = torch.cumsum(lens, 0)
    nnz = int(rowptr[-1])
    e = torch.arange(nnz, device=dev, dtype=torch.int64)
    idx = (_u01(e, 6, seed) ** 2 * dims).to(torch.int64).clamp_(0, dims - 1)
    val = ((_u01(e, 7, seed) - 0.5) * 4.0)
    row_of = torch.repeat_interleave(rows, lens)
    m = torch.zeros(n_rows, device=dev, dtype=torch.float64).index_add_(
        0, row_of, _planted_weight(idx, seed, 1.0) * val)
    if regression:
        label = m.to(torch.float32)
    else:
        label = (_u01(rows, 8, seed) < torch.sigmoid(m)).to(torch.float32)
    return Csr(rowptr, idx.to(torch.int32), val.to(torch.float32), label)


def auc(score: torch.Tensor, label: torch.Tensor) -> float:
    """Area under the ROC curve by rank statistic (ties broken by order)."""
    s = score.double().flatten()
    y = (label.flatten() > 0).double()
    order = torch.argsort(s)
    ranks = torch.empty_like(s)
    ranks[order] = torch.arange(1, s.numel() + 1, dtype=torch.float64, device=s.device)
    npos = float(y.sum())
    nneg = s.numel() - npos
    if npos == 0 or nneg == 0:
        return float("nan")
    return float(((ranks * y).sum() - npos * (npos + 1) / 2) / (npos * nneg))


def logloss(margin: torch.Tensor, label: torch.Tensor) -> float:
    m = margin.double().flatten()
    y = (label.flatten() > 0).double() * 2 - 1
    return float(torch.nn.functional.softplus(-y * m).mean())


__all__ = ["Csr", "criteo_like", "a9a_like", "kdd12_like", "ragged", "auc", "logloss", "math"]
