"""Host-side mirror of the UDTF interface of the hot path, over the C-ABI.

BASELINE.json's north_star fixes the surface: `train_classifier`, `train_regressor`,
`train_fm` UDTFs with option strings, the lifecycle initialize -> process(row)* ->
close -> forward(model rows), plus the `mhash` / `feature_hashing` UDFs.  The reference
mount has no source (/root/reference/README.md:20-22), so names follow north_star and
argument meaning follows include/hivemall_b200.h.  The reference host language (Java)
has no toolchain in this image; INTEGRATION.md shows the JNI stub over the same C-ABI.

Everything computes on the GPU through libhivemall_b200.so.  numpy is used for host
buffers only; torch is optional and only wraps device memory.
"""
from __future__ import annotations

import ctypes as C
from typing import Iterable, Sequence

import numpy as np

from . import _lib
from ._lib import HivemallError, HmConfig, HmStats, UDFArgumentError, check  # noqa: F401

DEFAULT_NUM_FEATURES = 1 << 24
DEFAULT_HASH_SEED = 0x9747B28C


def _ptr(a) -> C.c_void_p:
    if a is None:
        return C.c_void_p(0)
    if isinstance(a, np.ndarray):
        return C.c_void_p(a.ctypes.data)
    if isinstance(a, int):
        return C.c_void_p(a)
    if hasattr(a, "data_ptr"):  # torch tensor
        return C.c_void_p(a.data_ptr())
    raise TypeError(f"cannot take the address of {type(a)!r}")


def _pack_strings(strings: Sequence[str | bytes]):
    enc = [s.encode("utf-8") if isinstance(s, str) else bytes(s) for s in strings]
    offsets = np.zeros(len(enc) + 1, dtype=np.int64)
    if enc:
        offsets[1:] = np.cumsum([len(b) for b in enc])
    data = np.frombuffer(b"".join(enc), dtype=np.uint8) if enc else np.zeros(0, dtype=np.uint8)
    return np.ascontiguousarray(data), offsets


def parse_options(function_name: str, options: str = "") -> HmConfig:
    cfg = HmConfig()
    check(_lib.load().hm_parse_options(function_name.encode(), options.encode(), C.byref(cfg)))
    return cfg


def mhash(strings: Sequence[str | bytes], num_features: int = DEFAULT_NUM_FEATURES,
          seed: int = DEFAULT_HASH_SEED, device: int = 0) -> np.ndarray:
    """mhash UDF over a batch of feature names: MurmurHash3_x86_32 folded into
    [0, num_features) on the GPU.  Returns int32 indices."""
    data, offsets = _pack_strings(strings)
    out = np.empty(len(strings), dtype=np.int32)
    check(_lib.load().hm_mhash(_ptr(data), _ptr(offsets), len(strings), seed & 0xFFFFFFFF, num_features,
                               _ptr(out), device))
    return out


def feature_hashing(features: Sequence[str | bytes], num_features: int = DEFAULT_NUM_FEATURES,
                    seed: int = DEFAULT_HASH_SEED, device: int = 0):
    """feature_hashing UDF over "name:value" strings -> (int32 index, float32 value)."""
    data, offsets = _pack_strings(features)
    idx = np.empty(len(features), dtype=np.int32)
    val = np.empty(len(features), dtype=np.float32)
    check(_lib.load().hm_feature_hashing(_ptr(data), _ptr(offsets), len(features), seed & 0xFFFFFFFF,
                                         num_features, _ptr(idx), _ptr(val), device))
    return idx, val


BIAS_FEATURE = 0          # the bias clause "0:1.0": feature index 0 (a name made of digits is an index already)
BIAS_CLAUSE = "0:1.0"


def add_bias(features: Sequence[str | bytes]) -> list:
    """add_bias UDF: one row's feature strings with the bias clause "0:1.0" appended.  Row-level host logic
    that stays upstream of the UDTF boundary: the rows reach process() with the bias in
    them.  A row that already carries feature 0 gets a second one, as any repeated feature would."""
    out = list(features)
    out.append(BIAS_CLAUSE.encode() if out and isinstance(out[0], bytes) else BIAS_CLAUSE)
    return out


def add_bias_csr(rowptr, feature, value, bias_feature: int = BIAS_FEATURE, bias_value: float = 1.0):
    """add_bias for rows that are already (index, value) pairs in CSR form: every row gets one more element
    (bias_feature, bias_value) at its end.  value=None (index-only rows) stays None when bias_value is 1.
    Returns (rowptr, feature, value)."""
    rowptr = np.ascontiguousarray(rowptr, dtype=np.int64)
    feature = np.ascontiguousarray(feature)
    n = len(rowptr) - 1
    base = int(rowptr[0])
    lens = np.diff(rowptr)
    new_rp = np.zeros(n + 1, dtype=np.int64)
    new_rp[1:] = np.cumsum(lens + 1)
    # position of old element p of row i: p - base + i
    shift = np.repeat(np.arange(n, dtype=np.int64), lens)
    src = np.arange(base, base + int(lens.sum()), dtype=np.int64)
    new_f = np.empty(int(new_rp[-1]), dtype=feature.dtype)
    new_f[src - base + shift] = feature[src]
    new_f[new_rp[1:] - 1] = bias_feature
    if value is None and bias_value == 1.0:
        return new_rp, new_f, None
    new_v = np.empty(int(new_rp[-1]), dtype=np.float32)
    new_v[src - base + shift] = 1.0 if value is None else np.asarray(value, dtype=np.float32)[src]
    new_v[new_rp[1:] - 1] = bias_value
    return new_rp, new_f, new_v


def host_alloc(nbytes: int) -> tuple[int, np.ndarray]:
    """Pinned host buffer (address, uint8 view).  Free with host_free(address)."""
    p = C.c_void_p()
    check(_lib.load().hm_host_alloc(C.byref(p), nbytes))
    buf = (C.c_uint8 * max(nbytes, 1)).from_address(p.value)
    return p.value, np.frombuffer(buf, dtype=np.uint8, count=nbytes)


def host_free(address: int) -> None:
    check(_lib.load().hm_host_free(C.c_void_p(address)))


def bind_host_to_gpu(device: int) -> list[int] | None:
    """Pin the calling process to the CPUs NVML names as closest to GPU `device`, so that host buffers
    allocated afterwards (host_alloc: pinned, placed by first touch) land in that GPU's NUMA node and the
    thread that feeds the GPU runs next to it.  With one process per GPU and eight GPUs copying at once, buffers
    that all sit in one socket's memory cap the end-to-end rate.  Returns the CPU list, or None when NVML or
    the affinity call is unavailable (nothing is changed then)."""
    import os
    try:
        import pynvml
        import torch
        pynvml.nvmlInit()
        try:
            uuid = str(torch.cuda.get_device_properties(device).uuid)
            handle = pynvml.nvmlDeviceGetHandleByUUID(uuid if uuid.startswith("GPU-") else "GPU-" + uuid)
        except Exception:
            handle = pynvml.nvmlDeviceGetHandleByIndex(device)
        words = ((os.cpu_count() or 64) + 63) // 64
        mask = pynvml.nvmlDeviceGetCpuAffinity(handle, words)
        cpus = {i * 64 + b for i, w in enumerate(mask) for b in range(64) if (int(w) >> b) & 1}
        cpus = sorted(cpus & os.sched_getaffinity(0))
        if not cpus:
            return None
        os.sched_setaffinity(0, cpus)
        return cpus
    except Exception:
        return None


NOALLOC_MAX_DIMS = 1 << 26


def model_table_noalloc(d_w, dims: int, d_feature, d_weight, cap: int, d_n_rows, stream: int = 0) -> None:
    """Model table of a dense device weight array without any device allocation inside the
    library (for callers that must not call cudaMalloc: see hm_model_table_noalloc in the
    header).  Every argument is device memory the caller owns (torch CUDA tensors or raw
    addresses); the row count lands in d_n_rows[0] on the device.  Asynchronous on `stream`
    of the current device.  dims is limited to NOALLOC_MAX_DIMS; calls must not overlap."""
    check(_lib.load().hm_model_table_noalloc(_ptr(d_w), dims, _ptr(d_feature), _ptr(d_weight), cap,
                                             _ptr(d_n_rows), C.c_void_p(stream)))


class _DeviceArray:
    """Zero-copy view of a replica array for consumers of __cuda_array_interface__
    (torch.as_tensor, cupy.asarray)."""

    def __init__(self, ptr: int, n: int, owner):
        self._owner = owner
        self.__cuda_array_interface__ = {
            "shape": (n,), "typestr": "<f4", "data": (ptr, False), "version": 3, "strides": None}


class UDTF:
    """One UDTF instance = one model replica on one GPU."""

    function_name = ""

    def __init__(self, options: str = "", device: int = 0, function_name: str | None = None):
        self._lib = _lib.load()
        self._h = C.c_void_p()
        self.function_name = function_name or self.function_name
        self.options = options
        self.device = device
        self.config = parse_options(self.function_name, options)
        check(self._lib.hm_udtf_create(self.function_name.encode(), options.encode(), device, C.byref(self._h)))
        self._rows: list = []
        self._labels: list[float] = []
        self._row_flush = 4096
        self._keep = []  # host buffers that must outlive asynchronous copies

    # ---- lifecycle ---------------------------------------------------------------
    def process(self, features: Iterable, label: float) -> None:
        """process() of one row.  `features` holds "name:value" strings or (index, value)
        pairs.  Rows are buffered on the host and pushed in small batches; mini-batch
        boundaries are those of the row stream, not of the pushes."""
        self._rows.append(list(features))
        self._labels.append(float(label))
        if len(self._rows) >= self._row_flush:
            self._flush_rows()

    def _flush_rows(self) -> None:
        if not self._rows:
            return
        rows, labels = self._rows, np.asarray(self._labels, dtype=np.float32)
        self._rows, self._labels = [], []
        rowptr = np.zeros(len(rows) + 1, dtype=np.int64)
        rowptr[1:] = np.cumsum([len(r) for r in rows])
        flat = [f for r in rows for f in r]
        if flat and isinstance(flat[0], (str, bytes)):
            data, offsets = _pack_strings(flat)
            self.process_strings(rowptr, data, offsets, labels)
            self.sync()
        else:
            idx = np.asarray([f[0] for f in flat], dtype=np.int32)
            val = np.asarray([f[1] for f in flat], dtype=np.float32)
            self.process_batch(rowptr, idx, val, labels)
            self.sync()

    def process_batch(self, rowptr, feature, value, label) -> None:
        """process() of many rows at once from host CSR arrays (int64 rowptr, int32
        feature, float32 value, float32 label).  `value=None` means every value is 1.0
        (index-only features).  Asynchronous: keep the arrays alive and unmodified until
        sync()/close()."""
        rowptr = np.ascontiguousarray(rowptr, dtype=np.int64)
        feature = np.ascontiguousarray(feature, dtype=np.int32)
        value = None if value is None else np.ascontiguousarray(value, dtype=np.float32)
        label = np.ascontiguousarray(label, dtype=np.float32)
        n = len(rowptr) - 1
        if len(label) != n:
            raise UDFArgumentError(1, f"label has {len(label)} entries for {n} rows")
        if n and (rowptr[-1] > len(feature) or (value is not None and rowptr[-1] > len(value))):
            raise UDFArgumentError(1, "rowptr points past the end of feature/value")
        self._keep.append((rowptr, feature, value, label))
        check(self._lib.hm_udtf_process(self._h, n, _ptr(rowptr), _ptr(feature), _ptr(value), _ptr(label)))

    def process_pointers(self, n_rows: int, rowptr: int, feature: int, value: int | None, label: int) -> None:
        """process() from raw HOST addresses (pinned buffers from host_alloc); value None/0 = all 1.0."""
        check(self._lib.hm_udtf_process(self._h, n_rows, _ptr(rowptr), _ptr(feature), _ptr(value), _ptr(label)))

    def process_split(self, q_rowptr, q_feature, q_value, c_rowptr, c_feature, label) -> None:
        """process() of many rows whose features come in two groups: quantitative (index, value) and
        index-only (value 1.0).  Row i = its quantitative elements, then its index-only ones.  numpy arrays
        (kept alive until sync()/close()) or raw host addresses with `label` giving the row count via
        process_split_pointers."""
        q_rowptr = np.ascontiguousarray(q_rowptr, dtype=np.int64)
        c_rowptr = np.ascontiguousarray(c_rowptr, dtype=np.int64)
        q_feature = np.ascontiguousarray(q_feature, dtype=np.int32)
        q_value = np.ascontiguousarray(q_value, dtype=np.float32)
        c_feature = np.ascontiguousarray(c_feature, dtype=np.int32)
        label = np.ascontiguousarray(label, dtype=np.float32)
        n = len(label)
        if len(q_rowptr) != n + 1 or len(c_rowptr) != n + 1:
            raise UDFArgumentError(1, f"row pointers must have {n + 1} entries for {n} rows")
        if n and (q_rowptr[-1] > len(q_feature) or q_rowptr[-1] > len(q_value) or c_rowptr[-1] > len(c_feature)):
            raise UDFArgumentError(1, "a row pointer points past the end of its feature/value array")
        self._keep.append((q_rowptr, q_feature, q_value, c_rowptr, c_feature, label))
        check(self._lib.hm_udtf_process_split(self._h, n, _ptr(q_rowptr), _ptr(q_feature), _ptr(q_value),
                                              _ptr(c_rowptr), _ptr(c_feature), _ptr(label)))

    def process_split_pointers(self, n_rows: int, q_rowptr: int, q_feature: int, q_value: int, c_rowptr: int,
                               c_feature: int, label: int) -> None:
        """process_split() from raw HOST addresses (pinned buffers from host_alloc)."""
        check(self._lib.hm_udtf_process_split(self._h, n_rows, _ptr(q_rowptr), _ptr(q_feature), _ptr(q_value),
                                              _ptr(c_rowptr), _ptr(c_feature), _ptr(label)))

    def process_columns(self, q_feature, q_value, c_feature, label) -> None:
        """process() of many rows = a quantitative block whose n_q indices are the same in every row (`q_feature`,
        given once; `q_value` is the (rows, n_q) matrix) followed by n_c index-only features per row (`c_feature`,
        (rows, n_c)).  Asynchronous like process_batch."""
        q_feature = np.ascontiguousarray(q_feature, dtype=np.int32).reshape(-1)
        label = np.ascontiguousarray(label, dtype=np.float32)
        n = len(label)
        q_value = np.ascontiguousarray(q_value, dtype=np.float32).reshape(n, len(q_feature))
        c_feature = np.ascontiguousarray(c_feature, dtype=np.int32).reshape(n, -1)
        self._keep.append((q_feature, q_value, c_feature, label))
        check(self._lib.hm_udtf_process_columns(self._h, n, len(q_feature), _ptr(q_feature), _ptr(q_value),
                                                c_feature.shape[1], _ptr(c_feature), _ptr(label)))

    def process_columns_pointers(self, n_rows: int, n_q: int, q_feature: int, q_value: int, n_c: int, c_feature: int,
                                 label: int) -> None:
        """process_columns() from raw HOST addresses (pinned buffers from host_alloc)."""
        check(self._lib.hm_udtf_process_columns(self._h, n_rows, n_q, _ptr(q_feature), _ptr(q_value), n_c,
                                                _ptr(c_feature), _ptr(label)))

    def process_strings(self, rowptr, data, offsets, label) -> None:
        """process() of many rows whose features are still strings ("name:value" / "name"): int64 rowptr
        over strings, uint8 `data` holding the strings back to back, int64 offsets (len = strings + 1).
        The strings are hashed on the GPU straight into the learner's CSR.  Asynchronous like process_batch."""
        rowptr = np.ascontiguousarray(rowptr, dtype=np.int64)
        data = np.ascontiguousarray(data, dtype=np.uint8)
        offsets = np.ascontiguousarray(offsets, dtype=np.int64)
        label = np.ascontiguousarray(label, dtype=np.float32)
        n = len(rowptr) - 1
        if len(label) != n:
            raise UDFArgumentError(1, f"label has {len(label)} entries for {n} rows")
        if n and (rowptr[-1] >= len(offsets) or offsets[rowptr[-1]] > len(data)):
            raise UDFArgumentError(1, "rowptr / offsets point past the end of the strings")
        self._keep.append((rowptr, data, offsets, label))
        check(self._lib.hm_udtf_process_strings(self._h, n, _ptr(rowptr), _ptr(data), _ptr(offsets), _ptr(label)))

    def process_strings_pointers(self, n_rows: int, rowptr: int, data: int, offsets: int, label: int) -> None:
        """process_strings() from raw HOST addresses (pinned buffers from host_alloc)."""
        check(self._lib.hm_udtf_process_strings(self._h, n_rows, _ptr(rowptr), _ptr(data), _ptr(offsets), _ptr(label)))

    def process_device(self, n_rows: int, nnz: int, rowptr, feature, value, label) -> None:
        """process() of rows already in device memory (torch CUDA tensors or raw
        addresses): int64 rowptr, int32 feature, float32 value/label."""
        check(self._lib.hm_udtf_process_device(self._h, n_rows, nnz, _ptr(rowptr), _ptr(feature), _ptr(value),
                                               _ptr(label)))

    def close(self) -> int:
        """close(): trains what is buffered, runs the remaining -iters passes, waits for
        the GPU.  Returns the number of model-table rows."""
        self._flush_rows()
        n = C.c_int64()
        check(self._lib.hm_udtf_close(self._h, C.byref(n)))
        self._keep.clear()
        return n.value

    def forward(self):
        """forward(): the model table as numpy arrays (feature int32, weight float32[,
        Vif float32 (rows, factors)]), ascending feature order."""
        n = self.close()
        k = self.config.factors if self.function_name == "train_fm" else 0
        feat = np.empty(n, dtype=np.int32)
        wt = np.empty(n, dtype=np.float32)
        vif = np.empty((n, k), dtype=np.float32) if k else None
        m = C.c_int64()
        check(self._lib.hm_udtf_forward(self._h, _ptr(feat), _ptr(wt), _ptr(vif), n, C.byref(m)))
        if k:
            return feat[: m.value], wt[: m.value], vif[: m.value]
        return feat[: m.value], wt[: m.value]

    def save_model_table(self, path: str, binary: bool = False, delimiter: str = "\t") -> int:
        """close() + forward() into a file (text: feature<delim>weight[<delim>v1,...,vk] per line; or binary).
        Returns the number of rows written."""
        self._flush_rows()
        n = C.c_int64()
        check(self._lib.hm_udtf_save_model_table(self._h, path.encode(), int(binary), delimiter.encode()[:1], C.byref(n)))
        self._keep.clear()
        return n.value

    def load_model_table(self, path: str, binary: bool = False, delimiter: str = "\t") -> int:
        """Warm start from a model-table file: weights (train_fm: w0, w, V) exactly, optimizer slots from zero."""
        n = C.c_int64()
        check(self._lib.hm_udtf_load_model_table(self._h, path.encode(), int(binary), delimiter.encode()[:1], C.byref(n)))
        return n.value

    def destroy(self) -> None:
        if getattr(self, "_h", None) is not None and self._h:
            self._lib.hm_udtf_destroy(self._h)
            self._h = C.c_void_p()

    def __del__(self):
        try:
            self.destroy()
        except Exception:
            pass

    def __enter__(self):
        return self

    def __exit__(self, *exc):
        self.destroy()

    # ---- replica -----------------------------------------------------------------
    def sync(self) -> None:
        check(self._lib.hm_udtf_sync(self._h))
        self._keep.clear()

    @property
    def stream(self) -> int:
        return self._lib.hm_udtf_stream(self._h) or 0

    @property
    def step(self) -> int:
        t = C.c_int64()
        check(self._lib.hm_udtf_get_step(self._h, C.byref(t)))
        return t.value

    @step.setter
    def step(self, t: int) -> None:
        check(self._lib.hm_udtf_set_step(self._h, t))

    def arrays(self) -> dict:
        """name -> (device address, number of floats, interleave width)."""
        out = {}
        for i in range(self._lib.hm_udtf_num_arrays(self._h)):
            name, p, n, w = C.c_char_p(), C.c_void_p(), C.c_int64(), C.c_int32()
            check(self._lib.hm_udtf_array(self._h, i, C.byref(name), C.byref(p), C.byref(n), C.byref(w)))
            out[name.value.decode()] = (p.value, n.value, w.value)
        return out

    def array_view(self, name: str) -> _DeviceArray:
        p, n, _ = self.arrays()[name]
        return _DeviceArray(p, n, self)

    _LOGICAL = {"w": 0, "state": 1}
    _LOGICAL_FM = {"w": 0, "V": 1, "w0": 2, "gw": 3, "gV": 4, "gw0": 5}

    @property
    def state_width(self) -> int:
        """Optimizer slots per feature (0 SGD, 1 AdaGrad and the variance-keeping learners, 2 otherwise)."""
        return {0: 0, 1: 1, 5: 1}.get(self.config.opt, 2)

    def _logical(self, name: str) -> tuple[int, int]:
        """(index, length) of a logical model array."""
        d = self.config.dims
        if self.function_name == "train_fm":
            k = self.config.factors
            n = {"w": d, "V": d * k, "w0": 1, "gw": d, "gV": d * k, "gw0": 1}[name]
            return self._LOGICAL_FM[name], n
        return self._LOGICAL[name], d * (1 if name == "w" else self.state_width)

    def array_to_host(self, name: str) -> np.ndarray:
        """Copy a logical model array to the host.  Linear learners: "w" (dims floats),
        "state" (dims * state_width floats, a feature's slots adjacent).  train_fm: "w",
        "V" (dims * factors, row per feature), "w0", and with -opt adagrad the accumulators
        "gw", "gV", "gw0"."""
        which, n = self._logical(name)
        out = np.empty(n, dtype=np.float32)
        if n:
            check(self._lib.hm_udtf_read_array(self._h, which, _ptr(out), n))
        return out

    def array_from_host(self, name: str, values) -> None:
        """Overwrite a logical model array from the host (warm start)."""
        which, _ = self._logical(name)
        values = np.ascontiguousarray(values, dtype=np.float32).reshape(-1)
        check(self._lib.hm_udtf_write_array(self._h, which, _ptr(values), len(values)))

    def scale_replica(self, scale: float) -> None:
        check(self._lib.hm_udtf_scale_replica(self._h, scale))

    # ---- prediction --------------------------------------------------------------
    def predict(self, rowptr, feature, value) -> np.ndarray:
        rowptr = np.ascontiguousarray(rowptr, dtype=np.int64)
        feature = np.ascontiguousarray(feature, dtype=np.int32)
        value = np.ascontiguousarray(value, dtype=np.float32)
        out = np.empty(len(rowptr) - 1, dtype=np.float32)
        check(self._lib.hm_udtf_predict(self._h, len(out), _ptr(rowptr), _ptr(feature), _ptr(value), _ptr(out)))
        return out

    def predict_device(self, n_rows: int, rowptr, feature, value, out) -> None:
        check(self._lib.hm_udtf_predict_device(self._h, n_rows, _ptr(rowptr), _ptr(feature), _ptr(value), _ptr(out)))

    def last_margins(self, out: np.ndarray | None = None) -> np.ndarray:
        """Margins of the most recently trained mini-batch (device->host copy, waits)."""
        if out is None:
            out = np.empty(self.config.mini_batch, dtype=np.float32)
        n = C.c_int64()
        check(self._lib.hm_udtf_last_margins(self._h, _ptr(out), len(out), C.byref(n)))
        return out[: n.value]

    def hot_features(self) -> np.ndarray:
        """Hot features picked for the most recently sorted block (int32, at most 64, most frequent first)."""
        out = np.empty(64, dtype=np.int32)
        n = C.c_int32()
        check(self._lib.hm_udtf_hot_features(self._h, _ptr(out), 64, C.byref(n)))
        return out[: n.value]

    # ---- measurement ---------------------------------------------------------------
    def stats(self) -> HmStats:
        s = HmStats()
        check(self._lib.hm_udtf_stats(self._h, C.byref(s)))
        return s

    def set_profiling(self, on: bool) -> None:
        check(self._lib.hm_udtf_set_profiling(self._h, int(on)))

    def kernel_times(self) -> dict:
        """name -> dict(launches, ms, bytes) collected while profiling was on."""
        out = {}
        for i in range(self._lib.hm_udtf_num_kernels(self._h)):
            name, n, ms, b = C.c_char_p(), C.c_int64(), C.c_double(), C.c_double()
            check(self._lib.hm_udtf_kernel_time(self._h, i, C.byref(name), C.byref(n), C.byref(ms), C.byref(b)))
            out[name.value.decode()] = {"launches": n.value, "ms": ms.value, "bytes": b.value}
        return out


class TrainClassifierUDTF(UDTF):
    function_name = "train_classifier"


class TrainRegressorUDTF(UDTF):
    function_name = "train_regressor"


class FactorizationMachineUDTF(UDTF):
    function_name = "train_fm"


def train_classifier(options: str = "", device: int = 0) -> TrainClassifierUDTF:
    return TrainClassifierUDTF(options, device)


def train_regressor(options: str = "", device: int = 0) -> TrainRegressorUDTF:
    return TrainRegressorUDTF(options, device)


def train_fm(options: str = "", device: int = 0) -> FactorizationMachineUDTF:
    return FactorizationMachineUDTF(options, device)


# ---- online learners that share the mini-batch skeleton -----------------------------------
# Rosenblatt's perceptron and the passive-aggressive family of Crammer et al. (JMLR 2006): the function fixes the
# loss, the optimizer is plain SGD with eta = 1, the step size lives in the per-row coefficient.

class PerceptronUDTF(UDTF):
    function_name = "train_perceptron"


class PassiveAggressiveUDTF(UDTF):
    function_name = "train_pa"


class PassiveAggressive1UDTF(UDTF):
    function_name = "train_pa1"


class PassiveAggressive2UDTF(UDTF):
    function_name = "train_pa2"


class PassiveAggressive1RegressionUDTF(UDTF):
    function_name = "train_pa1_regr"


class PassiveAggressive2RegressionUDTF(UDTF):
    function_name = "train_pa2_regr"


def train_perceptron(options: str = "", device: int = 0) -> PerceptronUDTF:
    return PerceptronUDTF(options, device)


def train_pa(options: str = "", device: int = 0) -> PassiveAggressiveUDTF:
    return PassiveAggressiveUDTF(options, device)


def train_pa1(options: str = "", device: int = 0) -> PassiveAggressive1UDTF:
    return PassiveAggressive1UDTF(options, device)


def train_pa2(options: str = "", device: int = 0) -> PassiveAggressive2UDTF:
    return PassiveAggressive2UDTF(options, device)


def train_pa1_regr(options: str = "", device: int = 0) -> PassiveAggressive1RegressionUDTF:
    return PassiveAggressive1RegressionUDTF(options, device)


def train_pa2_regr(options: str = "", device: int = 0) -> PassiveAggressive2RegressionUDTF:
    return PassiveAggressive2RegressionUDTF(options, device)


# ---- learners that keep a variance per feature ---------------------------------------------
# Adaptive regularization of weights (Crammer, Kulesza, Dredze, NIPS 2009), diagonal covariance: array "state"
# holds the variances (1 at the start).

class AROWClassifierUDTF(UDTF):
    function_name = "train_arow"


class AROWRegressionUDTF(UDTF):
    function_name = "train_arow_regr"


class AROWeRegressionUDTF(UDTF):
    function_name = "train_arowe_regr"


def train_arow(options: str = "", device: int = 0) -> AROWClassifierUDTF:
    return AROWClassifierUDTF(options, device)


def train_arow_regr(options: str = "", device: int = 0) -> AROWRegressionUDTF:
    return AROWRegressionUDTF(options, device)


def train_arowe_regr(options: str = "", device: int = 0) -> AROWeRegressionUDTF:
    return AROWeRegressionUDTF(options, device)


# Confidence-weighted (Dredze, Crammer, Pereira 2008) and soft confidence-weighted (Wang, Zhao, Hoi 2012)
# learners: the same two passes as AROW with other closed forms for the step.

class ConfidenceWeightedUDTF(UDTF):
    function_name = "train_cw"


class SoftConfidenceWeighted1UDTF(UDTF):
    function_name = "train_scw"


class SoftConfidenceWeighted2UDTF(UDTF):
    function_name = "train_scw2"


def train_cw(options: str = "", device: int = 0) -> ConfidenceWeightedUDTF:
    return ConfidenceWeightedUDTF(options, device)


def train_scw(options: str = "", device: int = 0) -> SoftConfidenceWeighted1UDTF:
    return SoftConfidenceWeighted1UDTF(options, device)


def train_scw2(options: str = "", device: int = 0) -> SoftConfidenceWeighted2UDTF:
    return SoftConfidenceWeighted2UDTF(options, device)
