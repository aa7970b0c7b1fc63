"""Loader and checkers of the reference-vector slot (tests/golden/hivemall/, see its README.md).

Every checker takes the parsed JSON of one file and an implementation object -- OracleImpl (CPU
oracle) or GpuImpl (CUDA path through the C-ABI) -- and asserts the clause.  Test infrastructure."""
from __future__ import annotations

import glob
import json
import os

import numpy as np

import _oracle

HERE = os.path.dirname(os.path.abspath(__file__))
SLOT = os.path.join(HERE, "golden", "hivemall")
TEMPLATE_MARK = "oracle (NOT the reference)"
RTOL, ATOL = 1e-5, 1e-7


def files(kind: str, directory: str = SLOT, allow_templates: bool = False):
    out = []
    for path in sorted(glob.glob(os.path.join(directory, f"{kind}*.json"))):
        with open(path) as f:
            j = json.load(f)
        marked = isinstance(j, dict) and j.get("source") == TEMPLATE_MARK
        if marked and not allow_templates:
            continue
        out.append((path, j))
    return out


def dense(sp, n, dtype=np.float32):
    a = np.zeros(n, dtype=dtype)
    if sp:
        a[np.asarray(sp["index"], dtype=np.int64)] = np.asarray(sp["value"], dtype=np.float64).astype(dtype)
    return a


def sparse(a):
    idx = np.flatnonzero(a)
    return {"index": idx.tolist(), "value": [float(x) for x in a[idx]]}


def rows_of(j):
    r = j["rows"]
    return (np.asarray(r["rowptr"], np.int64), np.asarray(r["feature"], np.int32),
            np.asarray(r["value"], np.float32), np.asarray(r["label"], np.float32))


def close(got, want, what, rtol=RTOL):
    got, want = np.asarray(got, np.float64), np.asarray(want, np.float64)
    err = np.abs(got - want)
    tol = rtol * np.maximum(np.abs(got), np.abs(want)) + ATOL
    bad = err > tol
    assert not bad.any(), f"{what}: {int(bad.sum())} entries differ; worst at {int(np.argmax(err - tol))}: " \
                          f"got {got.flat[int(np.argmax(err - tol))]!r} want {want.flat[int(np.argmax(err - tol))]!r}"


# ---- implementations -------------------------------------------------------------------

class OracleImpl:
    name = "oracle"

    def mhash(self, strings, nf, seed):
        return _oracle.mhash(strings, nf, seed)

    def feature_hashing(self, strings, nf, seed):
        return None          # the oracle restates mhash only; "name:value" splitting is checked on the GPU side

    def config_fields(self, fn, options):
        c = _oracle.config_from_options(fn, options)
        return {n.rstrip("_") if n != "lambda_" else "lambda": getattr(c, n) for n, _ in _oracle.HmoConfig._fields_}

    def linear(self, fn, options, rows, before=None):
        cfg = _oracle.config_from_options(fn, options)
        d = int(cfg.dims)
        b = before or {}
        res = _oracle.train_linear(cfg, *rows, w=dense(b.get("w"), d), s1=dense(b.get("s1"), d), s2=dense(b.get("s2"), d),
                                   t=int(b.get("t", 0)))
        return {"t": res.t, "w": res.w, "s1": res.s1, "s2": res.s2}

    def fm_init(self, options, dims):
        cfg = _oracle.config_from_options("train_fm", options + f" -dims {dims}")
        return _oracle.fm_init_v(cfg).reshape(dims, int(cfg.factors))

    def table(self, fn, options, rows):
        cfg = _oracle.config_from_options(fn, options)
        if fn == "train_fm":
            r = _oracle.train_fm(cfg, *rows)
            feat = np.flatnonzero(r.touched).astype(np.int32)
            k = int(cfg.factors)
            return {"feature": np.concatenate([[-1], feat]), "weight": np.concatenate([r.w0, r.w[feat]]),
                    "vif": np.concatenate([np.zeros((1, k), np.float32), r.V.reshape(-1, k)[feat]])}
        res = _oracle.train_linear(cfg, *rows)
        f, w = _oracle.model_table(res.w)
        return {"feature": f, "weight": w}

    def mix(self, fn, options, replicas):
        outs = [self.linear(fn, options, r) for r in replicas]
        p = np.float32(1.0 / len(outs))
        return {k: sum(o[k] for o in outs).astype(np.float32) * p for k in ("w", "s1", "s2")}


class GpuImpl:
    name = "gpu"

    def __init__(self):
        import hivemall_b200 as hm
        self.hm = hm

    def mhash(self, strings, nf, seed):
        return self.hm.mhash(strings, num_features=nf, seed=seed)

    def feature_hashing(self, strings, nf, seed):
        return self.hm.feature_hashing(strings, num_features=nf, seed=seed)

    def config_fields(self, fn, options):
        c = self.hm.parse_options(fn, options)
        return {(n if n != "lambda_" else "lambda"): getattr(c, n) for n, _ in type(c)._fields_ if n != "reserved"}

    def _slots(self, u, d, b):
        """(s1, s2) of the file -> the learner's interleaved state array."""
        width = u.state_width
        if width == 0:
            return None
        s1, s2 = dense(b.get("s1"), d), dense(b.get("s2"), d)
        return s1 if width == 1 else np.stack([s1, s2], axis=1).reshape(-1)

    def linear(self, fn, options, rows, before=None):
        u = self.hm.UDTF(options, 0, function_name=fn)
        d = int(u.config.dims)
        if before:
            ftrl = u.config.opt == 4
            if not ftrl:
                u.array_from_host("w", dense(before.get("w"), d))
            st = self._slots(u, d, before)
            if st is not None:
                u.array_from_host("state", st)
            u.step = int(before.get("t", 0))
        u.process_batch(*rows)
        u.close()
        w, st, width = u.array_to_host("w"), u.array_to_host("state"), u.state_width
        out = {"t": u.step, "w": w,
               "s1": st if width == 1 else (st.reshape(-1, 2)[:, 0] if width == 2 else np.zeros(d, np.float32)),
               "s2": st.reshape(-1, 2)[:, 1] if width == 2 else np.zeros(d, np.float32)}
        u.destroy()
        return out

    def fm_init(self, options, dims):
        u = self.hm.train_fm(options + f" -dims {dims}")
        v = u.array_to_host("V").reshape(dims, u.config.factors)
        u.destroy()
        return v

    def table(self, fn, options, rows):
        u = self.hm.UDTF(options, 0, function_name=fn)
        u.process_batch(*rows)
        out = u.forward()
        u.destroy()
        return dict(zip(("feature", "weight", "vif"), out))

    def mix(self, fn, options, replicas):
        """P replicas trained one after the other on this GPU, then averaged the way mix.py averages
        them across GPUs: every replica array summed, then scaled by 1/P in fp32."""
        outs = [self.linear(fn, options, r) for r in replicas]
        p = np.float32(1.0 / len(outs))
        return {k: sum(o[k] for o in outs).astype(np.float32) * p for k in ("w", "s1", "s2")}


# ---- checkers ------------------------------------------------------------------------------

def check_hash(j, impl):
    nf, seed = int(j["num_features"]), int(j["seed"])
    plain = [c for c in j["cases"] if "value" not in c]
    if plain:
        got = impl.mhash([c["string"] for c in plain], nf, seed)
        want = np.asarray([c["index"] for c in plain])
        assert np.array_equal(got, want), f"mhash differs at {np.flatnonzero(got != want)[:10].tolist()}"
    valued = [c for c in j["cases"] if "value" in c]
    if valued:
        r = impl.feature_hashing([c["string"] for c in valued], nf, seed)
        if r is not None:
            assert np.array_equal(r[0], np.asarray([c["index"] for c in valued]))
            assert np.array_equal(r[1], np.asarray([c["value"] for c in valued], np.float32))


def check_options(j, impl):
    for case in j if isinstance(j, list) else j["cases"]:
        got = impl.config_fields(case["function"], case["options"])
        for k, v in case["fields"].items():
            if k not in got:
                continue        # (not an oracle field: hash_seed, mix_interval)
            if isinstance(v, float):
                assert abs(got[k] - v) <= 1e-6 * max(1.0, abs(v)), f"{case['options']!r}: {k} = {got[k]} want {v}"
            else:
                assert got[k] == v, f"{case['options']!r}: {k} = {got[k]} want {v}"


def check_step(j, impl):
    d = int(j["dims"])
    got = impl.linear(j["function"], j["options"], rows_of(j), j.get("before"))
    want = j["after"]
    assert got["t"] == int(want["t"])
    for k in ("w", "s1", "s2"):
        if k in want:
            close(got[k], dense(want[k], d), k, float(j.get("rtol", RTOL)))


def check_fm_init(j, impl):
    v = impl.fm_init(j["options"], int(j["dims"]))
    close(v[np.asarray(j["V"]["feature"], np.int64)], np.asarray(j["V"]["rows"], np.float32), "V")


def check_table(j, impl):
    got = impl.table(j["function"], j["options"], rows_of(j))
    want = j["table"]
    wf = np.asarray(want["feature"], np.int64)
    order = np.argsort(wf, kind="stable") if j.get("order") == "sorted" else np.arange(len(wf))
    assert np.array_equal(np.asarray(got["feature"], np.int64), wf[order]), "model-table features / order differ"
    close(got["weight"], np.asarray(want["weight"], np.float32)[order], "weight")
    if "vif" in want and want["vif"] is not None:
        close(got["vif"], np.asarray(want["vif"], np.float32)[order], "vif")


def check_mix(j, impl):
    d = int(j["dims"])
    reps = [rows_of(r) for r in j["replicas"]]
    got = impl.mix(j["function"], j["options"], reps)
    for k in ("w", "s1", "s2"):
        if k in j["after"]:
            close(got[k], dense(j["after"][k], d), k)


CHECKERS = {"hash": check_hash, "options": check_options, "step": check_step, "fm_init": check_fm_init,
            "table": check_table, "mix": check_mix}
