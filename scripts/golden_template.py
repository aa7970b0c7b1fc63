#!/usr/bin/env python
"""Write a template of every reference-vector file (tests/golden/hivemall/README.md) from the CPU
oracle's own output.  The files show the FORMAT to fill with the real jar's numbers; they carry
"source": "oracle (NOT the reference)" and the reference-vector tests ignore files so marked.

usage: python scripts/golden_template.py OUT_DIR"""
import json
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "tests"))
import _golden  # noqa: E402
import _oracle  # noqa: E402


def small_rows(n, dims, seed, per_row=6):
    rng = np.random.default_rng(seed)
    lens = rng.integers(0, per_row + 1, n)
    rowptr = np.zeros(n + 1, np.int64)
    rowptr[1:] = np.cumsum(lens)
    feature = rng.integers(0, dims, rowptr[-1]).astype(np.int32)
    value = np.round(rng.normal(0, 1, rowptr[-1]), 3).astype(np.float32)
    label = (rng.random(n) < 0.5).astype(np.float32)
    return rowptr, feature, value, label


def rows_json(r):
    return {"rowptr": r[0].tolist(), "feature": r[1].tolist(), "value": [float(x) for x in r[2]],
            "label": [float(x) for x in r[3]]}


def main(out_dir):
    os.makedirs(out_dir, exist_ok=True)
    impl = _golden.OracleImpl()
    mark = {"source": _golden.TEMPLATE_MARK}

    def dump(name, obj):
        with open(os.path.join(out_dir, name), "w") as f:
            json.dump(obj, f)

    strings = ["age", "workclass=Private", "", "height", "ユーザー"]
    seed, nf = 0x9747B28C, 1 << 24
    dump("hash_template.json", {**mark, "seed": seed, "num_features": nf,
                                "cases": [{"string": s, "index": int(i)} for s, i in zip(strings, impl.mhash(strings, nf, seed))]})
    dump("options_template.json", {**mark, "cases": [
        {"function": "train_classifier", "options": "-loss logloss -opt AdaGrad -mini_batch 128",
         "fields": {"loss": 0, "opt": 1, "mini_batch": 128, "eta0": 0.1, "eps": 1e-6, "dims": 1 << 24}},
        {"function": "train_fm", "options": "-factors 16 -classification", "fields": {"factors": 16, "loss": 0, "sigma": 0.1}},
        {"function": "train_pa1", "options": "-c 0.5 -mini_batch 4", "fields": {"loss": 5, "opt": 0, "eta0": 1.0, "pa_c": 0.5, "mini_batch": 4}},
        {"function": "train_pa2_regr", "options": "", "fields": {"loss": 8, "pa_c": 1.0, "pa_epsilon": 0.1}}]})
    dims = 64
    for opt in ("sgd", "adagrad", "adadelta", "adam", "ftrl"):
        options = f"-loss logloss -opt {opt} -mini_batch 16 -dims {dims}"
        first = impl.linear("train_classifier", options, small_rows(16, dims, 1))
        before = {"t": first["t"], **{k: _golden.sparse(first[k]) for k in ("w", "s1", "s2")}}
        rows = small_rows(16, dims, 2)
        after = impl.linear("train_classifier", options, rows, before)
        dump(f"step_{opt}_template.json", {**mark, "function": "train_classifier", "options": options, "dims": dims,
                                            "rows": rows_json(rows), "before": before,
                                            "after": {"t": after["t"], **{k: _golden.sparse(after[k]) for k in ("w", "s1", "s2")}}})
    # the online learners on the same skeleton, the further plain losses (3.3): one mini-batch step each
    for k, (fn, options) in enumerate((("train_classifier", f"-loss squared_hinge -opt sgd -mini_batch 16 -dims {dims}"),
                        ("train_classifier", f"-loss modified_huber -opt sgd -mini_batch 16 -dims {dims}"),
                        ("train_regressor", f"-loss huber -huber_c 0.5 -opt sgd -mini_batch 16 -dims {dims}"),
                        ("train_regressor", f"-loss quantile -quantile 0.8 -opt sgd -mini_batch 16 -dims {dims}"),
                        ("train_regressor", f"-loss epsilon_insensitive -epsilon 0.1 -opt sgd -mini_batch 16 -dims {dims}"),
                        ("train_perceptron", f"-mini_batch 16 -dims {dims}"),
                        ("train_pa", f"-mini_batch 16 -dims {dims}"),
                        ("train_pa1", f"-mini_batch 16 -dims {dims} -c 0.1"),
                        ("train_pa2", f"-mini_batch 16 -dims {dims} -c 0.1"),
                        ("train_pa1_regr", f"-mini_batch 16 -dims {dims} -c 0.1 -epsilon 0.05"),
                        ("train_pa2_regr", f"-mini_batch 16 -dims {dims} -c 0.1 -epsilon 0.05"),
                        # (s1 = the per-feature variance, 1 where the learner has seen nothing)
                        ("train_arow", f"-mini_batch 16 -dims {dims} -r 0.5"),
                        ("train_arow_regr", f"-mini_batch 16 -dims {dims} -r 0.5"),
                        ("train_arowe_regr", f"-mini_batch 16 -dims {dims} -r 0.5 -epsilon 0.05"),
                        ("train_cw", f"-mini_batch 16 -dims {dims} -phi 0.5"),
                        ("train_scw", f"-mini_batch 16 -dims {dims} -phi 0.8 -c 0.1"),
                        ("train_scw2", f"-mini_batch 16 -dims {dims} -c 0.1"))):
        first = impl.linear(fn, options, small_rows(16, dims, 1))
        before = {"t": first["t"], **{k: _golden.sparse(first[k]) for k in ("w", "s1", "s2")}}
        rows = small_rows(16, dims, 2)
        after = impl.linear(fn, options, rows, before)
        dump(f"step_{k:02d}_{fn}_template.json", {**mark, "function": fn, "options": options, "dims": dims,
                                           "rows": rows_json(rows), "before": before,
                                           "after": {"t": after["t"], **{k: _golden.sparse(after[k]) for k in ("w", "s1", "s2")}}})
    v = impl.fm_init("-factors 4 -seed 31 -sigma 0.1", dims)
    dump("fm_init_template.json", {**mark, "options": "-factors 4 -seed 31 -sigma 0.1", "dims": dims,
                                   "V": {"feature": [0, 5, 63], "rows": [[float(x) for x in v[i]] for i in (0, 5, 63)]}})
    for fn, options in (("train_classifier", f"-loss logloss -opt adagrad -mini_batch 8 -dims {dims}"),
                        ("train_fm", f"-factors 4 -classification -opt adagrad -mini_batch 8 -dims {dims}")):
        rows = small_rows(40, dims, 3)
        t = impl.table(fn, options, rows)
        dump(f"table_{fn}_template.json", {**mark, "function": fn, "options": options, "dims": dims, "rows": rows_json(rows),
                                           "table": {"feature": [int(x) for x in t["feature"]],
                                                     "weight": [float(x) for x in t["weight"]],
                                                     "vif": None if "vif" not in t else [[float(x) for x in r] for r in t["vif"]]}})
    options = f"-loss logloss -opt ftrl -mini_batch 8 -dims {dims} -alpha 0.5"
    reps = [small_rows(24, dims, 10 + r) for r in range(2)]
    after = impl.mix("train_classifier", options, reps)
    dump("mix_template.json", {**mark, "function": "train_classifier", "options": options, "dims": dims,
                               "replicas": [{"rows": rows_json(r)} for r in reps],
                               "after": {k: _golden.sparse(after[k]) for k in ("w", "s1", "s2")}})


if __name__ == "__main__":
    main(sys.argv[1] if len(sys.argv) > 1 else os.path.join(ROOT, "gpurun_out", "golden_template"))
