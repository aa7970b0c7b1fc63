import sys, os
sys.path.insert(0, os.path.join(os.path.dirname(__file__), "..")); sys.path.insert(0, os.path.join(os.path.dirname(__file__), "..", "tests"))
import numpy as np, torch
import hivemall_b200 as hm
from hivemall_b200 import synth
import _oracle
from _util import run_gpu_linear, run_oracle_linear
data = synth.ragged(3000, dims=5000, max_nnz=60, seed=3)
B = 256
for k in range(1, 13):
    n = min(3000, k * B)
    sub = synth.Csr(data.rowptr[: n + 1], data.feature, data.value, data.label[:n])
    options = f"-loss logloss -opt adagrad -mini_batch {B} -dims 5000 -eta0 0.05"
    u, w, st = run_gpu_linear("train_classifier", options, sub)
    ref, rst = run_oracle_linear("train_classifier", options, sub)
    err = np.abs(w.astype(np.float64) - ref.w)
    rel = err / np.maximum(np.maximum(np.abs(w), np.abs(ref.w)), 1e-30)
    i = int(np.argmax(err))
    print(k, "max abs err", err.max(), "at", i, "w", w[i], ref.w[i], "G", st[i], rst[i], "relmax", rel[err > 1e-7].max() if (err > 1e-7).any() else 0)
    if err.max() > 1e-7:
        rp, f, v, y = sub.numpy()
        # occurrences of feature i in the last minibatch
        r0 = (k - 1) * B
        occ = [(r, p) for r in range(r0, n) for p in range(rp[r], rp[r + 1]) if f[p] == i]
        print("  occurrences in last mb:", len(occ), [(r, float(v[p])) for r, p in occ][:20])
        # history of feature i
        for kk in range(1, k + 1):
            nn = min(3000, kk * B)
            s2 = synth.Csr(data.rowptr[: nn + 1], data.feature, data.value, data.label[:nn])
            u2, w2, st2 = run_gpu_linear("train_classifier", options, s2)
            r2, rst2 = run_oracle_linear("train_classifier", options, s2)
            print("   after mb", kk, "gpu", repr(w2[i]), repr(st2[i]), "ref", repr(r2.w[i]), repr(rst2[i]))
        break
