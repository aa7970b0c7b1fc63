"""configs[0] (a9a-shaped, -mini_batch 128): wall time of a warm process() call and its kernel times,
next to the scalar CPU oracle.  (The first call of a learner also pays for its buffers.)"""
import os, sys, time
sys.path.insert(0, os.path.join(os.path.dirname(__file__), "..")); sys.path.insert(0, os.path.join(os.path.dirname(__file__), "..", "tests"))
import numpy as np
import hivemall_b200 as hm
from hivemall_b200 import synth
import _oracle
opts = "-loss logloss -opt AdaGrad -mini_batch 128 -dims 124"
data = synth.a9a_like(32768, seed=0); rp, f, v, y = data.numpy()
u = hm.train_classifier(opts)
t0 = time.perf_counter(); u.process_batch(rp, f, v, y); u.sync(); t1 = time.perf_counter()
print("gpu first call (allocations included) %.2f ms" % ((t1 - t0) * 1e3))
for rep in range(3):
    t0 = time.perf_counter(); u.process_batch(rp, f, v, y); u.sync(); t1 = time.perf_counter()
    print("gpu warm call  %.2f ms  %.2f M rows/s" % ((t1 - t0) * 1e3, 32768 / (t1 - t0) / 1e6))
u.set_profiling(True); u.process_batch(rp, f, v, y); u.sync(); u.set_profiling(False)
print("kernels of one call:", {k: (v["launches"], round(v["ms"], 3)) for k, v in u.kernel_times().items() if v["launches"]})
u.destroy()
cfg = _oracle.config_from(hm.parse_options("train_classifier", opts))
for rep in range(2):
    t0 = time.perf_counter(); _oracle.train_linear(cfg, rp, f, v, y); t1 = time.perf_counter()
    print("cpu oracle, 1 thread  %.2f ms  %.2f M rows/s" % ((t1 - t0) * 1e3, 32768 / (t1 - t0) / 1e6))
