"""Times configs[0] (a9a-shaped, -mini_batch 128) on the GPU path and on the scalar oracle."""
import os, sys, time
sys.path.insert(0, os.path.join(os.path.dirname(__file__), "..")); sys.path.insert(0, os.path.join(os.path.dirname(__file__), "..", "tests"))
import numpy as np
import hivemall_b200 as hm
from hivemall_b200 import synth
import _oracle
opts = "-loss logloss -opt AdaGrad -mini_batch 128 -dims 124"
data = synth.a9a_like(32768, seed=0); rp, f, v, y = data.numpy()
for rep in range(3):
    u = hm.train_classifier(opts); u.sync()
    t0 = time.perf_counter(); u.process_batch(rp, f, v, y); u.close(); t1 = time.perf_counter()
    print("gpu  %.2f ms  %.2f M rows/s  launches %d" % ((t1 - t0) * 1e3, 32768 / (t1 - t0) / 1e6, u.stats().kernel_launches)); u.destroy()
cfg = _oracle.config_from(hm.parse_options("train_classifier", opts))
for rep in range(2):
    t0 = time.perf_counter(); _oracle.train_linear(cfg, rp, f, v, y); t1 = time.perf_counter()
    print("cpu  %.2f ms  %.2f M rows/s" % ((t1 - t0) * 1e3, 32768 / (t1 - t0) / 1e6))
for mb in (1024, 8192):
    o = f"-loss logloss -opt AdaGrad -mini_batch {mb} -dims {1<<20}"
    d = synth.criteo_like(1 << 20, dims=1 << 20, seed=1); a = d.numpy()
    u = hm.train_classifier(o); u.sync()
    t0 = time.perf_counter(); u.process_batch(*a); u.close(); t1 = time.perf_counter()
    print("gpu criteo mb=%d  %.2f ms  %.2f M rows/s" % (mb, (t1 - t0) * 1e3, (1 << 20) / (t1 - t0) / 1e6)); u.destroy()
