"""Sequence of the mini-batches of the LAST ingest block: when CTA 0 of every forward / update launch started, passed its
griddepcontrol.wait and ended (needs a -DHM_STAMP build; not part of the product).  usage: mb_sequence.py OPT MINI_BATCH ROWS"""
import ctypes as C, sys
import numpy as np, torch
import hivemall_b200 as hm
from hivemall_b200 import synth, _lib
opt = sys.argv[1] if len(sys.argv) > 1 else "ftrl"
mb = int(sys.argv[2]) if len(sys.argv) > 2 else 8192
rows = int(sys.argv[3]) if len(sys.argv) > 3 else 1 << 21
d = synth.criteo_like(rows, dims=1 << 24, seed=1, device="cuda")
u = hm.train_classifier(f"-opt {opt} -mini_batch {mb} -dims {1 << 24}")
for _ in range(3):
    u.process_device(d.rows, d.nnz, d.rowptr, d.feature, d.value, d.label)
u.sync()
e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
torch.cuda.synchronize(); import time; t0 = time.perf_counter()
u.process_device(d.rows, d.nnz, d.rowptr, d.feature, d.value, d.label)
t_host = time.perf_counter() - t0
u.sync()
dt = time.perf_counter() - t0
print(f"host returned from the call after {t_host * 1e3:.2f} ms = {t_host * 1e6 / (rows / mb):.2f} us per mini-batch of host work")
print(f"mini_batch {mb}: {rows} rows in {dt * 1e3:.2f} ms = {dt * 1e6 / (rows / mb):.2f} us per mini-batch (host clock, one call)")
buf = np.zeros(256 * 8, dtype=np.uint64)
assert _lib.load().hm_debug_read_seq(buf.ctypes.data_as(C.c_void_p), C.c_size_t(buf.size)) == 0
s = buf.reshape(256, 8).astype(np.int64)
s = s[(s[:, 0] > 0) & (s[:, 6] > 0)]
s = s[np.argsort(s[:, 1])]
print("mini-batches of the last block:", len(s))
cyc = np.diff(s[:, 1]) / 1e3
print("cycle (forward wait returned -> next forward wait returned), us: median %.2f  p10 %.2f  p90 %.2f  max %.2f" % (np.median(cyc), np.percentile(cyc, 10), np.percentile(cyc, 90), cyc.max()))
names = ["fwd start", "fwd wait returned", "fwd CTA0 end", "upd start", "upd wait returned", "upd CTA0 stepped", "upd merge done"]
rel = (s[:, :7] - s[:, 1:2]) / 1e3
for i, nme in enumerate(names):
    print(f"{nme:20s} median {np.median(rel[:, i]):7.2f}  p10 {np.percentile(rel[:, i], 10):7.2f}  p90 {np.percentile(rel[:, i], 90):7.2f}")
nxt = (s[1:, 1] - s[:-1, 6]) / 1e3
print("merge done -> next forward's wait returns: median %.2f  p90 %.2f" % (np.median(nxt), np.percentile(nxt, 90)))
