"""Timeline of the LAST mini-batch's forward and update launches from per-CTA globaltimer stamps (needs a -DHM_STAMP
build: scripts/build_variant.sh stamp -DHM_STAMP; not part of the product).  usage: mb_timeline.py OPT MINI_BATCH"""
import ctypes as C, sys
import numpy as np, torch
import hivemall_b200 as hm
from hivemall_b200 import synth, _lib
opt = sys.argv[1] if len(sys.argv) > 1 else "ftrl"
mb = int(sys.argv[2]) if len(sys.argv) > 2 else 8192
d = synth.criteo_like(mb * 24, dims=1 << 24, seed=1, device="cuda")
u = hm.train_classifier(f"-opt {opt} -mini_batch {mb} -dims {1 << 24}")
u.process_device(d.rows, d.nnz, d.rowptr, d.feature, d.value, d.label); u.sync()
lib = _lib.load()
def read(fn):
    buf = np.zeros(4096 * 8, dtype=np.uint64)
    rc = getattr(lib, fn)(buf.ctypes.data_as(C.c_void_p), C.c_size_t(buf.size)); assert rc == 0
    return buf.reshape(4096, 8).astype(np.int64)
f, s = read("hm_debug_read_fstamps"), read("hm_debug_read_stamps")
f = f[(f[:, 0] > 0) & (f[:, 5] > 0)]
s = s[(s[:, 0] > 0) & (s[:, 5] > 0)]
# keep the forward CTAs of the last launch only (an earlier, larger launch may have left stamps in higher rows)
t_last = f[:, 0].max()
f = f[f[:, 0] > t_last - 200_000]
T0 = np.median(f[:, 2])       # forward's griddepcontrol.wait returned: the previous update has finished
def rel(a): return np.round((np.percentile(a, [0, 50, 100]) - T0) / 1e3, 2)
print(f"mini_batch {mb}: forward CTAs {len(f)}, update CTAs {len(s)}  (us relative to the forward's wait returning; min / median / max)")
print("forward  start          ", rel(f[:, 0]))
print("forward  before wait    ", rel(f[:, 1]))
print("forward  wait returned  ", rel(f[:, 2]))
print("forward  prologue done  ", rel(f[:, 3]))
print("forward  rows done      ", rel(f[:, 4]))
print("forward  end            ", rel(f[:, 5]))
print("update   start          ", rel(s[:, 0]))
print("update   wait returned  ", rel(s[:, 7]))
print("update   tile staged    ", rel(s[:, 1]))
print("update   scanned        ", rel(s[:, 2]))
print("update   published      ", rel(s[:, 4]))
print("update   stepped        ", rel(s[:, 5]))
m = s[s[:, 6] > s[:, 0]]
if len(m): print("update   merge done     ", np.round((m[:, 6].max() - T0) / 1e3, 2))
