"""Per-CTA timeline of the last linear_update launch (needs a -DHM_STAMP build: scripts/build_variant.sh stamp hivemall_b200/csrc/hm_linear.cu -DHM_STAMP; not part of the product)."""
import ctypes as C, sys
import numpy as np, torch
import hivemall_b200 as hm
from hivemall_b200 import synth, _lib
opt = sys.argv[1] if len(sys.argv) > 1 else "ftrl"
mb = int(sys.argv[2]) if len(sys.argv) > 2 else 65536
d = synth.criteo_like(mb * 6, dims=1 << 24, seed=1, device="cuda")
u = hm.train_classifier(f"-opt {opt} -mini_batch {mb} -dims {1 << 24}")
u.process_device(d.rows, d.nnz, d.rowptr, d.feature, d.value, d.label); u.sync()
lib = _lib.load()
buf = np.zeros(4096 * 8, dtype=np.uint64)
rc = lib.hm_debug_read_stamps(buf.ctypes.data_as(C.c_void_p), C.c_size_t(buf.size)); assert rc == 0
s = buf.reshape(4096, 8).astype(np.int64)
act = (s[:, 0] > 0) & (s[:, 5] > 0)     # (the dense-column CTA leaves stamp 0 only)
s = s[act]; print("CTAs", len(s))
t0 = s[:, 0].min()
end = max(s[:, 5].max(), s[:, 6].max())
print("launch span us", (end - t0) / 1e3, " last non-merge end", (s[:, 5].max() - t0) / 1e3)
names = ["start->loaded", "loaded->warpscan", "warpscan->reduced", "reduced->published", "apply", ]
for i, nme in enumerate(names):
    dlt = (s[:, i + 1] - s[:, i]) / 1e3
    print(f"{nme:22s} median {np.median(dlt):7.2f}  p90 {np.percentile(dlt, 90):7.2f}  max {dlt.max():7.2f}")
tot = (s[:, 5] - s[:, 0]) / 1e3
print("CTA total median", np.median(tot), "p90", np.percentile(tot, 90))
st = np.sort((s[:, 0] - t0) / 1e3)
print("CTA start times (us) deciles", np.round(np.percentile(st, [0, 10, 20, 30, 40, 50, 60, 70, 80, 90, 100]), 1))
en = np.sort((s[:, 5] - t0) / 1e3)
print("CTA end   times (us) deciles", np.round(np.percentile(en, [0, 10, 20, 30, 40, 50, 60, 70, 80, 90, 100]), 1))
m = s[s[:, 6] > s[:, 0]]
if len(m): print("merge CTA: finished tile at", (m[0, 5] - t0) / 1e3, "merge done at", (m[0, 6] - t0) / 1e3)
