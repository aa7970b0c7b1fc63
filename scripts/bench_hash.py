"""Throughput of the mhash kernel on device-resident strings (CUDA events), vs the CPU oracle."""
import ctypes as C, os, sys, time
sys.path.insert(0, os.path.join(os.path.dirname(__file__), "..")); sys.path.insert(0, os.path.join(os.path.dirname(__file__), "..", "tests"))
import numpy as np, torch
from hivemall_b200 import _lib
import _oracle
lib = _lib.load()
n = 50_000_000
rng = np.random.default_rng(0)
lens = rng.integers(4, 21, size=n).astype(np.int64)          # "field=value"-like names, 4..20 bytes
off = np.zeros(n + 1, np.int64); off[1:] = np.cumsum(lens)
data = rng.integers(97, 123, size=int(off[-1]), dtype=np.uint8)
d_data = torch.from_numpy(data).cuda(); d_off = torch.from_numpy(off).cuda(); d_out = torch.empty(n, dtype=torch.int32, device="cuda")
st = torch.cuda.Stream()
def run():
    _lib.check(lib.hm_mhash_device(C.c_void_p(d_data.data_ptr()), C.c_void_p(d_off.data_ptr()), n, 0x9747B28C, 1 << 24,
                                   C.c_void_p(d_out.data_ptr()), C.c_void_p(st.cuda_stream)))
for _ in range(3): run()
st.synchronize()
a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
with torch.cuda.stream(st):
    a.record()
for _ in range(10): run()
with torch.cuda.stream(st):
    b.record()
st.synchronize()
ms = a.elapsed_time(b) / 10
alg = int(off[-1]) + 12 * n
print(f"mhash: {n / ms / 1e3:.1f} M strings/s, {ms:.2f} ms for {n} strings ({off[-1] / n:.1f} B avg); algorithmic {alg / ms / 1e6:.0f} GB/s")
m = 2_000_000
t0 = time.perf_counter(); ref = np.empty(m, np.int32)
_oracle.lib().hmo_mhash_batch(C.c_void_p(data.ctypes.data), C.c_void_p(off.ctypes.data), m, 0x9747B28C, 1 << 24, C.c_void_p(ref.ctypes.data))
dt = time.perf_counter() - t0
print(f"oracle (1 core): {m / dt / 1e6:.1f} M strings/s; first {m} equal: {np.array_equal(ref, d_out[:m].cpu().numpy())}")
