"""Dev aid: is the float-row e2e below the copy ceiling because of the pipeline or because of the buffer?  Raw H2D rate of the very
buffers the e2e uses (float rows, double rows), in chunk-sized copies, then the e2e rate on them; also with a write-combined buffer."""
import ctypes as C, os, sys, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np, torch
from flink_jpmml_b200 import capi, pmmlgen as g
eng = capi.Engine(0); L = capi.lib(); dev = torch.device("cuda:0")
pm, _ = g.gbt_pmml(); m = eng.compile(pm); F = 100; n = 1 << 21
X, pres, pb = g.events(1 << 16, F, 1, 0.1)
modes = eng.float_row_modes(m)
Xf = np.tile(eng.narrow_rows(X, modes), (n >> 16, 1)); Xd = np.tile(X, (n >> 16, 1)); P = np.tile(pb, (n >> 16, 1)).view(np.int32)
hXf = torch.from_numpy(Xf).pin_memory(); hXd = torch.from_numpy(Xd).pin_memory(); hP = torch.from_numpy(P).pin_memory()
hS = torch.empty(n, dtype=torch.float64).pin_memory(); hT = torch.empty(n, dtype=torch.uint8).pin_memory()
cudart = C.CDLL("libcudart.so.12")
def raw(h, chunk_rows):
    d = torch.empty_like(h[:chunk_rows], device=dev); torch.cuda.synchronize()
    best = 0
    for _ in range(3):
        t = time.perf_counter()
        for lo in range(0, h.shape[0], chunk_rows): d.copy_(h[lo:lo + chunk_rows], non_blocking=True)
        torch.cuda.synchronize(); best = max(best, h.numel() * h.element_size() / (time.perf_counter() - t) / 1e9)
    return best
print(f"raw H2D of the float rows {raw(hXf, 1 << 18):.1f} GB/s, of the double rows {raw(hXd, 1 << 18):.1f} GB/s, present words {raw(hP, 1 << 18):.1f} GB/s", flush=True)
def e2e(ptr, f32):
    def step():
        if f32: rc = L.fjp_score_batch_f32(eng._h, m._h, ptr, hP.data_ptr(), n, F, hS.data_ptr(), hT.data_ptr(), None)
        else: rc = L.fjp_score_batch(eng._h, m._h, ptr, hP.data_ptr(), None, n, F, hS.data_ptr(), hT.data_ptr(), None)
        assert rc == 0
    step(); step(); t = time.perf_counter()
    for _ in range(5): step()
    return n * 5 / (time.perf_counter() - t)
r = e2e(hXf.data_ptr(), True); print(f"e2e float rows {r:.3e} ev/s = {r * 416 / 1e9:.1f} GB/s", flush=True)
r = e2e(hXd.data_ptr(), False); print(f"e2e double rows {r:.3e} ev/s = {r * 816 / 1e9:.1f} GB/s", flush=True)
for rows in (65536, 131072):
    os.environ["FJP_CHUNK_ROWS"] = str(rows); r = e2e(hXf.data_ptr(), True); print(f"e2e float rows, chunks of {rows}: {r:.3e} ev/s = {r * 416 / 1e9:.1f} GB/s", flush=True)
os.environ.pop("FJP_CHUNK_ROWS")
# write-combined pinned buffer for the float rows
p = C.c_void_p(); rc = cudart.cudaHostAlloc(C.byref(p), C.c_size_t(Xf.nbytes), C.c_uint(4))   # cudaHostAllocWriteCombined
assert rc == 0
C.memmove(p, Xf.ctypes.data, Xf.nbytes)
r = e2e(p.value, True); print(f"e2e float rows from a write-combined buffer {r:.3e} ev/s = {r * 416 / 1e9:.1f} GB/s", flush=True)
# a second ordinary pinned float buffer (placement luck?)
hXf2 = torch.empty_like(hXf).pin_memory(); hXf2.copy_(hXf)
print(f"raw H2D of a second float buffer {raw(hXf2, 1 << 18):.1f} GB/s", flush=True)
r = e2e(hXf2.data_ptr(), True); print(f"e2e float rows, second buffer {r:.3e} ev/s = {r * 416 / 1e9:.1f} GB/s", flush=True)
