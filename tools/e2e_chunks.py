"""Dev aid: end-to-end rate of fjp_score_batch / fjp_score_batch_f32 on C4 for several pipeline chunk sizes (FJP_CHUNK_ROWS)."""
import ctypes as C, os, sys, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np, torch
from flink_jpmml_b200 import capi, pmmlgen as g
eng = capi.Engine(0); L = capi.lib()
pm, _ = g.gbt_pmml(); m = eng.compile(pm); F = 100; n = 1 << 21
X, pres, pb = g.events(1 << 16, F, 1, 0.1)
modes = eng.float_row_modes(m)
Xf = np.tile(eng.narrow_rows(X, modes), (n >> 16, 1)); Xd = np.tile(X, (n >> 16, 1)); P = np.tile(pb, (n >> 16, 1)).view(np.int32)
hXf = torch.from_numpy(Xf).pin_memory(); hXd = torch.from_numpy(Xd).pin_memory(); hP = torch.from_numpy(P).pin_memory()
hS = torch.empty(n, dtype=torch.float64).pin_memory(); hT = torch.empty(n, dtype=torch.uint8).pin_memory()
for rows in (0, 30000, 60000, 100000, 150000, 262144, 524288):
    if rows: os.environ["FJP_CHUNK_ROWS"] = str(rows)
    out = []
    for f32 in (True, False):
        def step():
            if f32: rc = L.fjp_score_batch_f32(eng._h, m._h, hXf.data_ptr(), hP.data_ptr(), n, F, hS.data_ptr(), hT.data_ptr(), None)
            else: rc = L.fjp_score_batch(eng._h, m._h, hXd.data_ptr(), hP.data_ptr(), None, n, F, hS.data_ptr(), hT.data_ptr(), None)
            assert rc == 0
        step(); step()
        t = time.perf_counter()
        for _ in range(5): step()
        out.append(n * 5 / (time.perf_counter() - t))
    print(f"chunk rows {rows or 'default'}: float rows {out[0]:.3e} ev/s ({out[0]*416/1e9:.1f} GB/s)  double rows {out[1]:.3e} ev/s ({out[1]*816/1e9:.1f} GB/s)", flush=True)
