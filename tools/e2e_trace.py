"""Dev aid: device timeline of one float-row and one double-row host-buffer call on C4 (FJP_TRACE_CHUNKS=1)."""
import os, sys
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
for f32 in (True, False):
    for it in range(3):
        if it == 2: os.environ["FJP_TRACE_CHUNKS"] = "1"; print("float rows" if f32 else "double rows", file=sys.stderr, flush=True)
        rc = L.fjp_score_batch_f32(eng._h, m._h, hXf.data_ptr(), hP.data_ptr(), n, F, hS.data_ptr(), hT.data_ptr(), None) if f32 else \
             L.fjp_score_batch(eng._h, m._h, hXd.data_ptr(), hP.data_ptr(), None, n, F, hS.data_ptr(), hT.data_ptr(), None)
        assert rc == 0
    os.environ.pop("FJP_TRACE_CHUNKS")
