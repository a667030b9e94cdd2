"""Device-resident throughput of a forest of deep trees (unpruned scikit-learn style), development aid."""
import ctypes as C, os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np, torch
from flink_jpmml_b200 import capi, pmmlgen as g
n, F = 1 << 20, 32
T, D = int(os.environ.get("T", 50)), int(os.environ.get("D", 13))
eng = capi.Engine(0)
for method in ("average", "majorityVote"):
    pm = g.rf_pmml(T=T, depth=D, F=F, n_labels=3, method=method, seed=9)[0]
    m = eng.compile(pm); print(m.describe[:170])
    X, pres, pb = g.events(n, F, 1, 0.0)
    dX = torch.from_numpy(X).cuda()
    dS = torch.empty(n, dtype=torch.float64, device="cuda"); dT = torch.empty(n, dtype=torch.uint8, device="cuda")
    st = torch.cuda.Stream(); torch.cuda.set_stream(st)
    run = lambda: capi.lib().fjp_score_device(eng._h, m._h, dX.data_ptr(), None, None, n, F, dS.data_ptr(), dT.data_ptr(), None, C.c_void_p(st.cuda_stream))
    for _ in range(2): assert run() == 0
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record(); [run() for _ in range(3)]; e1.record(); torch.cuda.synchronize()
    ms = e0.elapsed_time(e1) / 3
    print(f"{method} T={T} depth={D}: {ms:.2f} ms, {n / ms * 1e3:.3e} events/s")
