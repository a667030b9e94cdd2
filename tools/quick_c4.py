"""Quick device-resident timing of the C4 forest kernel (development aid; bench.py is the contract)."""
import ctypes as C
import sys, os, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
import torch
from flink_jpmml_b200 import capi, pmmlgen as g

n = int(sys.argv[1]) if len(sys.argv) > 1 else 1 << 20
eng = capi.Engine(0)
pm, arr = g.gbt_pmml()
m = eng.compile(pm)
print(m.describe)
X, pres, pb = g.events(n, 100, 1, 0.1)
dX = torch.from_numpy(X).cuda(); dP = torch.from_numpy(pb.view(np.int32)).cuda()
dS = torch.empty(n, dtype=torch.float64, device="cuda"); dT = torch.empty(n, dtype=torch.uint8, device="cuda")
stream = torch.cuda.Stream(); torch.cuda.set_stream(stream); st = stream.cuda_stream; assert st != 0
L = capi.lib()
def run():
    rc = L.fjp_score_device(eng._h, m._h, dX.data_ptr(), dP.data_ptr(), None, n, 100, dS.data_ptr(), dT.data_ptr(), None, C.c_void_p(st))
    assert rc == 0, L.fjp_last_error()
for _ in range(3): run()
torch.cuda.synchronize()
ev0, ev1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
reps = 5
ev0.record()
for _ in range(reps): run()
ev1.record(); torch.cuda.synchronize()
ms = ev0.elapsed_time(ev1) / reps
print(f"n={n} ms={ms:.3f} events/s={n/ms*1e3:.3e} gather-eq GB/s={n*32000/ms*1e3/1e9:.1f}")
ref = g.gbt_numpy(arr, X[:20000], pres[:20000])
print("match numpy:", np.array_equal(dS[:20000].cpu().numpy(), ref), "status0:", int((dT == 0).sum().item()) == n)
t = time.time(); r = eng.score(m, X, pb); dt = time.time() - t
print(f"e2e (pageable host buffers) events/s={n/dt:.3e}")
