"""A/B of the two forest paths on C4 (development aid): classic one-kernel walk vs rank-key pre-pass + compact walk.
Compares every score bit for bit and times both, device-resident.  usage: compact_ab.py [n] [CE,CTI,CG ...]"""
import ctypes as C
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
import torch
from flink_jpmml_b200 import capi, pmmlgen as g

n = int(sys.argv[1]) if len(sys.argv) > 1 else 1 << 22
eng = capi.Engine(0)
pm, arr = g.gbt_pmml()
m = eng.compile(pm)
print(m.describe)
X, pres, pb = g.events(n, 100, 1, 0.1)
dX = torch.from_numpy(X).cuda(); dP = torch.from_numpy(pb.view(np.int32)).cuda()
stream = torch.cuda.Stream(); torch.cuda.set_stream(stream); st = stream.cuda_stream; assert st != 0
L = capi.lib()

def run(dS, dT):
    rc = L.fjp_score_device(eng._h, m._h, dX.data_ptr(), dP.data_ptr(), None, n, 100, dS.data_ptr(), dT.data_ptr(), None, C.c_void_p(st))
    assert rc == 0, L.fjp_last_error()

def timed(label, env):
    for k in ("FJP_FOREST_COMPACT", "FJP_FOREST_CE", "FJP_FOREST_CTI", "FJP_FOREST_CG", "FJP_FOREST_CPASS"): os.environ.pop(k, None)
    os.environ.update(env)
    dS = torch.full((n,), -1.0, dtype=torch.float64, device="cuda"); dT = torch.full((n,), 9, dtype=torch.uint8, device="cuda")
    for _ in range(3): run(dS, dT)
    torch.cuda.synchronize()
    ev0, ev1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    ev0.record()
    for _ in range(5): run(dS, dT)
    ev1.record(); torch.cuda.synchronize()
    ms = ev0.elapsed_time(ev1) / 5
    print(f"{label:28s} ms={ms:8.3f} events/s={n / ms * 1e3:.3e}", flush=True)
    return dS, dT

base_S, base_T = timed("classic", {"FJP_FOREST_COMPACT": "0"})
ref = g.gbt_numpy(arr, X[:20000], pres[:20000])
print("classic matches numpy:", np.array_equal(base_S[:20000].cpu().numpy(), ref))
variants = sys.argv[2:] or ["640,8,0", "768,8,0", "1024,4,0", "640,4,0"]
for v in variants:
    ce, cti, cg, *rest = v.split(",")
    env = {"FJP_FOREST_COMPACT": "1", "FJP_FOREST_CE": ce, "FJP_FOREST_CTI": cti}
    if cg != "0": env["FJP_FOREST_CG"] = cg
    if rest: env["FJP_FOREST_CPASS"] = rest[0]
    S, T = timed(f"compact E={ce} TI={cti} G={cg}", env)
    same = bool(torch.equal(S.view(torch.int64), base_S.view(torch.int64))) and bool(torch.equal(T, base_T))
    print("   bit-identical to classic:", same, flush=True)
    if not same:
        bad = (S.view(torch.int64) != base_S.view(torch.int64)).nonzero().flatten()
        print("   mismatches:", bad.numel(), "first rows:", bad[:8].tolist(), S[bad[:4]].tolist(), base_S[bad[:4]].tolist())
