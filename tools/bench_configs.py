"""Device-resident throughput of every BASELINE.json config (C1..C5) + roofline fractions.
bench.py stays the contract for the headline (C4); this fills BASELINE.md §6 for the other configs.
usage: python tools/bench_configs.py [--cpu]  -> one JSON line per config"""
import argparse, ctypes as C, json, os, sys, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
import torch
from flink_jpmml_b200 import capi, pmmlgen as g

ap = argparse.ArgumentParser(); ap.add_argument("--cpu", action="store_true"); ap.add_argument("--only", default="")
args = ap.parse_args()
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
micro = json.load(open(os.path.join(ROOT, "profiles", "microbench_r01.json")))
peaks = json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json"))) if os.path.exists(os.path.join(ROOT, "MEASURED_PEAKS.json")) else {"hbm_gbs": 6650.0}
eng = capi.Engine(0); L = capi.lib()
stream = torch.cuda.Stream(); torch.cuda.set_stream(stream)

def time_model(pm, F, n, p_missing, reps=5):
    m = eng.compile(pm)
    gen = torch.Generator(device="cuda"); gen.manual_seed(1)
    X = torch.randn(n, F, generator=gen, device="cuda", dtype=torch.float64)
    P = None
    if p_missing > 0:
        Xh, pres, pb = g.events(1 << 16, F, 3, p_missing)      # present pattern tiled from a 65k-row sample
        P = torch.from_numpy(pb.view(np.int32)).cuda().repeat((n + (1 << 16) - 1) // (1 << 16), 1)[:n].contiguous()
    S = torch.empty(n, dtype=torch.float64, device="cuda"); T = torch.empty(n, dtype=torch.uint8, device="cuda")
    def run():
        rc = L.fjp_score_device(eng._h, m._h, X.data_ptr(), P.data_ptr() if P is not None else None, None, n, F, S.data_ptr(), T.data_ptr(), None, C.c_void_p(stream.cuda_stream))
        assert rc == 0, L.fjp_last_error()
    for _ in range(3): run()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record(stream)
    for _ in range(reps): run()
    e1.record(stream); torch.cuda.synchronize()
    ms = e0.elapsed_time(e1) / reps
    ok = int((T == 0).sum().item())
    return m, n / ms * 1e3, ms, ok, X, P

def cpu_rate(pm, X, P, n=4096):
    from oracle import oracle
    om = oracle.OracleModel(pm)
    xs = X[:n].cpu().numpy(); ps = P[:n].cpu().numpy().view(np.uint32) if P is not None else None
    t = time.perf_counter(); om.predict(xs[:512], ps[:512] if ps is not None else None, nthreads=1); r1 = 512 / (time.perf_counter() - t)
    cores = os.cpu_count()
    t = time.perf_counter(); om.predict(xs, ps, nthreads=cores); rn = n / (time.perf_counter() - t)
    return {"single_thread": r1, "all_cores": rn, "cores": cores}

configs = []
iris = open(os.path.join(ROOT, "tests", "golden", "iris_kmeans43.pmml"), "rb").read()
configs.append(("C1 iris kmeans k=4 F=4", iris, 4, 1 << 22, 0.0, {"flop": 48, "hbm": 40}))
configs.append(("C2 kmeans k=64 F=128", g.config_c2()[0], 128, 1 << 22, 0.0, {"dp_instr": 24576, "hbm": 1032}))
configs.append(("C2 kmeans k=64 F=128 5% missing", g.config_c2()[0], 128, 1 << 22, 0.05, {"dp_instr": 24576, "hbm": 1032}))
configs.append(("C3 multinomial logistic 1000x20", g.regression_pmml()[0], 1000, 1 << 20, 0.0, {"dp_instr": 40000, "hbm": 8008}))
configs.append(("C4 GBT 500x8 F=100 10% missing", g.gbt_pmml()[0], 100, 1 << 22, 0.10, {"gather": 32000, "hbm": 824}))
configs.append(("C4 GBT float fields (XGBoost style)", g.gbt_pmml(dtype="float")[0], 100, 1 << 22, 0.10, {"gather": 32000, "hbm": 824}))
g.FLOAT_THRESHOLDS = False
configs.append(("C4 GBT double thresholds (scikit-learn / LightGBM style)", g.gbt_pmml()[0], 100, 1 << 22, 0.10, {"gather": 32000, "hbm": 824}))
g.FLOAT_THRESHOLDS = True
configs.append(("C5 RF 200x8 F=64 majorityVote", g.rf_pmml()[0], 64, 1 << 22, 0.0, {"gather": 12800, "hbm": 528}))
configs.append(("C5 MLP 64-256-256-8", g.nn_pmml()[0], 64, 1 << 21, 0.0, {"dp_instr": 167936, "hbm": 520}))
# beyond BASELINE.json's configs: the two other common shapes of exported forests
configs.append(("RF 200x8 F=64 average of leaf distributions (scikit-learn style)", g.rf_pmml(method="average")[0], 64, 1 << 22, 0.0, {"gather": 12800, "hbm": 528}))
configs.append(("GBT 300 unbalanced trees, max depth 14, F=50 (leaf-wise growth)", g.gbt_pruned_pmml(T=300, max_depth=14, F=50, seed=3, p_leaf=0.3), 50, 1 << 21, 0.10, {"hbm": 416}))
for name, pm, F, n, pmiss, work in configs:
    if args.only and args.only not in name: continue
    fp0 = eng.fast_path_counts
    m, rate, ms, ok, X, P = time_model(pm, F, n, pmiss)
    fp1 = eng.fast_path_counts
    out = {"config": name, "events_per_s": rate, "ms": ms, "n": n, "rows_ok": ok, "model": m.describe,
           "hbm_frac": rate * work["hbm"] / 1e9 / peaks["hbm_gbs"]}
    if "gather" in work:
        out["gather_gbs"] = rate * work["gather"] / 1e9
        out["frac_l2_gather"] = out["gather_gbs"] / micro["l2_gather_8B_2MB_GBs"]; out["frac_smem_gather"] = out["gather_gbs"] / micro["smem_gather_8B_random_GBs"]
    if "dp_instr" in work:   # unfused double mul+add: 2 flop per (mul, add) pair; compare with the measured unfused FP64 rate
        out["fp64_unfused_tflops"] = rate * work["dp_instr"] / 1e12
        out["frac_fp64_unfused"] = out["fp64_unfused_tflops"] / micro["fp64_mul_add_unfused_TFLOPs"]
    if fp1[2] > fp0[2]: out["fma_path_rescored_fraction"] = (fp1[3] - fp0[3]) / (fp1[2] - fp0[2])
    if fp1[0] > fp0[0]: out["tensor_path_rescored_fraction"] = (fp1[1] - fp0[1]) / (fp1[0] - fp0[0])
    if args.cpu: out["cpu_oracle"] = cpu_rate(pm, X, P)
    print(json.dumps(out), flush=True)
    del X, P; torch.cuda.empty_cache()
