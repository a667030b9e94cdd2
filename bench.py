#!/usr/bin/env python3
"""bench.py -- headline benchmark: scored events/sec for the 500-tree GBT PMML (BASELINE.json config C4).

  python bench.py [--gpus N] [--steps K] [--warmup W] [--impl ours|reference]
  (N > 1: python -m torch.distributed.run --nnodes=1 --nproc-per-node N ... bench.py --gpus N ...)

A "step" is one pass of the hot path (PmmlModel.predict for every event of a micro-batch, reference:
flink-jpmml-scala/src/main/scala/io/radicalbit/flink/pmml/scala/api/PmmlModel.scala:109-119) over one batch of 2^22
synthetic events per GPU.  Events are sharded across ranks; there is no collective on the scoring path (weak scaling).
Rank 0 prints ONE JSON line.

  value      whole-job events/s, inputs resident in HBM, timed with CUDA events on the launching stream
  e2e        same metric through the C ABI with pinned HOST buffers (H2D + kernel + D2H inside the timed region)
  roofline   HBM roofline of the forest kernel (algorithmic 824 B/event) + the gather rooflines
  cpu_baseline  the oracle port (oracle/oracle.cc) on the box's host cores, bounded sample (reported, not the target)
"""
import argparse
import ctypes as C
import json
import os
import statistics
import subprocess
import sys
import threading
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

F, TREES, DEPTH, P_MISSING = 100, 500, 8, 0.10
EVENTS_PER_GPU = 1 << 22          # N = 2^22 per GPU
E2E_EVENTS_PER_GPU = 1 << 21      # per step through the host-buffer entry point
ALGO_HBM_BYTES = 824              # 100*8 features + 16 present bits + 8 score   
ALGO_GATHER_BYTES = 32000         # 500 trees * 8 levels * 8 B node record       
SEED = 20261106 + 4
WORKLOAD = "C4: MiningModel GBT sum, 500 complete trees depth 8, 100 double features, 10% missing"


def peaks():
    p = {"hbm_gbs": 6650.0, "source": "fallback (B200_PROFILING.md)"}
    try:
        with open(os.path.join(ROOT, "MEASURED_PEAKS.json")) as f:
            m = json.load(f)
        p = {"hbm_gbs": float(m["hbm_gbs"]), "source": "measured (MEASURED_PEAKS.json)"}
    except Exception:
        pass
    p["micro"] = {}
    for name in ("microbench_r01.json", "microbench_r02.json"):     # later rounds override / extend earlier figures
        try:
            with open(os.path.join(ROOT, "profiles", name)) as f:
                p["micro"].update(json.load(f))
        except Exception:
            pass
    try:
        p["bf16_tflops_sustained"] = float(m["bf16_tflops_sustained"]); p["bf16_tflops"] = float(m["bf16_tflops"])
    except Exception:
        p["bf16_tflops_sustained"], p["bf16_tflops"] = 1400.0, 1590.0
    return p


class ClockSampler:
    """nvidia-smi clocks / throttle reasons DURING the timed region (B200_PROFILING.md recipe)."""
    Q = "clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap"

    def __init__(self, index):
        self.index, self.rows, self.proc = index, [], None

    def start(self):
        try:
            self.proc = subprocess.Popen(["nvidia-smi", "-i", str(self.index), f"--query-gpu={self.Q}", "--format=csv,noheader,nounits", "-lms", "100"],
                                         stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
            self.t = threading.Thread(target=self._read, daemon=True)
            self.t.start()
        except Exception:
            self.proc = None

    def _read(self):
        for line in self.proc.stdout:
            self.rows.append([c.strip() for c in line.split(",")])

    def stop(self):
        if not self.proc:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        self.proc.terminate()
        try:
            self.proc.wait(timeout=5)
        except Exception:
            self.proc.kill()
        sm = [float(r[0]) for r in self.rows if r and r[0].replace(".", "").isdigit()]
        mx = [float(r[1]) for r in self.rows if len(r) > 1 and r[1].replace(".", "").isdigit()]
        pw = [float(r[2]) for r in self.rows if len(r) > 2 and r[2].replace(".", "").isdigit()]
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        reasons = [n for i, n in enumerate(names) if any(len(r) > 3 + i and r[3 + i].lower().startswith("active") for r in self.rows)]
        return {"sm_mhz": statistics.median(sm) if sm else None, "sm_max_mhz": max(mx) if mx else None,
                "power_w_max": max(pw) if pw else None, "samples": len(sm), "reasons": reasons}


def make_model_bytes():
    from flink_jpmml_b200 import pmmlgen
    return pmmlgen.gbt_pmml(T=TREES, depth=DEPTH, F=F, seed=SEED)


def device_events(torch, n, seed, device):
    """x ~ N(0,1)^100 doubles, each slot missing w.p. 0.10 (present-bit cleared), generated on the device."""
    g = torch.Generator(device=device)
    g.manual_seed(seed)
    X = torch.randn(n, F, generator=g, device=device, dtype=torch.float64)
    W = (F + 31) // 32
    P = torch.empty(n, W, dtype=torch.int32, device=device)
    shifts = torch.arange(32, device=device, dtype=torch.int64)
    step = 1 << 18
    for lo in range(0, n, step):
        hi = min(n, lo + step)
        bits = torch.zeros(hi - lo, W * 32, dtype=torch.int64, device=device)
        bits[:, :F] = (torch.rand(hi - lo, F, generator=g, device=device) >= P_MISSING).to(torch.int64)
        w = (bits.view(hi - lo, W, 32) << shifts).sum(-1)
        P[lo:hi] = torch.where(w >= 2 ** 31, w - 2 ** 32, w).to(torch.int32)
    return X, P


# ---------------------------------------------------------------------------------------------------------------------
# The other named shapes of BASELINE.json: device-resident events/s, the roofline that binds each
# per section 8(d), and parity of a sample against the oracle.  C4 is the headline above; these ride in `configs`.
def other_configs():
    """name -> (PMML bytes factory, F, events per GPU, p_missing, algorithmic work per event)"""
    from flink_jpmml_b200 import pmmlgen as g
    iris = lambda: open(os.path.join(ROOT, "tests", "golden", "iris_kmeans43.pmml"), "rb").read()
    return [
        ("C1", "Iris k-means k=4 F=4 (A/kmeans.xml)", iris, 4, 1 << 22, 0.0, {"flop": 48, "hbm": 40, "bound": "latency"}),
        ("C2", "k-means k=64 F=128, dense", lambda: g.config_c2()[0], 128, 1 << 22, 0.0, {"flop": 24576, "hbm": 1032, "bound": "fp64"}),
        ("C2_missing5", "k-means k=64 F=128, 5% missing", lambda: g.config_c2()[0], 128, 1 << 22, 0.05, {"flop": 24576, "hbm": 1032, "bound": "fp64"}),
        ("C3", "multinomial logistic 1000 features x 20 classes", lambda: g.regression_pmml()[0], 1000, 1 << 20, 0.0, {"flop": 40000, "hbm": 8008, "bound": "hbm"}),
        ("C5_RF", "random forest 200 trees depth 8 F=64 majorityVote", lambda: g.rf_pmml()[0], 64, 1 << 22, 0.0, {"gather": 12800, "hbm": 528, "bound": "gather", "wavefronts": "c5rf_step_smem_wavefronts_per_event"}),
        ("C5_MLP", "NeuralNetwork 64-256-256-8 rectifier + softmax", lambda: g.nn_pmml()[0], 64, 1 << 21, 0.0, {"flop": 167936, "hbm": 520, "bound": "tensor"}),
    ]


def measure_config(torch, np, eng, L, stream, dev, spec, seed, reps, want_parity):
    from flink_jpmml_b200 import pmmlgen as g
    key, desc, make, F_, n, pmiss, work = spec
    pm = make()
    m = eng.compile(pm)
    gen = torch.Generator(device=dev); gen.manual_seed(seed)
    if key == "C1":   # Iris-shaped events: U[0.2, 6.0] floored to one decimal (E/sources/IrisSource.scala:32-43)
        X = (torch.floor((torch.rand(n, F_, generator=gen, device=dev, dtype=torch.float64) * 5.8 + 0.2) * 10.0) / 10.0).contiguous()
    else:
        X = torch.randn(n, F_, generator=gen, device=dev, dtype=torch.float64)
    P = None
    if pmiss > 0:
        _, _, pb = g.events(1 << 16, F_, seed, pmiss)       # present pattern tiled from a 65k-row sample
        P = torch.from_numpy(pb.view(np.int32)).to(dev).repeat((n + (1 << 16) - 1) // (1 << 16), 1)[:n].contiguous()
    S = torch.empty(n, dtype=torch.float64, device=dev); T = torch.empty(n, dtype=torch.uint8, device=dev)

    def run():
        rc = L.fjp_score_device(eng._h, m._h, X.data_ptr(), P.data_ptr() if P is not None else None, None, n, F_, S.data_ptr(), T.data_ptr(), None, C.c_void_p(stream.cuda_stream))
        if rc != 0:
            raise RuntimeError(L.fjp_last_error().decode())
    for _ in range(3):
        run()
    torch.cuda.synchronize(dev)
    l0 = eng.launch_count
    t0 = eng.fast_path_counts
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    with torch.cuda.stream(stream):
        e0.record(stream)
        for _ in range(reps):
            run()
        e1.record(stream)
    torch.cuda.synchronize(dev)
    ms = e0.elapsed_time(e1) / reps
    t1 = eng.fast_path_counts
    out = {"workload": desc, "events_per_gpu_per_step": n, "ms_per_step": ms, "rows_scored_ok": int((T == 0).sum().item()),
           "gpu_launches_per_step": (eng.launch_count - l0) // reps, "model": m.describe, "work": work}
    if t1[0] > t0[0]:   # the tensor-core path took the batch: how many events its margin test handed back to the FP64 kernels
        out["tensor_path"] = {"events": (t1[0] - t0[0]) // reps, "rescored_fp64": (t1[1] - t0[1]) // reps,
                              "rescored_fraction": (t1[1] - t0[1]) / max(1, t1[0] - t0[0]),
                              "arithmetic": "tcgen05.mma kind::i8 on 4 base-256 digit planes per operand, exact int32 accumulation in TMEM; labels decided only outside the proven error bound"}
    if t1[2] > t0[2]:   # the fused-multiply-add path of the dense tile kernel took the batch
        out["fma_path"] = {"events": (t1[2] - t0[2]) // reps, "rescored_unfused": (t1[3] - t0[3]) // reps,
                           "rescored_fraction": (t1[3] - t0[3]) / max(1, t1[2] - t0[2]),
                           "arithmetic": "FP64 fused multiply-add; labels decided only outside the rounding bound between the two evaluation orders, the rest re-scored unfused in JPMML's order"}
    if want_parity:
        from oracle import oracle   # checker only: a 2048-event sample of this very batch against the CPU restatement
        ns = 2048
        xs = X[:ns].cpu().numpy(); ps = P[:ns].cpu().numpy().view(np.uint32) if P is not None else None
        o = oracle.OracleModel(pm).predict(xs, ps, nthreads=os.cpu_count() or 1)
        same = np.array_equal(o["status"], T[:ns].cpu().numpy()) and np.array_equal(o["score"], S[:ns].cpu().numpy(), equal_nan=True)
        out["parity_on_sample"] = "bit-exact (2048 events: status + score)" if same else "MISMATCH"
    del X, P, S, T
    torch.cuda.empty_cache()
    m.release()
    return out


def config_rooflines(rate, work, pk):
    """Fractions of the resource SURVEY.md section 8(d) names as binding for the shape (per GPU rate)."""
    micro = pk.get("micro", {}); r = {"hbm_frac": rate * work["hbm"] / 1e9 / pk["hbm_gbs"]}
    if "gather" in work:
        gbs = rate * work["gather"] / 1e9
        r.update({"gather_gbs": gbs, "frac_of_l2_gather_peak": gbs / micro["l2_gather_8B_2MB_GBs"] if micro.get("l2_gather_8B_2MB_GBs") else None,
                  "frac_of_smem_gather_peak": gbs / micro["smem_gather_8B_random_GBs"] if micro.get("smem_gather_8B_random_GBs") else None})
    if "flop" in work:
        tf = rate * work["flop"] / 1e12
        r.update({"algorithmic_tflops": tf, "frac_of_fp64_fma_peak": tf / micro["fp64_fma_TFLOPs"] if micro.get("fp64_fma_TFLOPs") else None})
        if work["bound"] == "tensor":
            # the path computes in int8 (dense int8 rate = 2 x the bf16 rate on this chip; MEASURED_PEAKS.json holds the bf16 figure).
            # Two readings: the algorithmic FP64-equivalent work against the bf16 peak, and the int8 operations the tensor pipe
            # really executes (10 digit-pair products per multiply-add at S = 4) against 2 x that peak.
            r["frac_of_bf16_tensor_sustained"] = tf / pk["bf16_tflops_sustained"]
            r["int8_tensor_tops_executed"] = tf * 10.0
            r["frac_of_int8_tensor_peak"] = tf * 10.0 / (2.0 * pk["bf16_tflops_sustained"])
            r["int8_peak_note"] = "2 x bf16_tflops_sustained of MEASURED_PEAKS.json (nominal dense int8 = 2 x bf16 on B200)"
    b = work["bound"]
    if micro.get(work.get("wavefronts", "")):   # tree walks gather from shared memory: the pipe issues one wavefront per SM and cycle
        wf = micro[work["wavefronts"]]; peak = 148 * (pk.get("sm_mhz_sampled") or 1965.0) * 1e6 / 1e9
        r.update({"smem_wavefronts_per_event": wf, "smem_wavefronts_source": micro.get(work["wavefronts"].replace("_per_event", "_note")),
                  "smem_pipe_gwavefronts": rate * wf / 1e9, "smem_pipe_peak_gwavefronts": peak, "frac_of_smem_pipe": rate * wf / 1e9 / peak})
        b = "smem_pipe"
    r["binding"] = b
    r["frac"] = {"hbm": r["hbm_frac"], "fp64": r.get("frac_of_fp64_fma_peak"), "gather": r.get("frac_of_smem_gather_peak"), "smem_pipe": r.get("frac_of_smem_pipe"),
                 "tensor": r.get("frac_of_int8_tensor_peak"), "latency": None}[b]
    return r


def run_reference(args):
    """--impl reference: the reference's CPU path.  JPMML/Flink cannot run here (no JVM on the box), so this is the
    oracle port (oracle/oracle.cc: literal per-event interpreter) on all host cores, same metric / config."""
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    import numpy as np
    from flink_jpmml_b200 import pmmlgen
    from oracle import oracle
    pm, _ = make_model_bytes()
    om = oracle.OracleModel(pm)
    cores = os.cpu_count() or 1
    # calibrate: ~3 s of CPU work per step
    X, pres, pb = pmmlgen.events(4096, F, SEED, P_MISSING)
    t = time.perf_counter(); om.predict(X, pb, nthreads=cores); rate = 4096 / (time.perf_counter() - t)
    n = int(max(4096, min(1 << 20, rate * 3.0)))
    X, pres, pb = pmmlgen.events(n, F, SEED + 1, P_MISSING)
    for _ in range(args.warmup):
        om.predict(X[:4096], pb[:4096], nthreads=cores)
    t0 = time.perf_counter()
    for _ in range(args.steps):
        om.predict(X, pb, nthreads=cores)
    dt = time.perf_counter() - t0
    v = n * args.steps / dt
    line = {"impl": "reference", "metric": "scored events/sec", "value": v, "unit": "events/s", "n_gpus": args.gpus, "steps": args.steps,
            "warmup": args.warmup, "ms_per_step": dt / args.steps * 1e3, "higher_is_better": True, "scaling": "weak", "vs_baseline": None,
            "dtype": "f64", "data": "synthetic", "config": {"workload": WORKLOAD, "events_per_step": n},
            "cpu_baseline": {"value": v, "unit": "events/s", "cores": cores, "kind": "port",
                             "sample": f"{n} events/step x {args.steps} steps, oracle/oracle.cc with {cores} threads (JPMML/Flink not runnable: no JVM)"},
            "e2e": {"value": v, "unit": "events/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0}}
    print(json.dumps(line), flush=True)


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=10)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--events", type=int, default=EVENTS_PER_GPU, help="events per GPU per step (default 2^22)")
    ap.add_argument("--no-latency", action="store_true")
    ap.add_argument("--no-cpu", action="store_true")
    ap.add_argument("--no-configs", action="store_true", help="skip the other BASELINE.json shapes (C1, C2, C3, C5)")
    args = ap.parse_args()
    args.warmup = max(args.warmup, 3) if args.impl == "ours" else args.warmup
    if args.impl == "reference":
        return run_reference(args)

    import numpy as np
    import torch
    from flink_jpmml_b200 import capi

    from flink_jpmml_b200 import dist as D
    rank, world, local = D.env_rank()
    torch.cuda.set_device(local)
    numa_node = D.bind_to_gpu_numa_node(local) if world > 1 else None   # pinned buffers next to this rank's GPU
    dev = torch.device("cuda", local)
    dist = D.init("nccl", dev)   # plumbing only: barrier + max-over-ranks of the timings; no collective while scoring
    L = capi.lib()
    eng = capi.Engine(local)
    pm, arrays = make_model_bytes()
    model = eng.compile(pm)

    n = args.events
    X, P = device_events(torch, n, SEED + 1000 * rank, dev)
    score = torch.empty(n, dtype=torch.float64, device=dev); status = torch.empty(n, dtype=torch.uint8, device=dev)
    stream = torch.cuda.Stream(device=dev)

    def step_device():
        rc = L.fjp_score_device(eng._h, model._h, X.data_ptr(), P.data_ptr(), None, n, F, score.data_ptr(), status.data_ptr(), None, C.c_void_p(stream.cuda_stream))
        if rc != 0:
            raise RuntimeError(L.fjp_last_error().decode())

    def barrier():
        if dist:
            dist.barrier()
        torch.cuda.synchronize(dev)

    # ------------------------------------------------------------------ device-resident throughput ("value")
    for _ in range(args.warmup):
        step_device()
    barrier()
    sampler = ClockSampler(local); sampler.start(); time.sleep(0.3)
    launches0 = eng.launch_count
    ev0, ev1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    with torch.cuda.stream(stream):
        ev0.record(stream)
        for _ in range(args.steps):
            step_device()
        ev1.record(stream)
    barrier()
    ms_dev = ev0.elapsed_time(ev1)
    launches = eng.launch_count - launches0
    ok_rows = int((status == 0).sum().item())

    # ------------------------------------------------------------------ end to end through the host-buffer entry point
    ne = min(E2E_EVENTS_PER_GPU, n); W = (F + 31) // 32
    hX = torch.empty(ne, F, dtype=torch.float64).pin_memory(); hP = torch.empty(ne, W, dtype=torch.int32).pin_memory()
    hX.copy_(X[:ne]); hP.copy_(P[:ne])
    hS = torch.empty(ne, dtype=torch.float64).pin_memory(); hT = torch.empty(ne, dtype=torch.uint8).pin_memory()

    def step_e2e():
        rc = L.fjp_score_batch(eng._h, model._h, hX.data_ptr(), hP.data_ptr(), None, ne, F, hS.data_ptr(), hT.data_ptr(), None)
        if rc != 0:
            raise RuntimeError(L.fjp_last_error().decode())

    def time_e2e(step):
        for _ in range(2):
            step()
        barrier()
        t0 = time.perf_counter()
        for _ in range(e2e_steps):
            step()
        torch.cuda.synchronize(dev)
        return time.perf_counter() - t0

    e2e_steps = max(3, min(args.steps, 10))
    t_e2e_f64 = time_e2e(step_e2e)
    e2e_match = bool(torch.equal(hS[:4096], score[:4096].cpu()))
    # float rows: the model's tests see every feature through ONE float (fjp_model_float_row_modes), so the binding can ship
    # 4-byte values rounded the way the model prescribes -- half the PCIe bytes, bit-identical scores
    modes = eng.float_row_modes(model)
    t_e2e, e2e_rows, e2e_f32_match = t_e2e_f64, "double", None
    if modes is not None:
        md = torch.from_numpy(modes.astype(np.int64)).to(dev)[None, :]
        Xd = X[:ne]; f = Xd.to(torch.float32); back = f.to(torch.float64)
        f = torch.where((back < Xd) & (md == 1), torch.nextafter(f, torch.full_like(f, float("inf"))), f)
        f = torch.where((back > Xd) & (md == 2), torch.nextafter(f, torch.full_like(f, float("-inf"))), f)
        hXf = torch.empty(ne, F, dtype=torch.float32).pin_memory(); hXf.copy_(f)
        del f, back, Xd
        hS2 = torch.empty(ne, dtype=torch.float64).pin_memory()

        def step_e2e_f32():
            rc = L.fjp_score_batch_f32(eng._h, model._h, hXf.data_ptr(), hP.data_ptr(), ne, F, hS2.data_ptr(), hT.data_ptr(), None)
            if rc != 0:
                raise RuntimeError(L.fjp_last_error().decode())
        t_e2e = time_e2e(step_e2e_f32); e2e_rows = "float"
        e2e_f32_match = bool(torch.equal(hS2, hS))          # every score of the float rows against the double rows
    # what THIS box copies from the very buffers the e2e used (pinned -> device, chunk-sized pieces, nothing else running): the
    # ceiling of the e2e figure; boxes of the pool differ by ~10 % here
    h2d_live = None
    try:
        src = hXf if modes is not None else hX
        rows_c = min(1 << 18, ne); dbuf = torch.empty_like(src[:rows_c], device=dev); torch.cuda.synchronize(dev)
        best = 0.0
        for _ in range(3):
            t0 = time.perf_counter()
            for lo in range(0, ne - rows_c + 1, rows_c):
                dbuf.copy_(src[lo:lo + rows_c], non_blocking=True)
            torch.cuda.synchronize(dev)
            best = max(best, (ne // rows_c) * rows_c * src.shape[1] * src.element_size() / (time.perf_counter() - t0) / 1e9)
        h2d_live = best; del dbuf
    except Exception:      # noqa: BLE001
        h2d_live = None
    clocks = sampler.stop()

    # ------------------------------------------------------------------ the other named shapes (all at N = 1; C5 halves at N > 1)
    cfg_out = {}
    if not args.no_configs:
        del X, P
        torch.cuda.empty_cache()
        for spec in other_configs():
            if world > 1 and spec[0] not in ("C5_RF", "C5_MLP"):
                continue
            cfg_out[spec[0]] = measure_config(torch, np, eng, L, stream, dev, spec, SEED + 17 + 1000 * rank, 5, rank == 0 and world == 1 and not args.no_cpu)
    cfg_keys = sorted(cfg_out)

    # ------------------------------------------------------------------ max over ranks
    maxed = D.max_over_ranks(dist, [ms_dev, t_e2e * 1e3, t_e2e_f64 * 1e3] + [cfg_out[k]["ms_per_step"] for k in cfg_keys], dev)
    ms_dev_max, ms_e2e_max, ms_e2e_f64_max = maxed[0], maxed[1], maxed[2]
    for k, v in zip(cfg_keys, maxed[3:]):
        cfg_out[k]["ms_per_step"] = v

    # ------------------------------------------------------------------ latency sweep + CPU baseline: rank 0 only
    latency, cpu, replicate = None, None, None
    # ------------------------------------------------------------------ model replication by peer copy (N >= 2, rank 0, after the timed regions)
    # The control stream reaches every subtask (S/package.scala:63-65); here a compiled model is copied GPU -> GPU over NVLink
    # (cudaMemcpyPeerAsync, no kernel reads a peer's memory) instead of being parsed and compiled once per GPU.
    if rank == 0 and world > 1 and torch.cuda.device_count() > 1:
        try:
            peer = (local + 1) % torch.cuda.device_count()
            eng2 = capi.Engine(peer)
            t1 = time.perf_counter(); m2 = eng2.replicate(model); t_rep = time.perf_counter() - t1
            t1 = time.perf_counter(); m3 = eng2.compile(pm); t_cmp = time.perf_counter() - t1
            nr = 1 << 16
            xs = hX[:nr].numpy(); ps = hP[:nr].numpy().view(np.uint32)
            ra = eng.score(model, xs, ps); rb = eng2.score(m2, xs, ps)
            replicate = {"bytes": model.device_bytes, "ms": t_rep * 1e3, "compile_instead_ms": t_cmp * 1e3, "from_gpu": local, "to_gpu": peer, "events_checked": nr,
                         "bit_identical": bool(np.array_equal(ra["score"], rb["score"]) and np.array_equal(ra["status"], rb["status"]))}
            m2.release(); m3.release(); eng2.close()
            torch.cuda.set_device(local)
        except Exception as ex:          # noqa: BLE001
            replicate = {"error": repr(ex)}
    if rank == 0 and not args.no_latency:
        latency = {}
        ticket = C.c_int()
        for b in (1, 64, 1024, 16384, 262144):
            reps = 1000 if b <= 16384 else 200
            ts = []
            for _ in range(reps):
                t1 = time.perf_counter()
                L.fjp_submit(eng._h, model._h, hX.data_ptr(), hP.data_ptr(), None, b, F, hS.data_ptr(), hT.data_ptr(), None, C.byref(ticket))
                L.fjp_wait(eng._h, ticket.value)
                ts.append((time.perf_counter() - t1) * 1e3)
            ts.sort()
            latency[str(b)] = {"p50_ms": round(ts[len(ts) // 2], 4), "p99_ms": round(ts[int(len(ts) * 0.99) - 1], 4)}
    if rank == 0 and world == 1 and not args.no_cpu:
        from oracle import oracle   # cpu_baseline leg: the one place bench.py may execute the oracle
        cores = os.cpu_count() or 1
        om = oracle.OracleModel(pm)
        xs = hX[:4096].numpy(); ps = hP[:4096].numpy().view(np.uint32)
        t1 = time.perf_counter(); om.predict(xs, ps, nthreads=cores); rate = 4096 / (time.perf_counter() - t1)
        ns = int(max(4096, min(ne, rate * 12.0)))
        xs = hX[:ns].numpy(); ps = hP[:ns].numpy().view(np.uint32)
        t1 = time.perf_counter(); o = om.predict(xs, ps, nthreads=cores); dt = time.perf_counter() - t1
        t1 = time.perf_counter(); om.predict(xs[:2048], ps[:2048], nthreads=1); dt1 = time.perf_counter() - t1
        cpu = {"value": ns / dt, "unit": "events/s", "cores": cores, "kind": "port",
               "sample": f"{ns} events of the same workload, oracle/oracle.cc with {cores} threads (JPMML/Flink not runnable on this box: no JVM)",
               "single_thread_value": 2048 / dt1,
               "parity_on_sample": "bit-exact" if (np.array_equal(o["score"], hS[:ns].numpy()) and np.array_equal(o["status"], hT[:ns].numpy())) else "MISMATCH"}

    # ------------------------------------------------------------------ through the operator (what a Flink job sees): rank 0, N = 1
    operator = None
    if rank == 0 and world == 1 and not args.no_latency:
        try:
            import tempfile
            from flink_jpmml_b200 import api
            d = tempfile.mkdtemp(); path = os.path.join(d, "c4.pmml")
            with open(path, "wb") as f:
                f.write(pm)
            xs = hX[: 1 << 20].numpy()

            def drive(n_threads):
                fns = [api.EvaluationFunction(eng, path, max_batch=16384) for _ in range(n_threads)]
                for fn in fns:
                    fn.processDense(xs[:16384]); fn.flush(); fn.collect_arrays()
                emitted = [0] * n_threads

                def task(k):
                    fns[k].processDense(xs); fns[k].flush(); emitted[k] = len(fns[k].collect_arrays()[0])
                ths = [threading.Thread(target=task, args=(k,)) for k in range(n_threads)]
                t1 = time.perf_counter(); [t.start() for t in ths]; [t.join() for t in ths]; dt = time.perf_counter() - t1
                for fn in fns:
                    fn.close()
                return sum(emitted) / dt, sum(emitted)
            r1, e1 = drive(1); r4, e4 = drive(4)
            import importlib.util
            spec = importlib.util.spec_from_file_location("bench_operators", os.path.join(os.path.dirname(os.path.abspath(__file__)), "tools", "bench_operators.py"))
            bo = importlib.util.module_from_spec(spec); spec.loader.exec_module(bo)
            core = bo.one_core_copy(F=F)
            names5, files5 = bo.c5_models(d)
            dyn_swaps = bo.c5_dynamic(eng, names5, files5, 0.050, passes=2)
            dyn_quiet = bo.c5_dynamic(eng, names5, files5, 1e9, passes=2)
            operator = {"fixed_model_C4_1_task_thread": {"events_per_s": r1, "emitted": e1, "frac_of_one_core_copy": r1 / core["events_per_s"]},
                        "fixed_model_C4_4_task_threads_one_engine": {"events_per_s": r4, "emitted": e4},
                        "one_core_copy_into_pinned": core,
                        "dynamic_C5_32_ids_hot_swap_every_50ms": dyn_swaps, "dynamic_C5_32_ids_no_control_traffic": dyn_quiet,
                        "max_batch": 16384, "api": "fjh_eval_function_process_dense / fjh_cofunction_process_dense -> pinned micro-batches -> fjp_submit / fjp_wait (EvaluationFunction / EvaluationCoFunction mirrors, S/api/functions/EvaluationFunction.scala:39-50, EvaluationCoFunction.scala:55-133)",
                        "note": "one task thread copies every 800-byte event into a pinned micro-batch: the rate is the host core's copy bandwidth (one_core_copy_into_pinned), not the GPU's; "
                                "events for a model whose background load is still running are parked and scored when it finishes, the other models' events keep flowing"}
        except Exception as ex:          # noqa: BLE001
            operator = {"error": repr(ex)}

    if rank == 0:
        pk = peaks()
        value = D.whole_job_rate(n, world, args.steps, ms_dev_max)
        per_launch_s = ms_dev_max / 1e3 / args.steps
        achieved = n * ALGO_HBM_BYTES / per_launch_s / 1e9
        micro = pk.get("micro", {})
        gather = n * ALGO_GATHER_BYTES / per_launch_s / 1e9
        # What binds the walk is the shared-memory pipe (one wavefront per cycle and SM): `frac` is the share of those cycles the
        # timed region used, from the wavefronts per event ncu counted for these kernels (profiles/) and the SM clock sampled
        # during the run.  The HBM roofline of SURVEY.md section 8(d) (824 algorithmic B/event) and its L2-gather figure follow as
        # secondary keys: neither binds (the node records are gathered from shared memory, not from L2).
        wf_event = micro.get("forest_step_smem_wavefronts_per_event")
        sm_mhz = clocks.get("sm_mhz") or pk.get("sm_max_mhz") or 1965.0
        wf_peak = 148 * sm_mhz * 1e6 / 1e9                      # G wavefronts/s the chip can issue
        wf_rate = (n * wf_event / per_launch_s / 1e9) if wf_event else None
        roof = {"bound": "smem_pipe", "achieved": wf_rate, "peak": wf_peak, "unit": "Gwavefronts/s", "frac": (wf_rate / wf_peak) if wf_rate else None,
                "wavefronts_per_event": wf_event, "wavefronts_source": micro.get("forest_step_smem_wavefronts_note"),
                # the same with only the wavefronts a conflict-free access pattern would need (ncu's "ideal"): the replays are work
                # the pipe really does, but not work the algorithm asks for
                "wavefronts_per_event_conflict_free": micro.get("forest_step_smem_wavefronts_ideal_per_event"),
                "frac_conflict_free": (n * micro["forest_step_smem_wavefronts_ideal_per_event"] / per_launch_s / 1e9 / wf_peak) if micro.get("forest_step_smem_wavefronts_ideal_per_event") else None,
                "peak_note": "148 SMs x SM clock sampled during the timed region (%.0f MHz): one shared-memory wavefront per SM and cycle" % sm_mhz,
                "traffic": (micro["forest_kernel_dram_bytes_per_event"] * n) if micro.get("forest_kernel_dram_bytes_per_event") else None,
                "traffic_source": micro.get("forest_kernel_dram_note"),
                "kernel": ("forest_rank_keys_kernel + forest_compact_kernel<8,8,0,640> (throughput path: %d launches per step, timed together)" % (launches // max(1, args.steps)))
                          if launches > args.steps else "forest_complete_kernel<8,8,0,false>",
                "hbm": {"algorithmic_bytes_per_event": ALGO_HBM_BYTES, "achieved_gbs": achieved, "peak_gbs": pk["hbm_gbs"], "frac": achieved / pk["hbm_gbs"], "peak_source": pk["source"]},
                "smem_pipe": {"walk_kernel_busy_frac_ncu": micro.get("forest_walk_smem_pipe_busy"), "walk_kernel_issue_slots_busy_frac_ncu": micro.get("forest_walk_issue_slots_busy"),
                              "source": micro.get("forest_walk_note")},
                "gather": {"algorithmic_bytes_per_event": ALGO_GATHER_BYTES, "achieved_gbs": gather,
                           "l2_gather_peak_gbs": micro.get("l2_gather_8B_2MB_GBs"), "smem_gather_peak_gbs": micro.get("smem_gather_8B_random_GBs"),
                           "frac_of_l2_gather": (gather / micro["l2_gather_8B_2MB_GBs"]) if micro.get("l2_gather_8B_2MB_GBs") else None,
                           "frac_of_smem_gather": (gather / micro["smem_gather_8B_random_GBs"]) if micro.get("smem_gather_8B_random_GBs") else None}}
        configs = {}; pk["sm_mhz_sampled"] = sm_mhz
        for k in cfg_keys:
            c = cfg_out[k]; per_gpu = c["events_per_gpu_per_step"] / (c["ms_per_step"] / 1e3)
            c["events_per_s"] = per_gpu * world; c["roofline"] = config_rooflines(per_gpu, c.pop("work"), pk)
            configs[k] = c
        e2e_value = D.whole_job_rate(ne, world, e2e_steps, ms_e2e_max)
        line = {"metric": "scored events/sec", "value": value, "unit": "events/s", "n_gpus": world, "steps": args.steps, "warmup": args.warmup,
                "ms_per_step": ms_dev_max / args.steps, "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "f64",
                "data": "synthetic (torch philox on device, seed 20261110+1000*rank)",
                "config": {"workload": WORKLOAD, "events_per_gpu_per_step": n, "l2_policy": "inputs (3.4 GB/GPU) larger than L2",
                           "model": model.describe, "parallelism": f"events sharded over {world} GPU(s), model replicated, no collective" + (f"; rank 0 bound to NUMA node {numa_node}" if numa_node is not None else "")},
                "e2e": {"value": e2e_value, "unit": "events/s", "h2d_bytes_per_step": ne * (F * (4 if e2e_rows == "float" else 8) + W * 4), "d2h_bytes_per_step": ne * 9,
                        "events_per_gpu_per_step": ne, "steps": e2e_steps, "matches_device_path": e2e_match, "rows": e2e_rows,
                        "api": ("fjp_score_batch_f32: float rows rounded per fjp_model_float_row_modes (the ensemble's tests see one float per value), "
                                "pinned host buffers, chunked copy/compute overlap" if e2e_rows == "float" else "fjp_score_batch (pinned host buffers, chunked copy/compute overlap)"),
                        "float_rows_match_double_rows": e2e_f32_match,
                        "double_rows": {"value": D.whole_job_rate(ne, world, e2e_steps, ms_e2e_f64_max), "h2d_bytes_per_step": ne * (F * 8 + W * 4), "api": "fjp_score_batch"},
                        "h2d_copy_GBs_this_box": h2d_live, "h2d_copy_GBs_reference": micro.get("h2d_pinned_GBs"),
                        "frac_of_h2d_copy_peak": (e2e_value / world * (F * (4 if e2e_rows == "float" else 8) + W * 4) / 1e9 / (h2d_live or micro.get("h2d_pinned_GBs") or float("nan")))
                                                 if (h2d_live or micro.get("h2d_pinned_GBs")) else None},
                "gpu_launches": launches, "rows_scored_ok": ok_rows, "roofline": roof, "cpu_baseline": cpu, "clocks": clocks, "configs": configs, "replicate": replicate, "operator": operator,
                "latency_ms": latency}
        print(json.dumps(line), flush=True)
    if dist:
        dist.barrier()
        dist.destroy_process_group()
    eng.close()


if __name__ == "__main__":
    main()
