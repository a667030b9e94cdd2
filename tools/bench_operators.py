"""Operator-level throughput through the host mirror (pinned micro-batches, submit/wait double buffering):
  C4  fixed-model operator (quickEvaluate shape), 2^20 events
  C5  dynamic serving: 32 ModelIds (16 RF 200x8 F=64 + 16 MLP 64-256-256-8), events addressed uniformly,
      every 50 ms a random id gets Del then Add(version + 1) while scoring runs (BASELINE.json config 5)
One JSON line per config.  Host-side packing runs on ONE task thread, as one Flink subtask would."""
import json, os, sys, tempfile, time, uuid
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
from flink_jpmml_b200 import api, capi, pmmlgen as g

eng = capi.Engine(0)
tmp = tempfile.mkdtemp()
rng = np.random.default_rng(20261106 + 5)

# ---------------- C4 through EvaluationFunction
pm, _ = g.gbt_pmml()
path = os.path.join(tmp, "c4.pmml"); open(path, "wb").write(pm)
X = rng.standard_normal((1 << 20, 100))
for mb in (16384, 262144):
    fn = api.EvaluationFunction(eng, path, max_batch=mb)
    fn.processDense(X[:mb]); fn.flush(); fn.collect()
    t = time.perf_counter(); fn.processDense(X); fn.flush(); dt = time.perf_counter() - t
    n_out = len(fn.collect()); fn.close()
    print(json.dumps({"config": "C4 fixed-model operator (dense events, 1 task thread)", "max_batch": mb, "events_per_s": len(X) / dt, "emitted": n_out}), flush=True)

# ---------------- C5 dynamic serving
names, files = [], []
for k in range(32):
    pmk = g.rf_pmml(seed=1000 + k)[0] if k < 16 else g.nn_pmml(seed=1000 + k)[0]
    p = os.path.join(tmp, f"m{k}.pmml"); open(p, "wb").write(pmk)
    names.append(str(uuid.UUID(int=7000 + k))); files.append(p)
version = [1] * 32
op = api.EvaluationCoFunction(eng, max_batch=65536)
for k in range(32): op.processElement2(api.AddMessage(names[k], 1, files[k]))
N, CH = 1 << 21, 32768
X = rng.standard_normal((N, 64)); idx = rng.integers(0, 32, N).astype(np.int32)
ids = lambda: [f"{names[k]}_{version[k]}" for k in range(32)]
op.processDense1(ids(), idx[:CH], X[:CH]); op.flush(); op.collect()      # warm-up: lazy loads of all 32 models
swaps, empties, emitted = 0, 0, 0
t0 = time.perf_counter(); last = t0
for lo in range(0, N, CH):
    now = time.perf_counter()
    if now - last >= float(os.environ.get("FJP_BENCH_SWAP_S", "0.050")):   # FJP_BENCH_SWAP_S=1e9: no hot swaps
        k = int(rng.integers(32)); op.processElement2(api.DelMessage(names[k], version[k])); version[k] += 1
        op.processElement2(api.AddMessage(names[k], version[k], files[k])); swaps += 1; last = now
    op.processDense1(ids(), idx[lo:lo + CH], X[lo:lo + CH])
op.flush(); dt = time.perf_counter() - t0
out = op.collect(); emitted = len(out); empties = sum(1 for p in out if p.value is api.EmptyScore)
print(json.dumps({"config": "C5 dynamic serving: 32 ModelIds (16 RF + 16 MLP), Del/Add(version+1) every 50 ms, 1 task thread", "events_per_s": N / dt,
                  "hot_swaps": swaps, "emitted": emitted, "empty_scores": empties, "seconds": dt}), flush=True)
op.close()
