import os, sys, time, uuid, tempfile, json
sys.path.insert(0, "/root/repo")
import numpy as np
from flink_jpmml_b200 import capi, api, pmmlgen as g
rng = np.random.default_rng(1)
eng = capi.Engine(0); tmp = tempfile.mkdtemp()
names, files = [], []
for k in range(32):
    pmk = g.rf_pmml(seed=1000 + k)[0] if k < 16 else g.nn_pmml(seed=1000 + k)[0]
    p = os.path.join(tmp, f"m{k}.pmml"); open(p, "wb").write(pmk); names.append(str(uuid.UUID(int=7000 + k))); files.append(p)
version = [1] * 32
op = api.EvaluationCoFunction(eng, max_batch=65536)
for k in range(32): op.processElement2(api.AddMessage(names[k], 1, files[k]))
N, CH = 1 << 20, 32768
X = rng.standard_normal((N, 64)); idx = rng.integers(0, 32, N).astype(np.int32)
ids = lambda: [f"{names[k]}_{version[k]}" for k in range(32)]
op.processDense1(ids(), idx[:CH], X[:CH]); op.flush(); op.collect()
for trial, k in enumerate([3, 20, 5, 25]):
    t0 = time.perf_counter(); op.processElement2(api.DelMessage(names[k], version[k])); t1 = time.perf_counter()
    version[k] += 1; op.processElement2(api.AddMessage(names[k], version[k], files[k])); t2 = time.perf_counter()
    if trial >= 2: time.sleep(0.4)      # give the background load time to finish
    t3 = time.perf_counter(); op.processDense1(ids(), idx[:CH], X[:CH]); t4 = time.perf_counter()
    op.processDense1(ids(), idx[CH:2*CH], X[CH:2*CH]); t5 = time.perf_counter()
    print(json.dumps({"model": "RF" if k < 16 else "MLP", "slept": trial >= 2, "del_ms": (t1-t0)*1e3, "add_ms": (t2-t1)*1e3, "first_chunk_ms": (t4-t3)*1e3, "next_chunk_ms": (t5-t4)*1e3}))
op.flush(); op.close()
